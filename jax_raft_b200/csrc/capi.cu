// ABI bookkeeping for libraftk: version and error strings.
#include "raftk_common.cuh"

extern "C" int raftk_version(void) { return 1000; }

extern "C" const char* raftk_error_string(int code) {
  switch (code) {
    case RAFTK_OK: return "ok";
    case RAFTK_ERR_NULL: return "raftk: a required pointer is NULL";
    case RAFTK_ERR_SHAPE: return "raftk: a size is non-positive or inconsistent";
    case RAFTK_ERR_TOO_SMALL:
      return "raftk: feature maps are too small to be down-sampled by the correlation pyramid";
    case RAFTK_ERR_UNSUPPORTED: return "raftk: no kernel in this build for the requested configuration";
    case RAFTK_ERR_SCRATCH: return "raftk: scratch buffer missing or too small";
    case RAFTK_ERR_ALIGN: return "raftk: a pointer is not aligned as documented";
    case RAFTK_ERR_STATIC_LIMIT:
      return "raftk: problem exceeds the built-in scratch (RAFTK_STATIC_SCRATCH_BYTES = 16 MiB); "
             "use raftk_corr_pyramid with a caller-provided scratch buffer";
    default:
      if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
      return "raftk: unknown error code";
  }
}

namespace raftk {
int& pdl_flag() {
  static int flag = [] {
    const char* e = getenv("RAFTK_PDL");
    return (e && e[0] == '0') ? 0 : 1;
  }();
  return flag;
}
}  // namespace raftk

extern "C" int raftk_set_pdl(int on) {
  int& f = raftk::pdl_flag();
  const int prev = f;
  f = on ? 1 : 0;
  return prev;
}
