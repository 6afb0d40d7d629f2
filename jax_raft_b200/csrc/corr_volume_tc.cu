// Kernel family 1, product entry point (interim: forwards to the fp32 FMA kernels until
// the tcgen05 kernel lands in this file).
#include "raftk_common.cuh"

extern "C" size_t raftk_corr_pyramid_scratch_bytes(int B, int nq, int h, int w, int D) {
  (void)B; (void)nq; (void)h; (void)w; (void)D;
  return 0;
}

extern "C" int raftk_corr_pyramid(const float* fmap1, const float* fmap2, float* const* levels,
                                  int num_levels, int B, int nq, int h, int w, int D, void* scratch,
                                  size_t scratch_bytes, void* stream) {
  (void)scratch; (void)scratch_bytes;
  return raftk_corr_pyramid_simt(fmap1, fmap2, levels, num_levels, B, nq, h, w, D, stream);
}
