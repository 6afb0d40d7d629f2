// Shared helpers for the raftk kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdlib>

#include "raftk.h"

namespace raftk {

// SMs of the current device (B200: 148 = 2 dies x 74), queried once per device; grids are sized from it.
static inline int sm_count() {
  static int cached[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!cached[dev]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;  // racing writers store the same value
  }
  return cached[dev];
}
#define RAFTK_SM_COUNT (::raftk::sm_count())

// Status after a <<<>>> launch: the last error of the runtime, PEEKED (not cleared).  It is this launch's own
// failure (bad configuration) or an earlier error of the context that nobody has collected yet; either way the
// host framework still sees it.
static inline int launch_status() {
  cudaError_t e = cudaPeekAtLastError();
  return e == cudaSuccess ? RAFTK_OK : static_cast<int>(e);
}

static inline bool levels_fit(int h, int w, int num_levels) {
  // model.py:431-436: the coarsest level must keep at least 2 rows and 2 columns.
  const int min_size = 2 * (1 << (num_levels - 1));
  return h >= min_size && w >= min_size;
}

__device__ __forceinline__ float ldg_f32(const float* p) { return __ldg(p); }

// Programmatic dependent launch: a kernel launched through launch_pdl may be scheduled while its
// predecessor in the stream (or graph) is still draining; pdl_wait() — the first thing such a kernel
// does — blocks until the predecessor has completed and its writes are visible.  Hides most of the
// ~2 us node-to-node launch latency of the short kernels between the library convolutions.  A no-op
// in a kernel launched the ordinary way.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// process-wide switch (capi.cu: raftk_set_pdl); initial value from RAFTK_PDL (default on)
int& pdl_flag();
static inline bool pdl_enabled() { return pdl_flag() != 0; }

// Returns the launch's own status (cudaLaunchKernelEx's return value).
template <typename... KArgs, typename... Args>
static inline int launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
  return e == cudaSuccess ? RAFTK_OK : static_cast<int>(e);
}

// Streaming 128-bit store / load: the volume is written once and read much later,
// the lookup output is consumed by a different kernel — neither wants L1 residency.
__device__ __forceinline__ void st_na_f4(float* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

__device__ __forceinline__ float2 ldg_nc_f2(const float* p) {
  float2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0, %1}, [%2];" : "=f"(r.x), "=f"(r.y) : "l"(p));
  return r;
}

__device__ __forceinline__ float4 ldg_nc_f4(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

// L2 eviction-priority hints.  The lookup touches the same ~19 MB of pyramid rows in every one of
// the 32 iterations of a pair (the flow moves little), while the convolutions in between stream
// hundreds of MB through the 126 MB L2: marking the footprint reads evict_last keeps them resident
// from one iteration to the next.
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ float4 ldg_f4_l2hint(const float4* p, uint64_t pol) {
  float4 r;
  asm volatile("ld.global.nc.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p), "l"(pol));
  return r;
}

}  // namespace raftk
