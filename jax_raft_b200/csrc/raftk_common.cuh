// Shared helpers for the raftk kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "raftk.h"

#define RAFTK_SM_COUNT 148  // B200: 2 dies x 74 SMs

namespace raftk {

// Launch-error check: peek so an earlier sticky error is reported but not cleared twice.
static inline int launch_status() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? RAFTK_OK : static_cast<int>(e);
}

static inline bool levels_fit(int h, int w, int num_levels) {
  // model.py:431-436: the coarsest level must keep at least 2 rows and 2 columns.
  const int min_size = 2 * (1 << (num_levels - 1));
  return h >= min_size && w >= min_size;
}

__device__ __forceinline__ float ldg_f32(const float* p) { return __ldg(p); }

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
