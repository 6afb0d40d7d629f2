// Host-model glue, NOT a row of SURVEY.md §8: y = act(x + bias[c]) in place on an NHWC conv output.
// The library convolutions are called without bias because torch adds a channels_last bias with a
// strided broadcast kernel (8.5 us per call, ~500 calls per 32-iteration step = 28 % of the step);
// this one pass also applies the activation that followed.  act: 0 none, 1 relu, 2 sigmoid, 3 tanh.

#include "raftk_common.cuh"

namespace raftk {

template <int ACT>
__device__ __forceinline__ float apply_act(float v) {
  if (ACT == 1) return fmaxf(v, 0.0f);
  if (ACT == 2) return 1.0f / (1.0f + expf(-v));
  if (ACT == 3) return tanhf(v);
  return v;
}

template <int ACT>
__global__ void __launch_bounds__(256)
bias_act_kernel(float* __restrict__ x, const float* __restrict__ bias, long long n4, int c4) {
  float4* x4 = reinterpret_cast<float4*>(x);
  const float4* b4 = reinterpret_cast<const float4*>(bias);
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n4; t += (long long)gridDim.x * blockDim.x) {
    float4 v = x4[t];
    const float4 b = __ldg(b4 + (int)(t % c4));
    v.x = apply_act<ACT>(v.x + b.x);
    v.y = apply_act<ACT>(v.y + b.y);
    v.z = apply_act<ACT>(v.z + b.z);
    v.w = apply_act<ACT>(v.w + b.w);
    x4[t] = v;
  }
}

template <int ACT>
__global__ void __launch_bounds__(256)
bias_act_scalar_kernel(float* __restrict__ x, const float* __restrict__ bias, long long n, int c) {
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x)
    x[t] = apply_act<ACT>(x[t] + __ldg(bias + (int)(t % c)));
}

}  // namespace raftk

extern "C" int raftk_bias_act_nhwc(float* x, const float* bias, long long npix, int c, int act, void* stream) {
  using namespace raftk;
  if (npix < 0 || c <= 0 || act < 0 || act > 3) return RAFTK_ERR_SHAPE;
  if (npix == 0) return RAFTK_OK;
  if (!x || !bias) return RAFTK_ERR_NULL;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const long long n = npix * c;
  const bool vec = (c % 4 == 0) && (((uintptr_t)x & 15) == 0) && (((uintptr_t)bias & 15) == 0);
  const long long work = vec ? n / 4 : n;
  long long blocks = (work + 255) / 256;
  if (blocks > (long long)RAFTK_SM_COUNT * 16) blocks = (long long)RAFTK_SM_COUNT * 16;
#define RAFTK_LAUNCH(A)                                                                      \
  if (vec) bias_act_kernel<A><<<(unsigned)blocks, 256, 0, s>>>(x, bias, n / 4, c / 4);        \
  else bias_act_scalar_kernel<A><<<(unsigned)blocks, 256, 0, s>>>(x, bias, n, c)
  switch (act) {
    case 0: RAFTK_LAUNCH(0); break;
    case 1: RAFTK_LAUNCH(1); break;
    case 2: RAFTK_LAUNCH(2); break;
    default: RAFTK_LAUNCH(3); break;
  }
#undef RAFTK_LAUNCH
  return launch_status();
}
