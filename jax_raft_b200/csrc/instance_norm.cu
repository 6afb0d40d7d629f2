// Host-model glue, NOT a row of SURVEY.md §8: InstanceNorm (+ ReLU) on an NHWC tensor, the norm
// layer of the feature encoder (model.py:706-707: nn.InstanceNorm(epsilon=1e-5, no scale / bias);
// applied inside ConvNormActivation, model.py:146-158).  torch runs it as two batch-norm kernels
// (~100 us per call at 220x512x64); this is one statistics pass, a tiny finalise and one normalise pass.
//
//   y[n, p, c] = (x[n, p, c] - mean[n, c]) / sqrt(var[n, c] + eps),  var = E[x^2] - mean^2 (biased),
// statistics accumulated in float64 so the one-pass variance does not cancel.

#include "raftk_common.cuh"

namespace raftk {

// stats[(n*C + c)*2 + {0,1}] += sum, sum of squares over a slice of the pixels of image n
__global__ void __launch_bounds__(256)
instnorm_stats_kernel(const float* __restrict__ x, double* __restrict__ stats, int npix, int c4, int slices) {
  const int n = blockIdx.y, slice = blockIdx.x;
  const int lanes_c = c4;                      // threads along channels (float4 each)
  const int rows = 256 / lanes_c;              // pixel rows handled per block iteration
  const int tc = threadIdx.x % lanes_c, tr = threadIdx.x / lanes_c;
  const float4* x4 = reinterpret_cast<const float4*>(x) + (size_t)n * npix * c4;
  const int p0 = (int)((long long)npix * slice / slices), p1 = (int)((long long)npix * (slice + 1) / slices);
  double s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0};
  if (tr < rows) {
    float fs[4] = {0, 0, 0, 0}, fq[4] = {0, 0, 0, 0};
    int cnt = 0;
    for (int p = p0 + tr; p < p1; p += rows) {
      const float4 v = __ldg(x4 + (size_t)p * c4 + tc);
      fs[0] += v.x; fs[1] += v.y; fs[2] += v.z; fs[3] += v.w;
      fq[0] += v.x * v.x; fq[1] += v.y * v.y; fq[2] += v.z * v.z; fq[3] += v.w * v.w;
      if (++cnt == 64) {  // spill the float partials into the double accumulators
#pragma unroll
        for (int i = 0; i < 4; ++i) { s[i] += fs[i]; q[i] += fq[i]; fs[i] = 0; fq[i] = 0; }
        cnt = 0;
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) { s[i] += fs[i]; q[i] += fq[i]; }
  }
  // block reduction over the pixel rows, then ONE atomic per channel statistic per block
  __shared__ double red[256][8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    red[threadIdx.x][2 * i] = s[i];
    red[threadIdx.x][2 * i + 1] = q[i];
  }
  __syncthreads();
  if (tr == 0) {
    double acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = red[tc][i];
    for (int r = 1; r < rows; ++r)
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] += red[r * lanes_c + tc][i];
    double* dst = stats + ((size_t)n * c4 * 4 + tc * 4) * 2;
#pragma unroll
    for (int i = 0; i < 8; ++i) atomicAdd(dst + i, acc[i]);
  }
}

// (sum, sum of squares) in float64 -> (mean, 1/sqrt(var + eps)) in float32, in place: slot i keeps
// its 16 bytes, the first 8 now hold the float pair.
__global__ void __launch_bounds__(256) instnorm_finalize_kernel(double* __restrict__ stats, int count, int npix, float eps) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const double mean = stats[2 * i] / npix;
  const double var = stats[2 * i + 1] / npix - mean * mean;
  const float2 mi = make_float2((float)mean, rsqrtf((float)var + eps));
  *reinterpret_cast<float2*>(stats + 2 * i) = mi;
}

template <bool RELU>
__global__ void __launch_bounds__(256)
instnorm_apply_kernel(float* __restrict__ x, const double* __restrict__ stats, int npix, int c4, long long total4) {
  float4* x4 = reinterpret_cast<float4*>(x);
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total4; t += (long long)gridDim.x * blockDim.x) {
    const int cc = (int)(t % c4);
    const long long n = t / ((long long)npix * c4);
    const double* st = stats + ((size_t)n * c4 * 4 + cc * 4) * 2;
    float4 v = x4[t];
    float* pv = reinterpret_cast<float*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 mi = __ldg(reinterpret_cast<const float2*>(st + 2 * i));
      const float y = (pv[i] - mi.x) * mi.y;
      pv[i] = RELU ? fmaxf(y, 0.0f) : y;
    }
    x4[t] = v;
  }
}

}  // namespace raftk

extern "C" size_t raftk_instance_norm_scratch_bytes(int n, int c) { return (size_t)(n > 0 ? n : 0) * (c > 0 ? c : 0) * 2 * sizeof(double); }

extern "C" int raftk_instance_norm_nhwc(float* x, int n, int npix, int c, float eps, int relu, void* scratch,
                                        size_t scratch_bytes, void* stream) {
  using namespace raftk;
  if (n < 0 || npix <= 0 || c <= 0) return RAFTK_ERR_SHAPE;
  if (n == 0) return RAFTK_OK;
  if (!x || !scratch) return RAFTK_ERR_NULL;
  if (c % 4 || c > 1024) return RAFTK_ERR_UNSUPPORTED;
  if (scratch_bytes < raftk_instance_norm_scratch_bytes(n, c)) return RAFTK_ERR_SCRATCH;
  if (((uintptr_t)x & 15) || ((uintptr_t)scratch & 7)) return RAFTK_ERR_ALIGN;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemsetAsync(scratch, 0, raftk_instance_norm_scratch_bytes(n, c), s);
  if (e != cudaSuccess) return (int)e;
  int slices = (RAFTK_SM_COUNT * 4 + n - 1) / n;
  if (slices > npix) slices = npix;
  instnorm_stats_kernel<<<dim3(slices, n), 256, 0, s>>>(x, static_cast<double*>(scratch), npix, c / 4, slices);
  const long long total4 = (long long)n * npix * (c / 4);
  long long blocks = (total4 + 255) / 256;
  if (blocks > (long long)RAFTK_SM_COUNT * 16) blocks = (long long)RAFTK_SM_COUNT * 16;
  instnorm_finalize_kernel<<<(n * c + 255) / 256, 256, 0, s>>>(static_cast<double*>(scratch), n * c, npix, eps);
  if (relu)
    instnorm_apply_kernel<true><<<(unsigned)blocks, 256, 0, s>>>(x, static_cast<const double*>(scratch), npix, c / 4, total4);
  else
    instnorm_apply_kernel<false><<<(unsigned)blocks, 256, 0, s>>>(x, static_cast<const double*>(scratch), npix, c / 4, total4);
  return launch_status();
}
