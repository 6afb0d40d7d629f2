// Host-model glue, NOT a row of SURVEY.md §8: InstanceNorm (+ ReLU) on an NHWC tensor, the norm
// layer of the feature encoder (model.py:706-707: nn.InstanceNorm(epsilon=1e-5, no scale / bias);
// applied inside ConvNormActivation, model.py:146-158).  torch runs it as two batch-norm kernels
// (~100 us per call at 220x512x64); this is one statistics pass and one normalise pass.
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
    auto add = [&](const float4 v) {
      fs[0] += v.x; fs[1] += v.y; fs[2] += v.z; fs[3] += v.w;
      fq[0] += v.x * v.x; fq[1] += v.y * v.y; fq[2] += v.z * v.z; fq[3] += v.w * v.w;
    };
    int p = p0 + tr;
    for (; p + 3 * rows < p1; p += 4 * rows) {  // four independent loads in flight per thread
      const float4 v0 = __ldg(x4 + (size_t)p * c4 + tc);
      const float4 v1 = __ldg(x4 + (size_t)(p + rows) * c4 + tc);
      const float4 v2 = __ldg(x4 + (size_t)(p + 2 * rows) * c4 + tc);
      const float4 v3 = __ldg(x4 + (size_t)(p + 3 * rows) * c4 + tc);
      add(v0); add(v1); add(v2); add(v3);
      if (++cnt == 16) {  // spill the float partials (64 pixels) into the double accumulators
#pragma unroll
        for (int i = 0; i < 4; ++i) { s[i] += fs[i]; q[i] += fq[i]; fs[i] = 0; fq[i] = 0; }
        cnt = 0;
      }
    }
    for (; p < p1; p += rows) add(__ldg(x4 + (size_t)p * c4 + tc));
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

// Thread (tr, tc) owns channel group tc (4 channels) of pixel rows tr, tr + rows, ... of one slice of
// image n: it turns its four (sum, sum of squares) float64 pairs into (mean, 1/std) once — no separate
// finalise launch — and the loop has no index arithmetic beyond adds.
// RES: out = relu(residual + y) on top (the residual block's tail, model.py:183-184).
template <bool RELU, bool RES>
__global__ void __launch_bounds__(256)
instnorm_apply_kernel(float* __restrict__ x, const float* __restrict__ residual, const double* __restrict__ stats, int npix,
                      int c4, int slices, float eps) {
  const int n = blockIdx.y, slice = blockIdx.x;
  const int rows = 256 / c4;
  const int tc = threadIdx.x % c4, tr = threadIdx.x / c4;
  if (tr >= rows) return;
  float4* x4 = reinterpret_cast<float4*>(x) + (size_t)n * npix * c4 + tc;
  const float4* r4 = RES ? reinterpret_cast<const float4*>(residual) + (size_t)n * npix * c4 + tc : nullptr;
  const double* st = stats + ((size_t)n * c4 * 4 + tc * 4) * 2;
  float2 mi[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const double mean = st[2 * i] / npix;
    const double var = st[2 * i + 1] / npix - mean * mean;
    mi[i] = make_float2((float)mean, rsqrtf((float)var + eps));
  }
  const int p0 = (int)((long long)npix * slice / slices), p1 = (int)((long long)npix * (slice + 1) / slices);
  auto norm = [&](float4 v, size_t off) {
    v.x = (v.x - mi[0].x) * mi[0].y;
    v.y = (v.y - mi[1].x) * mi[1].y;
    v.z = (v.z - mi[2].x) * mi[2].y;
    v.w = (v.w - mi[3].x) * mi[3].y;
    if (RELU) {
      v.x = fmaxf(v.x, 0.0f); v.y = fmaxf(v.y, 0.0f); v.z = fmaxf(v.z, 0.0f); v.w = fmaxf(v.w, 0.0f);
    }
    if (RES) {
      const float4 r = __ldg(r4 + off);
      v.x = fmaxf(v.x + r.x, 0.0f); v.y = fmaxf(v.y + r.y, 0.0f); v.z = fmaxf(v.z + r.z, 0.0f); v.w = fmaxf(v.w + r.w, 0.0f);
    }
    return v;
  };
  int p = p0 + tr;
  for (; p + 3 * rows < p1; p += 4 * rows) {
    const size_t o = (size_t)p * c4, step = (size_t)rows * c4;
    float4* a = x4 + o;
    const float4 v0 = a[0], v1 = a[step], v2 = a[2 * step], v3 = a[3 * step];
    a[0] = norm(v0, o); a[step] = norm(v1, o + step); a[2 * step] = norm(v2, o + 2 * step); a[3 * step] = norm(v3, o + 3 * step);
  }
  for (; p < p1; p += rows) x4[(size_t)p * c4] = norm(x4[(size_t)p * c4], (size_t)p * c4);
}

// x = relu(x + y), float4 grid-stride (the residual tail of a block whose norm is folded, model.py:183-184)
__global__ void __launch_bounds__(256) add_relu_kernel(float4* __restrict__ x, const float4* __restrict__ y, long long total4) {
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total4; t += (long long)gridDim.x * blockDim.x) {
    float4 a = x[t];
    const float4 b = __ldg(y + t);
    a.x = fmaxf(a.x + b.x, 0.0f); a.y = fmaxf(a.y + b.y, 0.0f); a.z = fmaxf(a.z + b.z, 0.0f); a.w = fmaxf(a.w + b.w, 0.0f);
    x[t] = a;
  }
}

}  // namespace raftk

extern "C" size_t raftk_instance_norm_scratch_bytes(int n, int c) { return (size_t)(n > 0 ? n : 0) * (c > 0 ? c : 0) * 2 * sizeof(double); }

static int instance_norm_impl(float* x, const float* residual, int n, int npix, int c, float eps, int relu, void* scratch,
                              size_t scratch_bytes, void* stream) {
  using namespace raftk;
  if (n < 0 || npix <= 0 || c <= 0) return RAFTK_ERR_SHAPE;
  if (n == 0) return RAFTK_OK;
  if (!x || !scratch) return RAFTK_ERR_NULL;
  if (c % 4 || c > 1024) return RAFTK_ERR_UNSUPPORTED;
  if (scratch_bytes < raftk_instance_norm_scratch_bytes(n, c)) return RAFTK_ERR_SCRATCH;
  if (((uintptr_t)x & 15) || ((uintptr_t)residual & 15) || ((uintptr_t)scratch & 7)) return RAFTK_ERR_ALIGN;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemsetAsync(scratch, 0, raftk_instance_norm_scratch_bytes(n, c), s);
  if (e != cudaSuccess) return (int)e;
  int slices = (RAFTK_SM_COUNT * 4 + n - 1) / n;
  if (slices > npix) slices = npix;
  instnorm_stats_kernel<<<dim3(slices, n), 256, 0, s>>>(x, static_cast<double*>(scratch), npix, c / 4, slices);
  int aslices = (RAFTK_SM_COUNT * 16 + n - 1) / n;
  if (aslices > npix) aslices = npix;
  const double* st = static_cast<const double*>(scratch);
  const dim3 grid(aslices, n);
  if (residual) {
    if (relu)
      instnorm_apply_kernel<true, true><<<grid, 256, 0, s>>>(x, residual, st, npix, c / 4, aslices, eps);
    else
      instnorm_apply_kernel<false, true><<<grid, 256, 0, s>>>(x, residual, st, npix, c / 4, aslices, eps);
  } else if (relu) {
    instnorm_apply_kernel<true, false><<<grid, 256, 0, s>>>(x, nullptr, st, npix, c / 4, aslices, eps);
  } else {
    instnorm_apply_kernel<false, false><<<grid, 256, 0, s>>>(x, nullptr, st, npix, c / 4, aslices, eps);
  }
  return launch_status();
}

extern "C" int raftk_instance_norm_nhwc(float* x, int n, int npix, int c, float eps, int relu, void* scratch,
                                        size_t scratch_bytes, void* stream) {
  return instance_norm_impl(x, nullptr, n, npix, c, eps, relu, scratch, scratch_bytes, stream);
}

extern "C" int raftk_instance_norm_residual_nhwc(float* x, const float* residual, int n, int npix, int c, float eps, int relu,
                                                 void* scratch, size_t scratch_bytes, void* stream) {
  if (n > 0 && !residual) return RAFTK_ERR_NULL;
  return instance_norm_impl(x, residual, n, npix, c, eps, relu, scratch, scratch_bytes, stream);
}

extern "C" int raftk_add_relu(float* x, const float* y, long long count, void* stream) {
  using namespace raftk;
  if (count < 0 || count % 4) return RAFTK_ERR_SHAPE;
  if (count == 0) return RAFTK_OK;
  if (!x || !y) return RAFTK_ERR_NULL;
  if (((uintptr_t)x | (uintptr_t)y) & 15) return RAFTK_ERR_ALIGN;
  long long blocks = (count / 4 + 255) / 256;
  if (blocks > (long long)RAFTK_SM_COUNT * 16) blocks = (long long)RAFTK_SM_COUNT * 16;
  add_relu_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<float4*>(x), reinterpret_cast<const float4*>(y), count / 4);
  return launch_status();
}
