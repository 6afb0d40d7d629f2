// Kernel family 1, diagnostic variant: fp32 FMA all-pairs volume + level-by-level pooling.
// Same contract as raftk_corr_pyramid (model.py:418-446, 472-481); no tensor cores.
// The tests use it to tell indexing/pooling-bound errors from split-precision error of
// the tcgen05 kernel, and the pooling kernel here is also the reference's "pool level
// l+1 from level l" order, bit-identical to a float32 CPU evaluation of
// ((a + b) + (c + d)) / 4.

#include "raftk_common.cuh"

namespace raftk {

constexpr int TM = 64, TN = 64, TK = 16;

// C[q, j] = sum_d A[q, d] * Bm[j, d] * scale;  A: (nq, D), Bm: (N, D), C: (nq, N) per batch.
__global__ void __launch_bounds__(256)
volume_simt_kernel(const float* __restrict__ f1, const float* __restrict__ f2, float* __restrict__ c,
                   int nq, int N, int D, float sqrt_d) {
  __shared__ float As[TK][TM + 1];
  __shared__ float Bs[TK][TN + 1];
  const int b = blockIdx.z;
  const float* A = f1 + (size_t)b * nq * D;
  const float* Bm = f2 + (size_t)b * N * D;
  float* C = c + (size_t)b * nq * N;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < D; k0 += TK) {
    for (int e = threadIdx.x; e < TM * TK; e += 256) {
      const int rr = e / TK, kk = e - rr * TK;
      const int m = m0 + rr, n = n0 + rr, k = k0 + kk;
      As[kk][rr] = (m < nq && k < D) ? __ldg(A + (size_t)m * D + k) : 0.0f;
      Bs[kk][rr] = (n < N && k < D) ? __ldg(Bm + (size_t)n * D + k) : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      float a[4], bb[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bb[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= nq) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n < N) C[(size_t)m * N + n] = acc[i][j] / sqrt_d;
    }
  }
}

// out[s, y, x] = ((in[s,2y,2x] + in[s,2y,2x+1]) + (in[s,2y+1,2x] + in[s,2y+1,2x+1])) / 4
__global__ void __launch_bounds__(256)
pool2x2_kernel(const float* __restrict__ in, float* __restrict__ out, long long slabs, int ih, int iw) {
  const int oh = ih >> 1, ow = iw >> 1;
  const long long total = slabs * oh * ow;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(t % ow);
    const int y = (int)((t / ow) % oh);
    const long long s = t / ((long long)ow * oh);
    const float* p = in + (s * ih + 2 * y) * (long long)iw + 2 * x;
    const float top = __fadd_rn(__ldg(p), __ldg(p + 1));
    const float bot = __fadd_rn(__ldg(p + iw), __ldg(p + iw + 1));
    out[t] = __fadd_rn(top, bot) * 0.25f;
  }
}

int check_pyramid_args(const float* fmap1, const float* fmap2, float* const* levels, int num_levels,
                       int B, int nq, int h, int w, int D) {
  if (num_levels < 1 || num_levels > RAFTK_MAX_LEVELS || B < 0 || nq < 0 || h <= 0 || w <= 0 || D <= 0)
    return RAFTK_ERR_SHAPE;
  if (!levels_fit(h, w, num_levels)) return RAFTK_ERR_TOO_SMALL;
  if (B == 0 || nq == 0) return RAFTK_OK;  // empty batch: pointers may be NULL
  if (!fmap1 || !fmap2 || !levels) return RAFTK_ERR_NULL;
  for (int l = 0; l < num_levels; ++l)
    if (!levels[l]) return RAFTK_ERR_NULL;
  return RAFTK_OK;
}

int launch_pool_levels(float* const* levels, int first_out_level, int num_levels, long long slabs,
                       int h, int w, cudaStream_t s) {
  for (int l = first_out_level; l < num_levels; ++l) {
    const int ih = h >> (l - 1), iw = w >> (l - 1);
    const long long total = slabs * (ih >> 1) * (iw >> 1);
    if (total == 0) continue;
    const int blocks = (int)((total + 255) / 256 < (long long)RAFTK_SM_COUNT * 16
                                 ? (total + 255) / 256
                                 : (long long)RAFTK_SM_COUNT * 16);
    pool2x2_kernel<<<blocks, 256, 0, s>>>(levels[l - 1], levels[l], slabs, ih, iw);
  }
  return launch_status();
}

}  // namespace raftk

extern "C" int raftk_corr_pyramid_simt(const float* fmap1, const float* fmap2, float* const* levels,
                                       int num_levels, int B, int nq, int h, int w, int D,
                                       void* stream) {
  using namespace raftk;
  int rc = check_pyramid_args(fmap1, fmap2, levels, num_levels, B, nq, h, w, D);
  if (rc != RAFTK_OK) return rc;
  if (B == 0 || nq == 0) return RAFTK_OK;
  if (B > 65535) return RAFTK_ERR_UNSUPPORTED;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int N = h * w;
  dim3 grid((N + TN - 1) / TN, (nq + TM - 1) / TM, B);
  // model.py:481 divides by np.sqrt(D): a float64 scalar that jax rounds to float32.
  const float sqrt_d = (float)sqrt((double)D);
  volume_simt_kernel<<<grid, 256, 0, s>>>(fmap1, fmap2, levels[0], nq, N, D, sqrt_d);
  rc = launch_status();
  if (rc != RAFTK_OK) return rc;
  return launch_pool_levels(levels, 1, num_levels, (long long)B * nq, h, w, s);
}
