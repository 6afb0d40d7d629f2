// Kernel family 2b: "alt_corr" — the window lookup computed on the fly from the feature
// maps, so the O(N^2) volume never exists.  No reference counterpart;
// defined by linearity of the average pool:
//   pool_l(f1 . f2^T)[q, Y, X] = f1[q] . pool_l(f2)[Y, X]
// so sampling level l of the volume (model.py:462-466) equals sampling dot products
// against fmap2 pooled l times with the rule of model.py:443-445.
//
// Layout: fmap1 (B, nq, D); f2_levels[l] (B, h>>l, w>>l, D); channels innermost, so one
// target pixel is D contiguous floats and a warp reads it with coalesced 16-byte loads.

#include "lookup_taps.cuh"
#include "raftk_common.cuh"

namespace raftk {

// out[b, y, x, :] = ((in[b,2y,2x] + in[b,2y,2x+1]) + (in[b,2y+1,2x] + in[b,2y+1,2x+1])) / 4
__global__ void __launch_bounds__(256)
pool_fmap_kernel(const float* __restrict__ in, float* __restrict__ out, int B, int ih, int iw, int D4) {
  const int oh = ih >> 1, ow = iw >> 1;
  const long long total = (long long)B * oh * ow * D4;
  const float4* in4 = reinterpret_cast<const float4*>(in);
  float4* out4 = reinterpret_cast<float4*>(out);
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(t % D4);
    const int x = (int)((t / D4) % ow);
    const int y = (int)((t / ((long long)D4 * ow)) % oh);
    const int b = (int)(t / ((long long)D4 * ow * oh));
    const float4* p = in4 + (((long long)b * ih + 2 * y) * iw + 2 * x) * D4 + d;
    const float4 a = __ldg(p), bb = __ldg(p + D4), c = __ldg(p + (long long)iw * D4),
                 dd = __ldg(p + (long long)iw * D4 + D4);
    float4 o;
    o.x = __fadd_rn(__fadd_rn(a.x, bb.x), __fadd_rn(c.x, dd.x)) * 0.25f;
    o.y = __fadd_rn(__fadd_rn(a.y, bb.y), __fadd_rn(c.y, dd.y)) * 0.25f;
    o.z = __fadd_rn(__fadd_rn(a.z, bb.z), __fadd_rn(c.z, dd.z)) * 0.25f;
    o.w = __fadd_rn(__fadd_rn(a.w, bb.w), __fadd_rn(c.w, dd.w)) * 0.25f;
    out4[t] = o;
  }
}

struct AltParams {
  const float* f1;
  const float* f2[RAFTK_MAX_LEVELS];
  int lh[RAFTK_MAX_LEVELS];
  int lw[RAFTK_MAX_LEVELS];
  const float* coords;
  float* out;
  int B, nq, D, L, r;
  float inv_sqrt_d;
};

constexpr int kAltWarps = 4;
constexpr int kAltMaxD4 = 4;  // per-lane float4 registers: D <= 32 * 4 * 4 = 512

// One warp per query.  f1[q] lives in registers (float4 per lane per 128 channels); every
// footprint cell is a D-long dot product with a pooled fmap2 pixel, reduced by shuffles.
template <int R>
__global__ void __launch_bounds__(kAltWarps * 32) altcorr_lookup_kernel(const AltParams p) {
  extern __shared__ float smem[];
  const int r = R > 0 ? R : p.r;
  const int S = 2 * r + 1, FP = 2 * r + 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long q = (long long)blockIdx.x * kAltWarps + warp;
  if (q >= (long long)p.B * p.nq) return;
  const int b = (int)(q / p.nq);
  float* F = smem + warp * (p.L * FP * FP);

  const int nd4 = p.D >> 7;  // float4 chunks of 128 channels
  float4 a[kAltMaxD4];
  const float4* f1q = reinterpret_cast<const float4*>(p.f1 + q * p.D);
#pragma unroll
  for (int t = 0; t < kAltMaxD4; ++t) a[t] = t < nd4 ? __ldg(f1q + t * 32 + lane) : make_float4(0, 0, 0, 0);

  const float cx = __ldg(p.coords + 2 * q), cy = __ldg(p.coords + 2 * q + 1);
  float sc = 1.0f;
#pragma unroll 1
  for (int l = 0; l < p.L; ++l, sc *= 0.5f) {
    const int lh = p.lh[l], lw = p.lw[l];
    const int bx = clamp_to_int(floorf(cx * sc)) - r;
    const int by = clamp_to_int(floorf(cy * sc)) - r;
    const float* f2b = p.f2[l] + (long long)b * lh * lw * p.D;
    float* Fl = F + l * FP * FP;
#pragma unroll 1
    for (int e0 = 0; e0 < FP * FP; e0 += 4) {
      float part[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e = e0 + u;
        const int ry = e / FP, rx = e - ry * FP;
        const int gy = by + ry, gx = bx + rx;
        float s = 0.0f;
        if (e < FP * FP && (unsigned)gy < (unsigned)lh && (unsigned)gx < (unsigned)lw) {
          const float4* t2 = reinterpret_cast<const float4*>(f2b + ((long long)gy * lw + gx) * p.D);
#pragma unroll
          for (int t = 0; t < kAltMaxD4; ++t) {
            if (t < nd4) {
              const float4 v = __ldg(t2 + t * 32 + lane);
              s = fmaf(a[t].x, v.x, s);
              s = fmaf(a[t].y, v.y, s);
              s = fmaf(a[t].z, v.z, s);
              s = fmaf(a[t].w, v.w, s);
            }
          }
        }
        part[u] = s;
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
        for (int u = 0; u < 4; ++u) part[u] += __shfl_xor_sync(0xffffffffu, part[u], off);
      }
      if (lane < 4 && e0 + lane < FP * FP) {
        const float v = lane == 0 ? part[0] : lane == 1 ? part[1] : lane == 2 ? part[2] : part[3];
        Fl[e0 + lane] = v * p.inv_sqrt_d;
      }
    }
  }
  __syncwarp();
  lookup_taps<R>(F, cx, cy, p.L, r, p.out + q * (p.L * S * S), lane);
}

}  // namespace raftk

extern "C" int raftk_altcorr_pool_fmap2(float* const* f2_levels, int num_levels, int B, int h, int w,
                                        int D, void* stream) {
  using namespace raftk;
  if (num_levels < 1 || num_levels > RAFTK_MAX_LEVELS || B < 0 || h <= 0 || w <= 0 || D <= 0)
    return RAFTK_ERR_SHAPE;
  if (!levels_fit(h, w, num_levels)) return RAFTK_ERR_TOO_SMALL;
  if (D % 4) return RAFTK_ERR_UNSUPPORTED;
  if (B == 0) return RAFTK_OK;  // empty batch: pointers may be NULL
  if (!f2_levels) return RAFTK_ERR_NULL;
  for (int l = 0; l < num_levels; ++l) {
    if (!f2_levels[l]) return RAFTK_ERR_NULL;
    if ((uintptr_t)f2_levels[l] & 15) return RAFTK_ERR_ALIGN;
  }
  if (B == 0) return RAFTK_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  for (int l = 1; l < num_levels; ++l) {
    const int ih = h >> (l - 1), iw = w >> (l - 1);
    const long long total = (long long)B * (ih >> 1) * (iw >> 1) * (D / 4);
    long long blocks = (total + 255) / 256;
    if (blocks > (long long)RAFTK_SM_COUNT * 16) blocks = (long long)RAFTK_SM_COUNT * 16;
    pool_fmap_kernel<<<(unsigned)blocks, 256, 0, s>>>(f2_levels[l - 1], f2_levels[l], B, ih, iw, D / 4);
  }
  return launch_status();
}

extern "C" int raftk_altcorr_lookup(const float* fmap1, const float* const* f2_levels, int num_levels,
                                    const float* coords, float* out, int B, int nq, int h, int w,
                                    int D, int radius, void* stream) {
  using namespace raftk;
  if (num_levels < 1 || num_levels > RAFTK_MAX_LEVELS || B < 0 || nq < 0 || h <= 0 || w <= 0 ||
      D <= 0 || radius < 0)
    return RAFTK_ERR_SHAPE;
  if ((h >> (num_levels - 1)) < 1 || (w >> (num_levels - 1)) < 1) return RAFTK_ERR_TOO_SMALL;
  if (D % 128 || D > 128 * kAltMaxD4 || radius > 7) return RAFTK_ERR_UNSUPPORTED;
  if ((long long)B * nq == 0) return RAFTK_OK;  // empty batch: pointers may be NULL
  if (!fmap1 || !f2_levels || !coords || !out) return RAFTK_ERR_NULL;
  if ((uintptr_t)fmap1 & 15) return RAFTK_ERR_ALIGN;
  AltParams p{};
  for (int l = 0; l < num_levels; ++l) {
    if (!f2_levels[l]) return RAFTK_ERR_NULL;
    if ((uintptr_t)f2_levels[l] & 15) return RAFTK_ERR_ALIGN;
    p.f2[l] = f2_levels[l];
    p.lh[l] = h >> l;
    p.lw[l] = w >> l;
  }
  const long long Q = (long long)B * nq;
  if (Q == 0) return RAFTK_OK;
  p.f1 = fmap1;
  p.coords = coords;
  p.out = out;
  p.B = B;
  p.nq = nq;
  p.D = D;
  p.L = num_levels;
  p.r = radius;
  p.inv_sqrt_d = 1.0f / (float)sqrt((double)D);
  const int FP = 2 * radius + 2;
  const size_t smem = (size_t)kAltWarps * num_levels * FP * FP * sizeof(float);
  const long long blocks = (Q + kAltWarps - 1) / kAltWarps;
  if (blocks > 0x7fffffffLL) return RAFTK_ERR_SHAPE;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (radius == 4)
    altcorr_lookup_kernel<4><<<(unsigned)blocks, kAltWarps * 32, smem, s>>>(p);
  else if (radius == 3)
    altcorr_lookup_kernel<3><<<(unsigned)blocks, kAltWarps * 32, smem, s>>>(p);
  else
    altcorr_lookup_kernel<0><<<(unsigned)blocks, kAltWarps * 32, smem, s>>>(p);
  return launch_status();
}
