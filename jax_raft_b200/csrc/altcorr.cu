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

// ------------------------------------------------------------------------------------------
// Tiled alt_corr: a block owns an 8x8 patch of query pixels at ONE pyramid level.  Neighbouring
// queries have overlapping windows (optical flow is smooth), so the block fetches the bounding
// box of its 64 windows once — 32 channels at a time into shared memory — instead of every warp
// re-reading its own 100 target pixels (1 KB each) from L2: 16x less L2 traffic in the common
// case.  Thread (query, sub) accumulates the dot products of cells sub, sub+4, ... of its query's
// (2r+2)^2 footprint, each as one sequential fma chain over d = 0..D-1, so the value of a cell
// does not depend on which tile or path computed it.  If the 64 windows do not fit the box
// (rough flow, or far-out-of-range coordinates) the block falls back to a per-query loop with
// the same summation order.  Taps are then evaluated by the shared lookup code, level by level.
// ------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T pick_level(const T* arr, int l) {  // block-uniform select: keeps params out of local memory
  T v = arr[0];
#pragma unroll
  for (int i = 1; i < RAFTK_MAX_LEVELS; ++i)
    if (l == i) v = arr[i];
  return v;
}

constexpr int kTileThreads = 256;
constexpr int kTileQ = 64;          // 8 x 8 queries
constexpr int kCellCap = 576;       // bounding-box cells held in shared memory (e.g. 24 x 24)
constexpr int kCellPitch = 36;      // floats per cell per 32-channel chunk (+4: bank spread, 16-byte aligned)

template <int R>
__global__ void __launch_bounds__(kTileThreads) altcorr_tile_kernel(const AltParams p, int rows_q, int w_q) {
  constexpr int FP = 2 * R + 2, S = 2 * R + 1, NCELL = FP * FP, NC = NCELL / 4;
  extern __shared__ float smem[];
  float* cellbuf = smem;                              // [kCellCap][kCellPitch]
  float* Fq = smem + kCellCap * kCellPitch;           // [kTileQ][NCELL]
  __shared__ int red[4][8];
  __shared__ int box[5];

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int qi = t >> 2, sub = t & 3;
  if (t == 0) box[4] = 0;
  __syncthreads();
  const int tiles_x = (w_q + 7) >> 3, tiles_y = (rows_q + 7) >> 3;
  int bid = blockIdx.x;
  const int l = bid % p.L;
  bid /= p.L;
  const int txi = bid % tiles_x;
  bid /= tiles_x;
  const int tyi = bid % tiles_y;
  const int b = bid / tiles_y;
  const int qy = tyi * 8 + (qi >> 3), qx = txi * 8 + (qi & 7);
  const bool qvalid = qy < rows_q && qx < w_q;
  const long long q = (long long)b * p.nq + (long long)qy * w_q + qx;

  const int lh = pick_level(p.lh, l), lw = pick_level(p.lw, l);
  const float* f2b = pick_level(p.f2, l) + (long long)b * lh * lw * p.D;
  const float sc = __int_as_float((127 - l) << 23);
  float cx = 0.0f, cy = 0.0f;
  if (qvalid) {
    cx = __ldg(p.coords + 2 * q);
    cy = __ldg(p.coords + 2 * q + 1);
  }
  const bool normal = fabsf(cx) <= 1.0e6f && fabsf(cy) <= 1.0e6f;
  const int bx = clamp_to_int(floorf(cx * sc)) - R, by = clamp_to_int(floorf(cy * sc)) - R;

  // ---- bounding box of the tile's windows ----
  {
    int mnx = qvalid ? bx : 0x7fffffff, mxx = qvalid ? bx : -0x7fffffff;
    int mny = qvalid ? by : 0x7fffffff, mxy = qvalid ? by : -0x7fffffff;
    int bad = (qvalid && !normal) ? 1 : 0;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      mnx = min(mnx, __shfl_xor_sync(0xffffffffu, mnx, off));
      mxx = max(mxx, __shfl_xor_sync(0xffffffffu, mxx, off));
      mny = min(mny, __shfl_xor_sync(0xffffffffu, mny, off));
      mxy = max(mxy, __shfl_xor_sync(0xffffffffu, mxy, off));
      bad |= __shfl_xor_sync(0xffffffffu, bad, off);
    }
    if (lane == 0) {
      red[0][warp] = mnx;
      red[1][warp] = mxx;
      red[2][warp] = mny;
      red[3][warp] = mxy;
      if (bad) atomicOr(&box[4], 1);
    }
  }
  __syncthreads();
  if (t == 0) {
    int mnx = red[0][0], mxx = red[1][0], mny = red[2][0], mxy = red[3][0];
    for (int i = 1; i < 8; ++i) {
      mnx = min(mnx, red[0][i]);
      mxx = max(mxx, red[1][i]);
      mny = min(mny, red[2][i]);
      mxy = max(mxy, red[3][i]);
    }
    box[0] = mnx;
    box[1] = mny;
    box[2] = mxx - mnx + FP;  // width in cells
    box[3] = mxy - mny + FP;  // height
  }
  __syncthreads();
  const int X0 = box[0], Y0 = box[1], BW = box[2], BH = box[3];
  const bool shared_path = box[4] == 0 && BW > 0 && BH > 0 && (long long)BW * BH <= kCellCap;

  float acc[NC];
#pragma unroll
  for (int i = 0; i < NC; ++i) acc[i] = 0.0f;
  const float* f1q = p.f1 + (qvalid ? q : 0) * p.D;

  if (shared_path) {
    const int ncell = BW * BH;
    const int rel = (by - Y0) * BW + (bx - X0);  // this query's window origin inside the box
    for (int c0 = 0; c0 < p.D; c0 += 32) {
      __syncthreads();  // the previous chunk has been consumed
      for (int idx = t; idx < ncell * 8; idx += kTileThreads) {
        const int cell = idx >> 3, part = idx & 7;
        const int cyy = cell / BW, cxx = cell - cyy * BW;
        const int gy = Y0 + cyy, gx = X0 + cxx;
        float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if ((unsigned)gy < (unsigned)lh && (unsigned)gx < (unsigned)lw)
          v = __ldg(reinterpret_cast<const float4*>(f2b + ((long long)gy * lw + gx) * p.D + c0) + part);
        *reinterpret_cast<float4*>(cellbuf + cell * kCellPitch + part * 4) = v;
      }
      __syncthreads();
      if (qvalid) {
        float4 a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = __ldg(reinterpret_cast<const float4*>(f1q + c0) + k);
#pragma unroll
        for (int i = 0; i < NC; ++i) {
          const int e = sub + 4 * i;
          const int ry = e / FP, rx = e - ry * FP;
          const float4* cp = reinterpret_cast<const float4*>(cellbuf + (rel + ry * BW + rx) * kCellPitch);
          float s = acc[i];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float4 v = cp[k];
            s = fmaf(a[k].x, v.x, s);
            s = fmaf(a[k].y, v.y, s);
            s = fmaf(a[k].z, v.z, s);
            s = fmaf(a[k].w, v.w, s);
          }
          acc[i] = s;
        }
      }
    }
  } else if (qvalid) {
    // rough flow: every thread walks its own cells in global memory, same summation order
#pragma unroll 1
    for (int i = 0; i < NC; ++i) {
      const int e = sub + 4 * i;
      const int ry = e / FP, rx = e - ry * FP;
      const int gy = by + ry, gx = bx + rx;
      float s = 0.0f;
      if (normal && (unsigned)gy < (unsigned)lh && (unsigned)gx < (unsigned)lw) {
        const float4* cp = reinterpret_cast<const float4*>(f2b + ((long long)gy * lw + gx) * p.D);
        const float4* ap = reinterpret_cast<const float4*>(f1q);
        for (int k = 0; k < p.D / 4; ++k) {
          const float4 v = __ldg(cp + k), a = __ldg(ap + k);
          s = fmaf(a.x, v.x, s);
          s = fmaf(a.y, v.y, s);
          s = fmaf(a.z, v.z, s);
          s = fmaf(a.w, v.w, s);
        }
      }
      acc[i] = s;
    }
  }
#pragma unroll
  for (int i = 0; i < NC; ++i) Fq[qi * NCELL + sub + 4 * i] = acc[i] * p.inv_sqrt_d;
  __syncthreads();

  // ---- taps of level l: warp w takes queries 8w .. 8w+7 of the tile ----
  const int C = p.L * S * S;
  for (int k = 0; k < 8; ++k) {
    const int qj = warp * 8 + k;
    const int yj = tyi * 8 + (qj >> 3), xj = txi * 8 + (qj & 7);
    if (yj >= rows_q || xj >= w_q) continue;
    const long long qq = (long long)b * p.nq + (long long)yj * w_q + xj;
    const float ccx = __ldg(p.coords + 2 * qq), ccy = __ldg(p.coords + 2 * qq + 1);
    lookup_taps_level<R>(Fq + qj * NCELL, ccx, ccy, l, p.out + qq * C, lane);
  }
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
  const bool tiled = (radius == 3 || radius == 4) && D % 32 == 0 && nq % w == 0;
  if (!tiled && (D % 128 || D > 128 * kAltMaxD4 || radius > 7)) return RAFTK_ERR_UNSUPPORTED;
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
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // Tiled kernel: queries form whole image rows of width w (the full map or a band of rows)
  if (tiled) {
    const int rows_q = nq / w;
    const int FPt = 2 * radius + 2;
    const size_t smem_t = (size_t)(kCellCap * kCellPitch + kTileQ * FPt * FPt) * sizeof(float);
    const long long nblk = (long long)B * ((rows_q + 7) / 8) * ((w + 7) / 8) * num_levels;
    if (nblk > 0x7fffffffLL) return RAFTK_ERR_SHAPE;
    if (radius == 4) {
      if (cudaFuncSetAttribute(altcorr_tile_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_t) != cudaSuccess)
        return launch_status();
      altcorr_tile_kernel<4><<<(unsigned)nblk, kTileThreads, smem_t, s>>>(p, rows_q, w);
    } else {
      if (cudaFuncSetAttribute(altcorr_tile_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_t) != cudaSuccess)
        return launch_status();
      altcorr_tile_kernel<3><<<(unsigned)nblk, kTileThreads, smem_t, s>>>(p, rows_q, w);
    }
    return launch_status();
  }
  const int FP = 2 * radius + 2;
  const size_t smem = (size_t)kAltWarps * num_levels * FP * FP * sizeof(float);
  const long long blocks = (Q + kAltWarps - 1) / kAltWarps;
  if (blocks > 0x7fffffffLL) return RAFTK_ERR_SHAPE;
  if (radius == 4)
    altcorr_lookup_kernel<4><<<(unsigned)blocks, kAltWarps * 32, smem, s>>>(p);
  else if (radius == 3)
    altcorr_lookup_kernel<3><<<(unsigned)blocks, kAltWarps * 32, smem, s>>>(p);
  else
    altcorr_lookup_kernel<0><<<(unsigned)blocks, kAltWarps * 32, smem, s>>>(p);
  return launch_status();
}
