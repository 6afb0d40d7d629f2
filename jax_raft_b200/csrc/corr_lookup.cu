// Kernel family 2: radius-r bilinear window lookup over the correlation pyramid.
// Replaces CorrBlock.index_pyramid (model.py:448-470) + grid_sample (model.py:24-34).
//
// Two kernels, same arithmetic, same bits:
//   corr_lookup_fast_kernel  L = 4, r in {3, 4} (the reference's configurations) — see below;
//   corr_lookup_kernel       other radii (r <= 7) and level counts (<= 6): one warp per query pixel; phase 1
//                            stages each level's (2r+2)^2 footprint in shared memory with lanes
//                            along footprint rows (4-byte loads), phase 2 gives each lane channels
//                            lane, lane+32, ... so the output row goes out as 128-byte warp stores.
//
// Arithmetic follows jax.scipy.ndimage.map_coordinates(order=1, mode='constant') as the
// reference calls it, PER TAP and in float32 with no FMA contraction:
//   s = c_l + d;  i0 = floor(s);  w1 = s - i0;  w0 = 1 - w1
//   out = (((wy0*wx0)*v00 + (wy0*wx1)*v01) + (wy1*wx0)*v10) + (wy1*wx1)*v11
// with v = 0 for a corner outside the level.  c_l = c / 2^l is exact.  Because
// fl(c_l + d) can round UP to an integer, a tap's i0 is floor(c_l)+d or floor(c_l)+d+1;
// in the second case w1 == 0 exactly, so the only corner that can fall outside the
// staged footprint carries weight 0 and reading it as 0 leaves the sum unchanged.

#include "raftk_common.cuh"
#include "lookup_taps.cuh"

namespace raftk {

struct LookupParams {
  const float* level[RAFTK_MAX_LEVELS];
  int lh[RAFTK_MAX_LEVELS];
  int lw[RAFTK_MAX_LEVELS];
  int aligned[RAFTK_MAX_LEVELS];  // level rows can be read as 16-byte chunks (pitch % 4 == 0, base % 16 == 0)
  const float* coords;
  float* out;
  int Q;
  int L;
  int r;
  // fused coordinate update (raftk_corr_lookup_update; fast path only): coords_rw += delta + dbias first
  const float* delta;   // NULL: plain lookup
  int ld_delta;
  const float* dbias;
  const float* coords0;
  float* coords_rw;
  float* flow;
  float* flow_pitched;  // may be NULL
  int ld_flow;
};

// Generic staging: per level the (2r+2)^2 footprint as [l][2r+2][2r+2], zeros outside the level.
__device__ __forceinline__ void stage_footprints_generic(const LookupParams& p, int q, float cx, float cy, int r,
                                                         float* F, int lane) {
  const int FP = 2 * r + 2;
  float sc = 1.0f;
#pragma unroll 1
  for (int l = 0; l < p.L; ++l, sc *= 0.5f) {
    const int lh = p.lh[l], lw = p.lw[l];
    const int bx = clamp_to_int(floorf(cx * sc)) - r;
    const int by = clamp_to_int(floorf(cy * sc)) - r;
    const float* slab = p.level[l] + (size_t)q * lh * lw;
    float* Fl = F + l * FP * FP;
#pragma unroll 4
    for (int e = lane; e < FP * FP; e += 32) {
      const int ry = e / FP, rx = e - ry * FP;
      const int gy = by + ry, gx = bx + rx;
      float v = 0.0f;
      if ((unsigned)gy < (unsigned)lh && (unsigned)gx < (unsigned)lw) v = __ldg(slab + gy * lw + gx);
      Fl[e] = v;
    }
  }
}

constexpr int kLookupWarps = 8;

// R > 0: compile-time radius; R == 0: radius from params (generic, slower).
template <int R>
__global__ void __launch_bounds__(kLookupWarps * 32) corr_lookup_kernel(const LookupParams p) {
  extern __shared__ float smem[];
  const int r = R > 0 ? R : p.r;
  const int S = 2 * r + 1;   // taps per side
  const int FP = 2 * r + 2;  // footprint side
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = blockIdx.x * kLookupWarps + warp;
  if (q >= p.Q) return;
  float* F = smem + warp * (p.L * FP * FP);

  const float cx = __ldg(p.coords + 2 * (size_t)q);
  const float cy = __ldg(p.coords + 2 * (size_t)q + 1);

  stage_footprints_generic(p, q, cx, cy, r, F, lane);
  __syncwarp();

  // ---- phase 2: taps ----
  lookup_taps<R>(F, cx, cy, p.L, r, p.out + (size_t)q * (p.L * S * S), lane);
}

// ------------------------------------------------------------------------------------------
// Fast path: L = 4, r in {3, 4}.  Levels whose row pitch is a multiple of 4 floats (and whose base
// is 16-byte aligned: 440x1024, 1088x1920 levels 0-2, ...) are read as 16-byte chunks, the others
// (e.g. every level of raft_small's 46x62 map) as the same chunks in four 4-byte loads.
//
// Persistent warps, one query at a time, software-pipelined: the chunk loads of a warp's next query
// are in flight while the taps of the current one are evaluated.  Phase 1 fetches each footprint row as the aligned
// 16-byte chunks that cover it (3 or 4 per row) — 5 LDG.128 per lane per query instead of 13
// scalar loads, touching the same 32-byte sectors — into a [2r+3][20] tile per level (16 used columns) whose
// column origin is floor4(x0): chunks outside the level or past the footprint are written as
// zeros, and a guard row below the footprint stays zero, so phase 2 needs no bounds checks.
// Phase 2 evaluates taps lane, lane+32, lane+64 of each level with per-lane tap constants
// hoisted out of the query loop; arithmetic is the same per-tap float32 sequence as
// lookup_taps (bit-identical results).  Coordinates beyond +-1e6 or non-finite take the generic
// per-query routine (warp-uniform branch).
// ------------------------------------------------------------------------------------------
// One footprint chunk: floats [gx, gx+4) of row gy of a level slab (gx % 4 == 0), zeros where the
// chunk leaves the level.  Aligned levels use one 16-byte load; others four 4-byte loads.
__device__ __forceinline__ float4 load_chunk(const float* slab, int lh, int lw, int gy, int gx, bool aligned,
                                             uint64_t pol) {
  float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  if ((unsigned)gy >= (unsigned)lh) return v;
  const float* row = slab + (size_t)gy * lw;
  if (aligned) {
    if ((unsigned)gx < (unsigned)lw) v = ldg_f4_l2hint(reinterpret_cast<const float4*>(row + gx), pol);
  } else {
    if ((unsigned)gx < (unsigned)lw) v.x = __ldg(row + gx);
    if ((unsigned)(gx + 1) < (unsigned)lw) v.y = __ldg(row + gx + 1);
    if ((unsigned)(gx + 2) < (unsigned)lw) v.z = __ldg(row + gx + 2);
    if ((unsigned)(gx + 3) < (unsigned)lw) v.w = __ldg(row + gx + 3);
  }
  return v;
}

constexpr int kFastWarps = 8;

// Footprint chunks of one query, in registers between the loads and the shared-memory tile.
struct Chunks {
  float4 v[5];
};

template <int R>
__global__ void __launch_bounds__(kFastWarps * 32) corr_lookup_fast_kernel(const LookupParams p) {
  // PITCH 20 (not 16): rows land 20 banks apart, so the taps of a warp (9 rows x 4 columns) read
  // 32 different banks; 16 would fold them onto 8
  constexpr int S = 2 * R + 1, FP = 2 * R + 2, ROWS = FP + 1, PITCH = 20, TILE = ROWS * PITCH;
  constexpr int NTAP = (S * S + 31) / 32;
  extern __shared__ float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* F = smem + warp * (4 * TILE);
  pdl_wait();  // the coordinates come from the kernel before this one
  // guard row of each level tile (never written again); all other cells are rewritten per query
  for (int e = lane; e < 4 * PITCH; e += 32) F[(e / PITCH) * TILE + FP * PITCH + (e % PITCH)] = 0.0f;
  __syncwarp();

  // tap constants (independent of the query)
  // (a lane whose third tap does not exist evaluates tap 0 again and skips the store)
  float di[NTAP], dj[NTAP];
#pragma unroll
  for (int k = 0; k < NTAP; ++k) {
    const int t = lane + 32 * k < S * S ? lane + 32 * k : 0;
    const int i = t / S, j = t - i * S;
    di[k] = (float)(i - R);
    dj[k] = (float)(j - R);
  }
  const bool last_tap_valid = lane + 32 * (NTAP - 1) < S * S;
  // load-task constants: rows 0..7 of level k in iteration k; for r = 4 the last two rows of
  // all four levels share a fifth iteration (8 lanes per level)
  const int row_u = lane >> 2, chunk = lane & 3;
  const int lvx = lane >> 3, rowx = 8 + ((lane >> 2) & 1);
  // this lane's level for that fifth iteration, selected once (no dynamic indexing of params)
  const float* levelx = lvx == 0 ? p.level[0] : lvx == 1 ? p.level[1] : lvx == 2 ? p.level[2] : p.level[3];
  const int lhx = lvx == 0 ? p.lh[0] : lvx == 1 ? p.lh[1] : lvx == 2 ? p.lh[2] : p.lh[3];
  const int lwx = lvx == 0 ? p.lw[0] : lvx == 1 ? p.lw[1] : lvx == 2 ? p.lw[2] : p.lw[3];
  const bool alignedx = (lvx == 0 ? p.aligned[0] : lvx == 1 ? p.aligned[1] : lvx == 2 ? p.aligned[2] : p.aligned[3]) != 0;
  constexpr int C = 4 * S * S;  // the fast path is instantiated for L = 4 only
  const uint64_t pol = l2_policy_evict_last();
  const int nwarps = gridDim.x * kFastWarps;

  // phase 1a: the aligned 16-byte chunks covering query q's four footprints -> registers
  auto fetch = [&](int q, float cx, float cy, Chunks& c) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      c.v[k] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
      const float sc = __int_as_float((127 - k) << 23);
      const int bx = (int)floorf(cx * sc) - R, by = (int)floorf(cy * sc) - R;
      const int lh = p.lh[k], lw = p.lw[k];
      const int gy = by + row_u, gx = (bx & ~3) + 4 * chunk;
      if (row_u < FP && gx <= bx + FP - 1)
        c.v[k] = load_chunk(p.level[k] + (size_t)q * lh * lw, lh, lw, gy, gx, p.aligned[k] != 0, pol);
    }
    c.v[4] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (FP > 8) {
      const float sc = __int_as_float((127 - lvx) << 23);
      const int bx = (int)floorf(cx * sc) - R, by = (int)floorf(cy * sc) - R;
      const int gy = by + rowx, gx = (bx & ~3) + 4 * chunk;
      if (gx <= bx + FP - 1) c.v[4] = load_chunk(levelx + (size_t)q * lhx * lwx, lhx, lwx, gy, gx, alignedx, pol);
    }
  };
  auto near = [](float cx, float cy) { return fabsf(cx) <= 1.0e6f && fabsf(cy) <= 1.0e6f; };

  // Software pipeline over this warp's queries: the chunk loads of query q + nwarps are issued before
  // the taps of query q are evaluated, so DRAM latency and tap arithmetic overlap inside one warp
  // (at batch 1 every warp has only one or two queries and no other warp's phase to hide behind).
  // the query's coordinates; with the fused update (UpdateCell's tail of the PREVIOUS iteration,
  // model.py:504-506) they are first advanced, coords1 += delta + bias, and written back with the flow
  auto load_coords = [&](int qq, float& x, float& y) {
    if (p.delta == nullptr) {
      x = __ldg(p.coords + 2 * (size_t)qq);
      y = __ldg(p.coords + 2 * (size_t)qq + 1);
      return;
    }
    // lane 0 alone reads, advances and writes the query's coordinates (one reader-writer per address), then
    // hands them to the warp
    x = 0.0f;
    y = 0.0f;
    if (lane == 0) {
      const float2 c1 = *reinterpret_cast<const float2*>(p.coords_rw + 2 * (size_t)qq);
      const float2 d = *reinterpret_cast<const float2*>(p.delta + (size_t)qq * p.ld_delta);
      const float2 b = __ldg(reinterpret_cast<const float2*>(p.dbias));
      const float2 c0 = __ldg(reinterpret_cast<const float2*>(p.coords0) + qq);
      x = __fadd_rn(c1.x, __fadd_rn(d.x, b.x));
      y = __fadd_rn(c1.y, __fadd_rn(d.y, b.y));
      const float2 f = make_float2(__fadd_rn(x, -c0.x), __fadd_rn(y, -c0.y));
      *reinterpret_cast<float2*>(p.coords_rw + 2 * (size_t)qq) = make_float2(x, y);
      *reinterpret_cast<float2*>(p.flow + 2 * (size_t)qq) = f;
      if (p.flow_pitched) *reinterpret_cast<float2*>(p.flow_pitched + (size_t)qq * p.ld_flow) = f;
    }
    x = __shfl_sync(0xffffffffu, x, 0);
    y = __shfl_sync(0xffffffffu, y, 0);
  };
  int q = blockIdx.x * kFastWarps + warp;
  float cx = 0.0f, cy = 0.0f;
  Chunks cur;
  if (q < p.Q) {
    load_coords(q, cx, cy);
    if (near(cx, cy)) fetch(q, cx, cy, cur);
  }
  while (q < p.Q) {
    const int qn = q + nwarps;
    const bool fast = near(cx, cy);
    if (fast) {
      // ---- phase 1b: registers -> shared-memory tiles ----
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (row_u < FP) *reinterpret_cast<float4*>(F + k * TILE + row_u * PITCH + chunk * 4) = cur.v[k];
      if (FP > 8) *reinterpret_cast<float4*>(F + lvx * TILE + rowx * PITCH + chunk * 4) = cur.v[4];
      __syncwarp();
    }
    // ---- prefetch the next query of this warp ----
    float ncx = 0.0f, ncy = 0.0f;
    if (qn < p.Q) {
      load_coords(qn, ncx, ncy);
      if (near(ncx, ncy)) fetch(qn, ncx, ncy, cur);
    }
    if (!fast) {
      // far outside or non-finite: generic staging (its own layout), then restore the zero tile
      stage_footprints_generic(p, q, cx, cy, R, F, lane);
      __syncwarp();
      lookup_taps<R>(F, cx, cy, p.L, R, p.out + (size_t)q * C, lane);
      __syncwarp();
      for (int e = lane; e < 4 * TILE; e += 32) F[e] = 0.0f;  // generic layout overwrote the guard rows
      __syncwarp();
    } else {
      // ---- phase 2: taps ----
      float* obase = p.out + (size_t)q * C + lane;
#pragma unroll
      for (int l = 0; l < 4; ++l) {
        const float sc = __int_as_float((127 - l) << 23);
        const float cxl = cx * sc, cyl = cy * sc;
        const float fbxa = (float)(((int)floorf(cxl) - R) & ~3);  // tile column origin
        const float fby = floorf(cyl) - (float)R;                 // tile row origin
        const float* Fl = F + l * TILE;
#pragma unroll
        for (int k = 0; k < NTAP; ++k) {
          const float sx = __fadd_rn(cxl, di[k]), sy = __fadd_rn(cyl, dj[k]);
          const float fx0 = floorf(sx), fy0 = floorf(sy);
          const float wx1 = __fadd_rn(sx, -fx0), wy1 = __fadd_rn(sy, -fy0);
          const float wx0 = __fadd_rn(1.0f, -wx1), wy0 = __fadd_rn(1.0f, -wy1);
          const int off = __float2int_rn(fmaf(fy0 - fby, (float)PITCH, fx0 - fbxa));
          const float* f = Fl + off;
          float acc = __fmul_rn(__fmul_rn(wy0, wx0), f[0]);
          acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy0, wx1), f[1]));
          acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx0), f[PITCH]));
          acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx1), f[PITCH + 1]));
          if (k < NTAP - 1 || last_tap_valid) obase[l * S * S + 32 * k] = acc;
        }
      }
      __syncwarp();
    }
    q = qn;
    cx = ncx;
    cy = ncy;
  }
}

}  // namespace raftk

struct raftk_lookup_update_args {
  const float* delta;
  int ld_delta;
  const float* bias;
  const float* coords0;
  float* coords1;
  float* flow;
  float* flow_pitched;
  int ld_flow;
};

static int corr_lookup_launch(const float* const* levels, int num_levels, const float* coords, float* out, int Q, int h,
                              int w, int radius, const raftk_lookup_update_args* upd, void* stream) {
  using namespace raftk;
  if (Q < 0 || h <= 0 || w <= 0 || radius < 0 || num_levels < 1 || num_levels > RAFTK_MAX_LEVELS)
    return RAFTK_ERR_SHAPE;
  if ((h >> (num_levels - 1)) < 1 || (w >> (num_levels - 1)) < 1) return RAFTK_ERR_TOO_SMALL;
  if (radius > 7) return RAFTK_ERR_UNSUPPORTED;
  if (Q == 0) return RAFTK_OK;  // empty batch: nothing to read, pointers may be NULL
  if (!levels || !coords || !out) return RAFTK_ERR_NULL;
  LookupParams p{};
  for (int l = 0; l < num_levels; ++l) {
    if (!levels[l]) return RAFTK_ERR_NULL;
    p.level[l] = levels[l];
    p.lh[l] = h >> l;
    p.lw[l] = w >> l;
  }
  p.coords = coords;
  p.out = out;
  p.Q = Q;
  p.L = num_levels;
  p.r = radius;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool fast = (radius == 3 || radius == 4) && num_levels == 4;
  if (upd) {
    if (!fast) return RAFTK_ERR_UNSUPPORTED;
    p.delta = upd->delta;
    p.ld_delta = upd->ld_delta;
    p.dbias = upd->bias;
    p.coords0 = upd->coords0;
    p.coords_rw = upd->coords1;
    p.flow = upd->flow;
    p.flow_pitched = upd->flow_pitched;
    p.ld_flow = upd->ld_flow;
  }
  for (int l = 0; l < num_levels; ++l)
    p.aligned[l] = (p.lw[l] % 4 == 0) && (((uintptr_t)p.level[l] & 15) == 0);
  if (fast) {
    const int tile = (2 * radius + 3) * 20;  // ROWS * PITCH of the kernel
    const size_t smem = (size_t)kFastWarps * 4 * tile * sizeof(float);
    // two queries per warp even at batch 1, so the kernel's software pipeline has something to overlap
    // (measured: 8.5 us against 9.0 us with one); persistent beyond 8 blocks per SM
    int blocks = (Q + 2 * kFastWarps - 1) / (2 * kFastWarps);
    if (blocks > RAFTK_SM_COUNT * 8) blocks = RAFTK_SM_COUNT * 8;
    if (radius == 4) return launch_pdl(corr_lookup_fast_kernel<4>, blocks, kFastWarps * 32, smem, s, p);
    return launch_pdl(corr_lookup_fast_kernel<3>, blocks, kFastWarps * 32, smem, s, p);
  }
  const int FP = 2 * radius + 2;
  const size_t smem = (size_t)kLookupWarps * num_levels * FP * FP * sizeof(float);
  const int blocks = (Q + kLookupWarps - 1) / kLookupWarps;
  if (radius == 4)
    corr_lookup_kernel<4><<<blocks, kLookupWarps * 32, smem, s>>>(p);
  else if (radius == 3)
    corr_lookup_kernel<3><<<blocks, kLookupWarps * 32, smem, s>>>(p);
  else
    corr_lookup_kernel<0><<<blocks, kLookupWarps * 32, smem, s>>>(p);
  return launch_status();
}

extern "C" int raftk_corr_lookup(const float* const* levels, int num_levels, const float* coords,
                                 float* out, int Q, int h, int w, int radius, void* stream) {
  return corr_lookup_launch(levels, num_levels, coords, out, Q, h, w, radius, nullptr, stream);
}

extern "C" int raftk_corr_lookup_update(const float* const* levels, int num_levels, const float* delta_pre, int ld_delta,
                                        const float* bias, const float* coords0, float* coords1, float* flow,
                                        float* flow_pitched, int ld_flow, float* out, int Q, int h, int w, int radius,
                                        void* stream) {
  if (Q > 0) {
    if (!delta_pre || !bias || !coords0 || !coords1 || !flow) return RAFTK_ERR_NULL;
    if (ld_delta < 2 || ld_delta % 2 || (flow_pitched && (ld_flow < 2 || ld_flow % 2))) return RAFTK_ERR_SHAPE;
    if (((uintptr_t)delta_pre | (uintptr_t)bias | (uintptr_t)coords0 | (uintptr_t)coords1 | (uintptr_t)flow |
         (uintptr_t)flow_pitched) & 7)
      return RAFTK_ERR_ALIGN;
  }
  const raftk_lookup_update_args upd{delta_pre, ld_delta, bias, coords0, coords1, flow, flow_pitched, ld_flow};
  return corr_lookup_launch(levels, num_levels, coords1, out, Q, h, w, radius, &upd, stream);
}
