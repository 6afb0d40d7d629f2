// Kernel family 2: radius-r bilinear window lookup over the correlation pyramid.
// Replaces CorrBlock.index_pyramid (model.py:448-470) + grid_sample (model.py:24-34).
//
// One warp per query pixel.  Phase 1 stages, per level, the (2r+2)x(2r+2) footprint
// that all (2r+1)^2 taps of that level touch into shared memory with lanes running along
// footprint rows (coalesced 4-byte loads; rows are 4-byte aligned only).  Phase 2 gives
// each lane output channels lane, lane+32, ... so the 4*(2r+1)^2*4-byte output row is
// written with full 128-byte warp stores.
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
  const float* coords;
  float* out;
  int Q;
  int L;
  int r;
};

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

  // ---- phase 1: footprints -> shared memory ----
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
  __syncwarp();

  // ---- phase 2: taps ----
  lookup_taps<R>(F, cx, cy, p.L, r, p.out + (size_t)q * (p.L * S * S), lane);
}

}  // namespace raftk

extern "C" int raftk_corr_lookup(const float* const* levels, int num_levels, const float* coords,
                                 float* out, int Q, int h, int w, int radius, void* stream) {
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
  const int FP = 2 * radius + 2;
  const size_t smem = (size_t)kLookupWarps * num_levels * FP * FP * sizeof(float);
  const int blocks = (Q + kLookupWarps - 1) / kLookupWarps;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (radius == 4)
    corr_lookup_kernel<4><<<blocks, kLookupWarps * 32, smem, s>>>(p);
  else if (radius == 3)
    corr_lookup_kernel<3><<<blocks, kLookupWarps * 32, smem, s>>>(p);
  else
    corr_lookup_kernel<0><<<blocks, kLookupWarps * 32, smem, s>>>(p);
  return launch_status();
}
