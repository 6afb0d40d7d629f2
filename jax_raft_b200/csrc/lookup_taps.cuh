// Tap evaluation shared by the volume lookup and the on-the-fly (alt_corr) lookup.
// Reference: CorrBlock.index_pyramid (model.py:448-470) calling grid_sample (model.py:24-34).
#pragma once

#include "raftk_common.cuh"

namespace raftk {

// Coordinates far outside any level select nothing; clamp so float->int is defined.
__device__ __forceinline__ int clamp_to_int(float f) {
  return static_cast<int>(fminf(fmaxf(f, -1048576.0f), 1048576.0f));
}

// One warp evaluates all L*(2r+1)^2 taps of one query from the staged footprints
// F[l][(2r+2)][(2r+2)] (footprint origin = floor(c_l) - r, zeros outside the level) and
// writes them to orow[c], c = l*(2r+1)^2 + i*(2r+1) + j with i the x offset.
//
// Per tap, in float32 with no FMA contraction (jax map_coordinates order=1):
//   s = c_l + d;  i0 = floor(s);  w1 = s - i0;  w0 = 1 - w1
//   out = (((wy0*wx0)*v00 + (wy0*wx1)*v01) + (wy1*wx0)*v10) + (wy1*wx1)*v11
// fl(c_l + d) can round UP to an integer, so i0 is floor(c_l)+d or floor(c_l)+d+1; in
// the second case w1 == 0 exactly and the one corner that may leave the footprint has
// weight 0, so reading it as 0 leaves the sum unchanged.
template <int R>
__device__ __forceinline__ void lookup_taps(const float* F, float cx, float cy, int L, int r_rt,
                                            float* orow, int lane) {
  const int r = R > 0 ? R : r_rt;
  const int S = 2 * r + 1, FP = 2 * r + 2;
  const int C = L * S * S;
  for (int c = lane; c < C; c += 32) {
    const int l = c / (S * S);
    const int k = c - l * S * S;
    const int i = k / S, j = k - i * S;                 // i: x offset (slow), j: y offset (fast)
    const float scl = __int_as_float((127 - l) << 23);  // 2^-l, exact
    const float cxl = cx * scl, cyl = cy * scl;
    const int bx = clamp_to_int(floorf(cxl)) - r;
    const int by = clamp_to_int(floorf(cyl)) - r;
    const float sx = __fadd_rn(cxl, (float)(i - r));
    const float sy = __fadd_rn(cyl, (float)(j - r));
    const float fx0 = floorf(sx), fy0 = floorf(sy);
    const float wx1 = __fadd_rn(sx, -fx0), wy1 = __fadd_rn(sy, -fy0);
    const float wx0 = __fadd_rn(1.0f, -wx1), wy0 = __fadd_rn(1.0f, -wy1);
    const int rx = clamp_to_int(fx0) - bx, ry = clamp_to_int(fy0) - by;
    const float* Fl = F + l * FP * FP;
    const bool x0in = (unsigned)rx < (unsigned)FP, x1in = (unsigned)(rx + 1) < (unsigned)FP;
    const bool y0in = (unsigned)ry < (unsigned)FP, y1in = (unsigned)(ry + 1) < (unsigned)FP;
    const float v00 = (y0in && x0in) ? Fl[ry * FP + rx] : 0.0f;
    const float v01 = (y0in && x1in) ? Fl[ry * FP + rx + 1] : 0.0f;
    const float v10 = (y1in && x0in) ? Fl[(ry + 1) * FP + rx] : 0.0f;
    const float v11 = (y1in && x1in) ? Fl[(ry + 1) * FP + rx + 1] : 0.0f;
    float acc = __fmul_rn(__fmul_rn(wy0, wx0), v00);
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy0, wx1), v01));
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx0), v10));
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx1), v11));
    orow[c] = acc;
  }
}

// The taps of ONE level l from that level's staged footprint Fl[(2r+2)][(2r+2)]; same arithmetic
// as lookup_taps, channels l*(2r+1)^2 ... of orow.
template <int R>
__device__ __forceinline__ void lookup_taps_level(const float* Fl, float cx, float cy, int l, float* orow, int lane) {
  constexpr int S = 2 * R + 1, FP = 2 * R + 2;
  const float scl = __int_as_float((127 - l) << 23);  // 2^-l, exact
  const float cxl = cx * scl, cyl = cy * scl;
  const int bx = clamp_to_int(floorf(cxl)) - R;
  const int by = clamp_to_int(floorf(cyl)) - R;
  for (int k = lane; k < S * S; k += 32) {
    const int i = k / S, j = k - i * S;
    const float sx = __fadd_rn(cxl, (float)(i - R));
    const float sy = __fadd_rn(cyl, (float)(j - R));
    const float fx0 = floorf(sx), fy0 = floorf(sy);
    const float wx1 = __fadd_rn(sx, -fx0), wy1 = __fadd_rn(sy, -fy0);
    const float wx0 = __fadd_rn(1.0f, -wx1), wy0 = __fadd_rn(1.0f, -wy1);
    const int rx = clamp_to_int(fx0) - bx, ry = clamp_to_int(fy0) - by;
    const bool x0in = (unsigned)rx < (unsigned)FP, x1in = (unsigned)(rx + 1) < (unsigned)FP;
    const bool y0in = (unsigned)ry < (unsigned)FP, y1in = (unsigned)(ry + 1) < (unsigned)FP;
    const float v00 = (y0in && x0in) ? Fl[ry * FP + rx] : 0.0f;
    const float v01 = (y0in && x1in) ? Fl[ry * FP + rx + 1] : 0.0f;
    const float v10 = (y1in && x0in) ? Fl[(ry + 1) * FP + rx] : 0.0f;
    const float v11 = (y1in && x1in) ? Fl[(ry + 1) * FP + rx + 1] : 0.0f;
    float acc = __fmul_rn(__fmul_rn(wy0, wx0), v00);
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy0, wx1), v01));
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx0), v10));
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx1), v11));
    orow[l * S * S + k] = acc;
  }
}

}  // namespace raftk
