/*
 * CPU oracle in C (OpenMP) — TEST INFRASTRUCTURE, not shipped.
 *
 * A multi-threaded restatement of the same reference functions as oracle/corr_oracle.py, used as
 * the timed CPU baseline of bench.py (a NumPy gather is single-threaded and would flatter the
 * GPU/CPU ratio) and cross-checked against the NumPy oracle in tests/test_oracle.py.
 * Citations are into /root/reference/jax_raft/model.py.  Float32 throughout, compiled with
 * -ffp-contract=off so the arithmetic order below is the arithmetic performed.
 *
 * Build: gcc -O3 -fopenmp -ffp-contract=off -shared -fPIC -o oracle/_build/libcorr_oracle.so oracle/corr_oracle.c -lm
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>

/* model.py:444 — nn.avg_pool 2x2 stride 2 VALID on (n, h, w): ((a + b) + (c + d)) / 4. */
void oracle_avg_pool2(const float* in, float* out, long n, int h, int w) {
  const int oh = h / 2, ow = w / 2;
#pragma omp parallel for schedule(static)
  for (long s = 0; s < n; ++s) {
    const float* p = in + (size_t)s * h * w;
    float* o = out + (size_t)s * oh * ow;
    for (int y = 0; y < oh; ++y)
      for (int x = 0; x < ow; ++x) {
        const float* q = p + (size_t)(2 * y) * w + 2 * x;
        o[(size_t)y * ow + x] = ((q[0] + q[1]) + (q[w] + q[w + 1])) / 4.0f;
      }
  }
}

/* model.py:24-34 via jax.scipy.ndimage.map_coordinates(order=1, mode='constant'):
 * one bilinear sample of a (h, w) slab at (x, y), each corner zero outside. */
static inline float sample(const float* slab, int h, int w, float sx, float sy) {
  const float fx0 = floorf(sx), fy0 = floorf(sy);
  const float wx1 = sx - fx0, wy1 = sy - fy0;
  const float wx0 = 1.0f - wx1, wy0 = 1.0f - wy1;
  /* clamp before the int conversion; anything this far out is outside every level */
  const float cx = fminf(fmaxf(fx0, -1048576.0f), 1048576.0f);
  const float cy = fminf(fmaxf(fy0, -1048576.0f), 1048576.0f);
  const long x0 = (long)cx, y0 = (long)cy;
  const int vx0 = x0 >= 0 && x0 < w, vx1 = x0 + 1 >= 0 && x0 + 1 < w;
  const int vy0 = y0 >= 0 && y0 < h, vy1 = y0 + 1 >= 0 && y0 + 1 < h;
  const float v00 = (vy0 && vx0) ? slab[y0 * w + x0] : 0.0f;
  const float v01 = (vy0 && vx1) ? slab[y0 * w + x0 + 1] : 0.0f;
  const float v10 = (vy1 && vx0) ? slab[(y0 + 1) * w + x0] : 0.0f;
  const float v11 = (vy1 && vx1) ? slab[(y0 + 1) * w + x0 + 1] : 0.0f;
  float acc = (wy0 * wx0) * v00;
  acc = acc + (wy0 * wx1) * v01;
  acc = acc + (wy1 * wx0) * v10;
  acc = acc + (wy1 * wx1) * v11;
  return acc;
}

/* model.py:448-470 — CorrBlock.index_pyramid.  levels[l]: (Q, h>>l, w>>l); coords (Q, 2) as (x, y);
 * out (Q, L*(2r+1)^2), channel l*(2r+1)^2 + i*(2r+1) + j samples (x/2^l + i - r, y/2^l + j - r). */
void oracle_index_pyramid(const float* const* levels, int num_levels, const float* coords, float* out, long Q,
                          int h, int w, int radius) {
  const int S = 2 * radius + 1;
#pragma omp parallel for schedule(static)
  for (long q = 0; q < Q; ++q) {
    float cx = coords[2 * q], cy = coords[2 * q + 1];
    float* o = out + (size_t)q * num_levels * S * S;
    for (int l = 0; l < num_levels; ++l) {
      const int lh = h >> l, lw = w >> l;
      const float* slab = levels[l] + (size_t)q * lh * lw;
      for (int i = 0; i < S; ++i)
        for (int j = 0; j < S; ++j)
          o[l * S * S + i * S + j] = sample(slab, lh, lw, cx + (float)(i - radius), cy + (float)(j - radius));
      cx = cx / 2.0f; /* model.py:466 */
      cy = cy / 2.0f;
    }
  }
}

/* model.py:85-98 — convex upsampling, factor 8, 2 channels.  flow (B,h,w,2), mask (B,h,w,576),
 * out (B,8h,8w,2); softmax over k as flax: exp(x - max) / sum. */
void oracle_convex_upsample(const float* flow, const float* mask, float* out, int B, int h, int w) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int b = 0; b < B; ++b)
    for (int y = 0; y < h; ++y)
      for (int x = 0; x < w; ++x) {
        const float* m = mask + (((size_t)b * h + y) * w + x) * 576;
        float fx[9], fy[9];
        for (int k = 0; k < 9; ++k) {
          const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
          if (yy >= 0 && yy < h && xx >= 0 && xx < w) {
            const float* f = flow + (((size_t)b * h + yy) * w + xx) * 2;
            fx[k] = 8.0f * f[0];
            fy[k] = 8.0f * f[1];
          } else {
            fx[k] = 0.0f;
            fy[k] = 0.0f;
          }
        }
        for (int s = 0; s < 64; ++s) {
          float mx = m[s];
          for (int k = 1; k < 9; ++k) mx = fmaxf(mx, m[k * 64 + s]);
          float e[9], sum = 0.0f;
          for (int k = 0; k < 9; ++k) {
            e[k] = expf(m[k * 64 + s] - mx);
            sum += e[k];
          }
          float ox = 0.0f, oy = 0.0f;
          for (int k = 0; k < 9; ++k) {
            const float p = e[k] / sum;
            ox += p * fx[k];
            oy += p * fy[k];
          }
          const int sy = s / 8, sx = s % 8;
          float* o = out + ((((size_t)b * 8 * h + 8 * y + sy) * (8 * (size_t)w)) + 8 * x + sx) * 2;
          o[0] = ox;
          o[1] = oy;
        }
      }
}
