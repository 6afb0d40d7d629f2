// grid_sample: bilinear sampling of an NHWC image at absolute pixel coordinates, zero outside.
// Replaces grid_sample (model.py:24-34), i.e. jax.scipy.ndimage.map_coordinates(order=1,
// mode='constant') vmapped over batch and channel.  The correlation lookup has this fused
// (lookup_taps.cuh); this entry point is the stand-alone form of the reference's helper.
//
// One thread per (sample, channel), scalar loads: C = 1 on the correlation path (one pyramid level), where there
// is nothing to vectorise over; a wide-C caller pays strided loads.
// Arithmetic per sample, float32, no FMA contraction, as the reference:
//   i0 = floor(c); w1 = c - i0; w0 = 1 - w1;
//   out = (((wy0*wx0)*v00 + (wy0*wx1)*v01) + (wy1*wx0)*v10) + (wy1*wx1)*v11,  v = 0 outside.

#include "lookup_taps.cuh"
#include "raftk_common.cuh"

namespace raftk {

__global__ void __launch_bounds__(256)
grid_sample_kernel(const float* __restrict__ img, const float* __restrict__ grid, float* __restrict__ out,
                   long long n_samples, int samples_per_image, int h, int w, int c) {
  const long long total = n_samples * c;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const long long s = t / c;
    const int ch = (int)(t - s * c);
    const long long n = s / samples_per_image;
    const float gx = __ldg(grid + 2 * s), gy = __ldg(grid + 2 * s + 1);  // (x, y): model.py:28 flips to (y, x)
    const float fx0 = floorf(gx), fy0 = floorf(gy);
    const float wx1 = __fadd_rn(gx, -fx0), wy1 = __fadd_rn(gy, -fy0);
    const float wx0 = __fadd_rn(1.0f, -wx1), wy0 = __fadd_rn(1.0f, -wy1);
    const int x0 = clamp_to_int(fx0), y0 = clamp_to_int(fy0);
    const float* base = img + n * (long long)h * w * c + ch;
    const bool vx0 = (unsigned)x0 < (unsigned)w, vx1 = (unsigned)(x0 + 1) < (unsigned)w;
    const bool vy0 = (unsigned)y0 < (unsigned)h, vy1 = (unsigned)(y0 + 1) < (unsigned)h;
    const float v00 = (vy0 && vx0) ? __ldg(base + ((long long)y0 * w + x0) * c) : 0.0f;
    const float v01 = (vy0 && vx1) ? __ldg(base + ((long long)y0 * w + x0 + 1) * c) : 0.0f;
    const float v10 = (vy1 && vx0) ? __ldg(base + ((long long)(y0 + 1) * w + x0) * c) : 0.0f;
    const float v11 = (vy1 && vx1) ? __ldg(base + ((long long)(y0 + 1) * w + x0 + 1) * c) : 0.0f;
    float acc = __fmul_rn(__fmul_rn(wy0, wx0), v00);
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy0, wx1), v01));
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx0), v10));
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx1), v11));
    out[t] = acc;
  }
}

}  // namespace raftk

extern "C" int raftk_grid_sample(const float* img, const float* grid, float* out, int n, int h, int w, int c,
                                 int gh, int gw, void* stream) {
  using namespace raftk;
  if (n < 0 || h <= 0 || w <= 0 || c <= 0 || gh < 0 || gw < 0) return RAFTK_ERR_SHAPE;
  const long long n_samples = (long long)n * gh * gw;
  if (n_samples == 0) return RAFTK_OK;
  if (!img || !grid || !out) return RAFTK_ERR_NULL;
  const long long total = n_samples * c;
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)RAFTK_SM_COUNT * 32) blocks = (long long)RAFTK_SM_COUNT * 32;
  grid_sample_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(img, grid, out, n_samples,
                                                                                    gh * gw, h, w, c);
  return launch_status();
}
