// Kernel family 3: 9-neighbour softmax convex 8x flow upsampling.
// Replaces the up_mask branch of upsample_flow (model.py:85-98).
//
// Half a warp per coarse pixel.  The 576 mask logits of a pixel are 9 planes (k) of 64
// sub-pixel values (fy*8+fx); lane s of the half-warp owns sub-pixels 4s..4s+3, so each of the
// nine plane reads is one 256-byte coalesced LDG.128 per half-warp (all nine in flight at once).
// The softmax over k (exp(x - max) / sum, as flax) and the convex combination stay in
// registers: one reciprocal per sub-pixel, ex2-based exponentials.  A lane's four sub-pixels x
// two flow channels are 32 contiguous output bytes; two lanes fill a 64-byte row segment.
//
// Memory-bound: 2304 B mask + 8 B flow in, 512 B out per coarse pixel.

#include "raftk_common.cuh"

namespace raftk {

constexpr int kUpsWarps = 8;

// HAS_BIAS: the logits are mask + bias[channel] (the mask predictor's last convolution without its
// bias pass; the sum is rounded exactly as a separate bias kernel would round it).
template <bool HAS_BIAS>
__global__ void __launch_bounds__(kUpsWarps * 32)
convex_upsample_kernel(const float* __restrict__ flow, const float* __restrict__ mask, const float* __restrict__ bias,
                       float* __restrict__ out, int B, int h, int w) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sub = lane & 15;  // owns sub-pixels 4*sub .. 4*sub+3
  const long long npix = (long long)B * h * w;
  const long long pix = ((long long)blockIdx.x * kUpsWarps + warp) * 2 + (lane >> 4);
  if (pix >= npix) return;
  const int x = (int)(pix % w);
  const int y = (int)((pix / w) % h);
  const int b = (int)(pix / ((long long)w * h));

  // mask logits: 9 x float4 per lane
  const float4* mrow = reinterpret_cast<const float4*>(mask + pix * 576) + sub;
  float4 m[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) m[k] = ldg_nc_f4(mrow + k * 16);
  if (HAS_BIAS) {
    const float4* brow = reinterpret_cast<const float4*>(bias) + sub;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const float4 bk = __ldg(brow + k * 16);
      m[k].x += bk.x; m[k].y += bk.y; m[k].z += bk.z; m[k].w += bk.w;
    }
  }

  // 8 * flow of the 9 neighbours (zero outside the image: SAME padding, model.py:91-92)
  float fxk[9], fyk[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
    float2 f = make_float2(0.0f, 0.0f);
    if ((unsigned)yy < (unsigned)h && (unsigned)xx < (unsigned)w)
      f = __ldg(reinterpret_cast<const float2*>(flow) + ((long long)b * h + yy) * w + xx);
    fxk[k] = 8.0f * f.x;
    fyk[k] = 8.0f * f.y;
  }

  float o[8];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float v[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) v[k] = c == 0 ? m[k].x : c == 1 ? m[k].y : c == 2 ? m[k].z : m[k].w;
    float mx = v[0];
#pragma unroll
    for (int k = 1; k < 9; ++k) mx = fmaxf(mx, v[k]);
    float s = 0.0f, ax = 0.0f, ay = 0.0f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const float e = __expf(v[k] - mx);
      s += e;
      ax = fmaf(e, fxk[k], ax);
      ay = fmaf(e, fyk[k], ay);
    }
    const float inv = 1.0f / s;
    o[2 * c] = ax * inv;
    o[2 * c + 1] = ay * inv;
  }
  // sub-pixels 4*sub.. -> (fy, fx) = (sub / 2, (sub % 2) * 4)
  const int fy = sub >> 1, fx = (sub & 1) * 4;
  const long long W8 = 8LL * w;
  float* orow = out + (((long long)b * 8 * h + 8 * y + fy) * W8 + 8 * x + fx) * 2;
  st_na_f4(orow, make_float4(o[0], o[1], o[2], o[3]));
  st_na_f4(orow + 4, make_float4(o[4], o[5], o[6], o[7]));
}

}  // namespace raftk

static int convex_upsample_launch(const float* flow, const float* mask, const float* bias, bool has_bias, float* out,
                                  int B, int h, int w, void* stream) {
  using namespace raftk;
  if (B < 0 || h <= 0 || w <= 0) return RAFTK_ERR_SHAPE;
  if (B == 0) return RAFTK_OK;  // empty batch: pointers may be NULL
  if (!flow || !mask || !out || (has_bias && !bias)) return RAFTK_ERR_NULL;
  if (((uintptr_t)flow & 7) || ((uintptr_t)mask & 15) || ((uintptr_t)out & 15) || ((uintptr_t)bias & 15))
    return RAFTK_ERR_ALIGN;
  const long long npix = (long long)B * h * w;
  if (npix == 0) return RAFTK_OK;
  const long long blocks = (npix + 2 * kUpsWarps - 1) / (2 * kUpsWarps);
  if (blocks > 0x7fffffffLL) return RAFTK_ERR_SHAPE;
  auto s = static_cast<cudaStream_t>(stream);
  if (has_bias)
    convex_upsample_kernel<true><<<(unsigned)blocks, kUpsWarps * 32, 0, s>>>(flow, mask, bias, out, B, h, w);
  else
    convex_upsample_kernel<false><<<(unsigned)blocks, kUpsWarps * 32, 0, s>>>(flow, mask, nullptr, out, B, h, w);
  return launch_status();
}

extern "C" int raftk_convex_upsample(const float* flow, const float* mask, float* out, int B,
                                     int h, int w, void* stream) {
  return convex_upsample_launch(flow, mask, nullptr, false, out, B, h, w, stream);
}

extern "C" int raftk_convex_upsample_bias(const float* flow, const float* mask, const float* mask_bias, float* out,
                                          int B, int h, int w, void* stream) {
  return convex_upsample_launch(flow, mask, mask_bias, true, out, B, h, w, stream);
}
