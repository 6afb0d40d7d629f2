// Kernel family 3: 9-neighbour softmax convex 8x flow upsampling.
// Replaces the up_mask branch of upsample_flow (model.py:85-98).
//
// One warp per coarse pixel.  The 576 mask logits of a pixel are 9 planes (k) of 64
// sub-pixel values (fy*8+fx); lane t owns sub-pixels 2t and 2t+1, so each of the nine
// plane reads is one 256-byte coalesced warp load (float2 per lane).  The softmax over
// k and the convex combination stay in registers.  A lane's two sub-pixels x two flow
// channels are 16 contiguous output bytes; four lanes fill one 64-byte row segment.
//
// Memory-bound: 2304 B mask + 8 B flow in, 512 B out per coarse pixel.

#include "raftk_common.cuh"

namespace raftk {

constexpr int kUpsWarps = 8;

__global__ void __launch_bounds__(kUpsWarps * 32)
convex_upsample_kernel(const float* __restrict__ flow, const float* __restrict__ mask,
                       float* __restrict__ out, int B, int h, int w) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long pix = (long long)blockIdx.x * kUpsWarps + warp;
  const long long npix = (long long)B * h * w;
  if (pix >= npix) return;
  const int x = (int)(pix % w);
  const int y = (int)((pix / w) % h);
  const int b = (int)(pix / ((long long)w * h));

  // mask logits: 9 x float2 per lane
  const float* mrow = mask + pix * 576 + 2 * lane;
  float2 m[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) m[k] = ldg_nc_f2(mrow + k * 64);

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

  // softmax over k (flax: exp(x - max) / sum), one per owned sub-pixel
  float mx0 = m[0].x, mx1 = m[0].y;
#pragma unroll
  for (int k = 1; k < 9; ++k) {
    mx0 = fmaxf(mx0, m[k].x);
    mx1 = fmaxf(mx1, m[k].y);
  }
  float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    m[k].x = expf(m[k].x - mx0);
    m[k].y = expf(m[k].y - mx1);
    s0 += m[k].x;
    s1 += m[k].y;
  }
  float4 o = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const float p0 = m[k].x / s0, p1 = m[k].y / s1;
    o.x += p0 * fxk[k];
    o.y += p0 * fyk[k];
    o.z += p1 * fxk[k];
    o.w += p1 * fyk[k];
  }
  // sub-pixel 2*lane -> (fy, fx) = (lane / 4, (lane % 4) * 2)
  const int fy = lane >> 2, fx = (lane & 3) * 2;
  const long long W8 = 8LL * w;
  float* orow = out + (((long long)b * 8 * h + 8 * y + fy) * W8 + 8 * x + fx) * 2;
  st_na_f4(orow, o);
}

}  // namespace raftk

extern "C" int raftk_convex_upsample(const float* flow, const float* mask, float* out, int B,
                                     int h, int w, void* stream) {
  using namespace raftk;
  if (B < 0 || h <= 0 || w <= 0) return RAFTK_ERR_SHAPE;
  if (B == 0) return RAFTK_OK;  // empty batch: pointers may be NULL
  if (!flow || !mask || !out) return RAFTK_ERR_NULL;
  if (((uintptr_t)flow & 7) || ((uintptr_t)mask & 7) || ((uintptr_t)out & 15)) return RAFTK_ERR_ALIGN;
  const long long npix = (long long)B * h * w;
  if (npix == 0) return RAFTK_OK;
  const long long blocks = (npix + kUpsWarps - 1) / kUpsWarps;
  if (blocks > 0x7fffffffLL) return RAFTK_ERR_SHAPE;
  convex_upsample_kernel<<<(unsigned)blocks, kUpsWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      flow, mask, out, B, h, w);
  return launch_status();
}
