// MaskPredictor's last 1x1 convolution (256 -> 576, with the 0.25 factor folded
// into its weights; model.py:394-400) fused with the convex 8x upsampler (model.py:85-98), so the
// 576-channel mask (16.2 MB per 440x1024 pair and iteration) is never written to memory:
//
//   mask[p, k*64 + s] = sum_c W[k*64 + s, c] * m[p, c] + bias[k*64 + s]        (TF32 product)
//   out[b, 8y+fy, 8x+fx, :] = sum_k softmax_k(mask[p, k, s]) * 8*flow[b, y+dy, x+dx, :],  s = fy*8 + fx
//
// One CTA = 128 coarse pixels (UMMA M = 128, cta_group::1; accumulator row = TMEM lane = pixel).  The 576
// output channels are produced in four passes of 144 columns: pass g holds, for the 16 sub-pixels
// s in [16g, 16g+16), all nine neighbour planes k (column k*16 + s - 16g), so the softmax over k of a
// sub-pixel is complete inside one pass.  The host permutes the weight rows and the bias into that order.
//   warp 8   TMA producer: the A tile (128 pixels x K floats, K <= 256, 8 K-blocks, resident) once — one barrier
//            per K-block, so the first pass starts on K-block 0 while the rest of the tile is still arriving —
//            then the weights, one (pass, K-block) slab of 144 rows x 128 B per stage of a 4-stage ring;
//   warp 9   one thread issues tcgen05.mma.kind::tf32 (K = 8 each) into one of two TMEM accumulators
//            (columns 0 and 256), so the MMAs of pass g+1 run under the epilogue of pass g;
//   warps 0-7 epilogue, thread = (pixel, 8 of the pass's 16 sub-pixels; TMEM lane quarter = warp % 4): nine
//            tcgen05.ld.32x32b.x8 bring all planes of its sub-pixels into registers, the accumulator is
//            released at once, then max / exp / sum / convex combination (the arithmetic of
//            convex_upsample.cu) and one 64-byte output row segment per pass.
// The CTAs are persistent (one per SM at most, tiles blockIdx.x, + gridDim.x, ...): weight ring, accumulator pair and
// barrier phases carry over the tile boundary, and the NEXT tile's A blocks are loaded K-block by K-block as the last
// pass of the current tile releases them (a_empty barriers), so that the 128 KB A tile — which comes from HBM at large
// batch — arrives under MMAs instead of in front of them; TMEM allocation and the bias staging happen once per CTA.
// Small batches: the passes are independent (different output rows), so when the pixel tiles alone would leave
// more than half of the SMs idle (batch 1 at 440x1024: 55 tiles on 148 SMs) the grid gets a second dimension and
// each CTA runs two of the four passes of its tile (110 CTAs; the A tile is read twice, from L2).
// Status: DESIGN.md §4.7.

#include <cuda.h>

#include <cstring>

#include "raftk_common.cuh"
#include "tc_ptx.cuh"

namespace raftk {
namespace mu {

using namespace raftk::ptx;

constexpr int TM = 128;             // pixels per CTA
constexpr int NP = 144;             // columns per pass: 9 planes x 16 sub-pixels
constexpr int PASSES = 4;
constexpr int STAGES = 4;
constexpr int A_BLOCK_BYTES = TM * 128;   // 16 KB: 128 rows x 32 floats
constexpr int B_STAGE_BYTES = NP * 128;   // 18 KB: 144 rows x 32 floats
constexpr int EPI_WARPS = 8;         // two per TMEM lane quarter: sub-pixels 0-7 and 8-15 of a pass
constexpr int NUM_THREADS = (EPI_WARPS + 2) * 32;
constexpr int SMEM_LIMIT = 232448;

struct Params {
  const float* bias;  // [576], pass order
  const float* flow;  // (B, h, w, 2)
  float* out;         // (B, 8h, 8w, 2)
  int B, h, w, K;
  long long Q;
};

__global__ void __launch_bounds__(NUM_THREADS, 1)
mask_upsample_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int KB = p.K / 32;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + KB * A_BLOCK_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + STAGES * B_STAGE_BYTES);
  uint64_t* a_full = bars;                // [8]: one per K-block
  uint64_t* a_empty = bars + 8;           // [8]: the last pass of a tile is done with this K-block of A
  uint64_t* b_full = bars + 16;           // [STAGES]
  uint64_t* b_empty = b_full + STAGES;    // [STAGES]
  uint64_t* acc_full = b_empty + STAGES;  // [2]
  uint64_t* acc_empty = acc_full + 2;     // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  float* smem_bias = reinterpret_cast<float*>(bars + 32);  // [576]
  // this CTA's share of the four passes (gridDim.y = 1: all; 2: two each)
  const int npass = PASSES / (int)gridDim.y, pass0 = (int)blockIdx.y * npass;
  // persistent over the pixel tiles blockIdx.x, blockIdx.x + gridDim.x, ...: the roles below run one continuous
  // pipeline across tile boundaries (weight ring, accumulator pair and barrier phases carry over), and the A tile
  // of the NEXT tile arrives K-block by K-block under the last pass of the current one
  const long long ntiles = (p.Q + TM - 1) / TM;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == EPI_WARPS && lane == 0) {
    for (int kb = 0; kb < 8; ++kb) {
      mbar_init(&a_full[kb], 1);
      mbar_init(&a_empty[kb], 1);
    }
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], EPI_WARPS);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == EPI_WARPS + 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == EPI_WARPS) {
    // ===================== TMA producer: the whole warp walks the loops, one elected lane issues =====================
    // (uniform control flow keeps tensor-map pointers, barriers and coordinates in uniform registers: tc_ptx.cuh)
    int stage = 0;
    uint32_t phase = 0, tile_par = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, tile_par ^= 1) {
      const int q0 = (int)(tile * TM);
      const bool first_tile = tile == (long long)blockIdx.x;
      for (int gi = 0; gi < npass; ++gi)
        for (int kb = 0; kb < KB; ++kb) {
          if (gi == 0 && kb < STAGES) {
            // this tile's A blocks: block kb just ahead of the first weight slab that needs it and, once the ring is
            // primed (kb = STAGES - 1), all the remaining ones, so that they do not queue behind ring slots.  A
            // block's buffer is free once the previous tile's last pass has consumed it.
            const int k_end = (kb == STAGES - 1) ? KB : kb + 1;
            for (int k = kb; k < k_end; ++k) {
              if (!first_tile) mbar_wait(&a_empty[k], tile_par ^ 1);
              if (elect_one()) {
                mbar_expect_tx(&a_full[k], (uint32_t)A_BLOCK_BYTES);  // rows past Q are zero-filled and still counted
                tma_load_2d(&tmA, &a_full[k], smem_a + k * A_BLOCK_BYTES, k * 32, q0);
              }
              __syncwarp();
            }
          }
          mbar_wait(&b_empty[stage], phase ^ 1);
          if (elect_one()) {
            mbar_expect_tx(&b_full[stage], B_STAGE_BYTES);
            tma_load_2d(&tmW, &b_full[stage], smem_b + stage * B_STAGE_BYTES, kb * 32, (pass0 + gi) * NP);
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
    }
  } else if (warp == EPI_WARPS + 1) {
    // ===================== MMA issuer: the whole warp walks the loops, one elected lane issues =====================
    // kind::tf32, D = f32, A = B = tf32, K-major, N = 144, M = 128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(NP >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
    // descriptors differ from their base only in the address field (bytes >> 4; no carry out of it below 256 KB)
    const uint64_t a_desc0 = umma_desc(smem_u32(smem_a)), b_desc0 = umma_desc(smem_u32(smem_b));
    int stage = 0;
    uint32_t phase = 0, tile_par = 0;
    int gp = 0;  // passes issued so far (accumulator gp & 1)
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, tile_par ^= 1)
      for (int gi = 0; gi < npass; ++gi, ++gp) {
        const int buf = gp & 1;
        mbar_wait(&acc_empty[buf], ((gp >> 1) & 1) ^ 1);  // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)buf * 256;
        for (int kb = 0; kb < KB; ++kb) {
          if (gi == 0) mbar_wait(&a_full[kb], tile_par);
          mbar_wait(&b_full[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            tc_mma_tf32_x4(tmem_d, a_desc0 + (uint64_t)(kb * (A_BLOCK_BYTES >> 4)), b_desc0 + (uint64_t)(stage * (B_STAGE_BYTES >> 4)),
                           idesc, kb != 0);
            tc_commit(&b_empty[stage]);
            if (gi == npass - 1 && tile + gridDim.x < ntiles) tc_commit(&a_empty[kb]);  // the next tile's A block may land here
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (elect_one()) tc_commit(&acc_full[buf]);
        __syncwarp();
      }
  } else {
    // ===================== epilogue: thread = (pixel, half of the pass's 16 sub-pixels) =====================
    const int quarter = warp & 3, half = warp >> 2;
    const long long W8 = 8LL * p.w;
    int gp = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const long long pix = tile * TM + 32 * quarter + lane;
      const bool valid = pix < p.Q;
      const int x = valid ? (int)(pix % p.w) : 0;
      const int y = valid ? (int)((pix / p.w) % p.h) : 0;
      const int b = valid ? (int)(pix / ((long long)p.w * p.h)) : 0;
      // 8 * flow of the nine neighbours, zero outside the image (SAME padding, model.py:91-92)
      float fxk[9], fyk[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
        float2 f = make_float2(0.0f, 0.0f);
        if (valid && (unsigned)yy < (unsigned)p.h && (unsigned)xx < (unsigned)p.w)
          f = __ldg(reinterpret_cast<const float2*>(p.flow) + ((long long)b * p.h + yy) * p.w + xx);
        fxk[k] = 8.0f * f.x;
        fyk[k] = 8.0f * f.y;
      }
      if (tile == (long long)blockIdx.x) {
        // the bias of this CTA's passes -> shared memory, once, by the epilogue warps alone and after the common
        // barrier: the producer and the MMA issuer do not wait for these global loads, which complete (like the flow
        // loads above) under the first pass's MMAs
        for (int i = threadIdx.x; i < npass * NP; i += EPI_WARPS * 32) smem_bias[pass0 * NP + i] = __ldg(p.bias + pass0 * NP + i);
        asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");
      }
      for (int gi = 0; gi < npass; ++gi, ++gp) {
        const int g = pass0 + gi, buf = gp & 1;
        mbar_wait(&acc_full[buf], (gp >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(32 * quarter) << 16) + (uint32_t)buf * 256 + 8 * half;
        const float* sb = smem_bias + g * NP + 8 * half;
        // all nine planes of this thread's eight sub-pixels: 72 registers, one wait
        float v[9][8];
#pragma unroll
        for (int k = 0; k < 9; ++k) tmem_ld_32x8(taddr + k * 16, v[k]);
        tmem_ld_wait();
        // the accumulator has been read: hand it back to the MMA warp
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_empty[buf]);
        float o[16];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float mx = -INFINITY;
#pragma unroll
          for (int k = 0; k < 9; ++k) {
            v[k][i] += sb[k * 16 + i];
            mx = fmaxf(mx, v[k][i]);
          }
          float s = 0.0f, ax = 0.0f, ay = 0.0f;
#pragma unroll
          for (int k = 0; k < 9; ++k) {
            const float e = __expf(v[k][i] - mx);
            s += e;
            ax = fmaf(e, fxk[k], ax);
            ay = fmaf(e, fyk[k], ay);
          }
          const float inv = 1.0f / s;
          o[2 * i] = ax * inv;
          o[2 * i + 1] = ay * inv;
        }
        if (valid) {
          // sub-pixels 16g + 8*half + i -> output row fy = 2g + half, fx = i: 8 pixels x 2 channels = 64 bytes
          float* orow = p.out + (((long long)b * 8 * p.h + 8 * y + 2 * g + half) * W8 + 8 * x) * 2;
#pragma unroll
          for (int j = 0; j < 16; j += 4) st_na_f4(orow + j, make_float4(o[j], o[j + 1], o[j + 2], o[j + 3]));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == EPI_WARPS + 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

}  // namespace mu
}  // namespace raftk

extern "C" int raftk_mask_upsample(const float* m, const float* weight, const float* bias, const float* flow, float* out,
                                   int B, int h, int w, int K, void* stream) {
  using namespace raftk;
  using namespace raftk::mu;
  if (B < 0 || h <= 0 || w <= 0 || K <= 0) return RAFTK_ERR_SHAPE;
  if (K % 32 || K > 256) return RAFTK_ERR_UNSUPPORTED;
  if (B == 0) return RAFTK_OK;
  if (!m || !weight || !bias || !flow || !out) return RAFTK_ERR_NULL;
  if (((uintptr_t)m & 15) || ((uintptr_t)weight & 15) || ((uintptr_t)bias & 3) || ((uintptr_t)flow & 7) || ((uintptr_t)out & 15))
    return RAFTK_ERR_ALIGN;
  const long long Q = (long long)B * h * w;
  if (Q > 0x7fffffffLL - TM) return RAFTK_ERR_SHAPE;
  EncodeTiledFn encode = get_encode_fn();
  if (!encode) return RAFTK_ERR_UNSUPPORTED;
  CUtensorMap tmA, tmW;
  {
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)Q};
    const cuuint64_t strides[1] = {(cuuint64_t)K * 4};
    const cuuint32_t box[2] = {32, TM};
    const cuuint32_t estr[2] = {1, 1};
    if (encode(&tmA, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(m), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return RAFTK_ERR_UNSUPPORTED;
  }
  {
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)(PASSES * NP)};
    const cuuint64_t strides[1] = {(cuuint64_t)K * 4};
    const cuuint32_t box[2] = {32, NP};
    const cuuint32_t estr[2] = {1, 1};
    if (encode(&tmW, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(weight), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return RAFTK_ERR_UNSUPPORTED;
  }
  Params p{};
  p.bias = bias;
  p.flow = flow;
  p.out = out;
  p.B = B;
  p.h = h;
  p.w = w;
  p.K = K;
  p.Q = Q;
  const size_t smem = 1024 + (size_t)(K / 32) * A_BLOCK_BYTES + (size_t)STAGES * B_STAGE_BYTES + 256 + PASSES * NP * 4;
  if (smem > (size_t)SMEM_LIMIT) return RAFTK_ERR_UNSUPPORTED;
  cudaError_t e = cudaFuncSetAttribute(mask_upsample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  const long long blocks = (Q + TM - 1) / TM;
  // pixel tiles alone would leave more than half of the SMs idle: two CTAs per tile, two passes each
  const unsigned split = 2 * blocks <= (long long)sm_count() ? 2u : 1u;
  // persistent: one CTA per SM at most, each walking the tiles blockIdx.x, blockIdx.x + gridDim.x, ...
  const long long grid_x = blocks < (long long)sm_count() ? blocks : (long long)sm_count();
  mask_upsample_kernel<<<dim3((unsigned)grid_x, split), NUM_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(tmA, tmW, p);
  return launch_status();
}
