// the window lookup (CorrBlock.index_pyramid, model.py:448-470) fused with
// MotionEncoder.convcorr1 (1x1 convolution 324 -> 256 + ReLU, model.py:275-276), so the 324-channel
// correlation features never leave the SM.
//
//   out[q, n] = act( sum_c W[n, c] * lookup(q)[c] + bias[n] )
//
// One CTA = 64 query pixels (UMMA M = 64, cta_group::1), all N <= 256 output channels:
//   warps 0-15 lookup, four queries per warp, LEVEL-MAJOR: the float32 tap arithmetic of corr_lookup.cu
//              (bit-identical values), one level of all four queries at a time, with the next level's
//              footprint chunks in flight; the taps go into a K-major, 128-byte-swizzled A tile in shared
//              memory (12 K-blocks of 64 rows x 32 floats for r = 4; each level's 81 channels are padded to
//              96, so level l is exactly K-blocks 3l..3l+2 and its MMAs can start when its taps are done);
//   warp 16    TMA producer: the prepared weights W' (N rows x K' floats) stream through a 3-stage ring,
//              one 32-float K-block (N x 128 B) per stage;
//   warp 17    one thread issues tcgen05.mma.kind::tf32 (K = 8 per instruction), level by level, into a
//              64 x N float32 accumulator in TMEM — the MMAs of level l run under the taps of level l + 1;
//   warps 0-15 epilogue: tcgen05.ld (lane quarter = warp % 4, N columns split over the four warps of a
//              quarter), + bias (staged in shared memory), ReLU, 128-byte row segments to global memory.
// The multiplication is TF32 (10-bit mantissa operands, float32 accumulation) — what the library
// convolution it replaces uses in the default precision mode.
//
// Status: parity-green, but still slower than the two-kernel path (batch 1: 21.4 us against
// 15.4 us for raftk_corr_lookup + the library's 1x1 convolution back to back), so the host model does not use
// it by default.  Phase elimination at batch 1: launch + TMEM + epilogue skeleton 8.2 us, + taps 10.8 us,
// + weights and MMAs 3.2 us (mostly hidden).  The taps are the problem: 64 queries per SM on 16 warps issue at
// 40 % (ncu: 1073 instructions per query against 779 in the stand-alone kernel, 89 KB of unrolled code with
// instruction-fetch stalls, four warps per scheduler), where the stand-alone kernel spreads the same
// arithmetic over 148 SMs x 48 warps.

#include <cuda.h>

#include <cstring>

#include "raftk_common.cuh"
#include "lookup_taps.cuh"
#include "tc_ptx.cuh"

namespace raftk {
namespace lc {

constexpr int TM = 64;  // queries per CTA
constexpr int LOOKUP_WARPS = 16;
constexpr int NUM_THREADS = (LOOKUP_WARPS + 2) * 32;
constexpr int STAGES = 3;
constexpr int PITCH = 20;
constexpr int A_BLOCK_BYTES = TM * 128;  // one K-block of the A tile: 64 rows x 32 floats
constexpr int SMEM_LIMIT = 232448;       // 227 KB

struct Params {
  const float* level[4];
  int lh[4], lw[4], aligned[4];
  const float* coords;
  const float* bias;
  float* out;
  int Q, N, relu;
};

using namespace raftk::ptx;

// One footprint chunk (see corr_lookup.cu): floats [gx, gx+4) of row gy, zeros outside the level.
__device__ __forceinline__ float4 load_chunk(const float* slab, int lh, int lw, int gy, int gx, bool aligned, uint64_t pol) {
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

template <int R>
__global__ void __launch_bounds__(NUM_THREADS, 1)
lookup_conv_kernel(const __grid_constant__ CUtensorMap tmW, const Params p) {
  constexpr int S = 2 * R + 1, SS = S * S, FP = 2 * R + 2, ROWS = FP + 1, TILE = ROWS * PITCH;
  constexpr int NTAP = (SS + 31) / 32;
  constexpr int KL = NTAP * 32;     // channels per level in the A tile: 96 (r = 4), 64 (r = 3)
  constexpr int KB = 4 * KL / 32;   // K-blocks: 12 / 8
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int b_stage_bytes = p.N * 128;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + KB * A_BLOCK_BYTES;
  float* smem_f = reinterpret_cast<float*>(smem_b + STAGES * b_stage_bytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_f + LOOKUP_WARPS * TILE);
  uint64_t* b_full = bars;             // [STAGES]
  uint64_t* b_empty = bars + STAGES;   // [STAGES]
  uint64_t* a_ready = bars + 2 * STAGES;  // [4]: level l's K-blocks of the A tile are written
  uint64_t* acc_full = a_ready + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);
  float* smem_bias = reinterpret_cast<float*>(bars + 16);  // [N], 128 bytes past the barriers

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * TM;
  const uint32_t tmem_cols = p.N <= 32 ? 32u : p.N <= 64 ? 64u : p.N <= 128 ? 128u : 256u;

  if (warp == LOOKUP_WARPS && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], 1);
    }
    for (int l = 0; l < 4; ++l) mbar_init(&a_ready[l], LOOKUP_WARPS);
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == LOOKUP_WARPS + 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  for (int i = threadIdx.x; i < p.N; i += NUM_THREADS) smem_bias[i] = __ldg(p.bias + i);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < LOOKUP_WARPS) {
    // ===================== lookup: taps -> A tile, level-major =====================
    // Each warp owns QPW rows of the tile (queries q0 + warp + LOOKUP_WARPS * j) and walks them level by
    // level, so that level l's K-blocks are complete — and their MMAs can run — while the taps of level
    // l + 1 are evaluated.  One (query, level) footprint is in flight ahead of the one being evaluated.
    constexpr int QPW = TM / LOOKUP_WARPS;
    float* F = smem_f + warp * TILE;  // one level's footprint tile
    for (int e = lane; e < PITCH; e += 32) F[FP * PITCH + e] = 0.0f;  // guard row
    // padding columns of the A tile (channels SS..KL-1 of every level) are zero
    for (int e = lane; e < QPW * 4 * (KL - SS); e += 32) {
      const int rr = e / (4 * (KL - SS)), rem = e % (4 * (KL - SS));
      const int l = rem / (KL - SS), c = l * KL + SS + rem % (KL - SS);
      const int r = warp + LOOKUP_WARPS * rr;
      *reinterpret_cast<float*>(smem_a + (c >> 5) * A_BLOCK_BYTES + r * 128 + ((((c & 31) >> 2) ^ (r & 7)) << 4) + (c & 3) * 4) = 0.0f;
    }
    __syncwarp();

    float di[NTAP], dj[NTAP];
#pragma unroll
    for (int k = 0; k < NTAP; ++k) {
      const int t = lane + 32 * k < SS ? lane + 32 * k : 0;
      const int i = t / S, j = t - i * S;
      di[k] = (float)(i - R);
      dj[k] = (float)(j - R);
    }
    const bool last_tap_valid = lane + 32 * (NTAP - 1) < SS;
    const int row_u = lane >> 2, chunk = lane & 3;
    const uint64_t pol = l2_policy_evict_last();

    float cx[QPW], cy[QPW];
    bool fastq[QPW], validq[QPW];
#pragma unroll
    for (int j = 0; j < QPW; ++j) {
      const int q = q0 + warp + LOOKUP_WARPS * j;
      validq[j] = q < p.Q;
      cx[j] = validq[j] ? __ldg(p.coords + 2 * (size_t)q) : 0.0f;
      cy[j] = validq[j] ? __ldg(p.coords + 2 * (size_t)q + 1) : 0.0f;
      fastq[j] = validq[j] && fabsf(cx[j]) <= 1.0e6f && fabsf(cy[j]) <= 1.0e6f;
    }

    // the chunks of one level's footprints of all QPW queries (rows 0-7 by all lanes, rows 8.. by lanes 0-7),
    // fetched one level ahead of the taps: DRAM latency hides behind a whole level of tap arithmetic
    float4 cur0[QPW], cur1[QPW], nxt0[QPW], nxt1[QPW];
    auto fetch_level = [&](int l, float4* a0, float4* a1) {
      const float sc = __int_as_float((127 - l) << 23);
      const int lh = p.lh[l], lw = p.lw[l];
      const bool al = p.aligned[l] != 0;
#pragma unroll
      for (int j = 0; j < QPW; ++j) {
        a0[j] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        a1[j] = a0[j];
        if (!fastq[j]) continue;
        const int q = q0 + warp + LOOKUP_WARPS * j;
        const int bx = (int)floorf(cx[j] * sc) - R, by = (int)floorf(cy[j] * sc) - R;
        const float* slab = p.level[l] + (size_t)q * lh * lw;
        const int gx = (bx & ~3) + 4 * chunk;
        if (gx <= bx + FP - 1) {
          a0[j] = load_chunk(slab, lh, lw, by + row_u, gx, al, pol);
          if (FP > 8 && row_u < FP - 8) a1[j] = load_chunk(slab, lh, lw, by + 8 + row_u, gx, al, pol);
        }
      }
    };

    fetch_level(0, cur0, cur1);
#pragma unroll
    for (int l = 0; l < 4; ++l) {
      const float sc = __int_as_float((127 - l) << 23);
      if (l + 1 < 4) fetch_level(l + 1, nxt0, nxt1);
#pragma unroll
      for (int j = 0; j < QPW; ++j) {
        const int r = warp + LOOKUP_WARPS * j;
        if (fastq[j]) {
          *reinterpret_cast<float4*>(F + row_u * PITCH + chunk * 4) = cur0[j];
          if (FP > 8 && row_u < FP - 8) *reinterpret_cast<float4*>(F + (8 + row_u) * PITCH + chunk * 4) = cur1[j];
        }
        __syncwarp();
        uint8_t* arow = smem_a + r * 128;
        const int sw = r & 7;
        const float cxl = cx[j] * sc, cyl = cy[j] * sc;
        const float fbxa = (float)(((int)floorf(cxl) - R) & ~3);
        const float fby = floorf(cyl) - (float)R;
#pragma unroll
        for (int k = 0; k < NTAP; ++k) {
          float acc;
          if (fastq[j]) {
            const float sx = __fadd_rn(cxl, di[k]), sy = __fadd_rn(cyl, dj[k]);
            const float fx0 = floorf(sx), fy0 = floorf(sy);
            const float wx1 = __fadd_rn(sx, -fx0), wy1 = __fadd_rn(sy, -fy0);
            const float wx0 = __fadd_rn(1.0f, -wx1), wy0 = __fadd_rn(1.0f, -wy1);
            const int off = __float2int_rn(fmaf(fy0 - fby, (float)PITCH, fx0 - fbxa));
            const float* f = F + off;
            acc = __fmul_rn(__fmul_rn(wy0, wx0), f[0]);
            acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy0, wx1), f[1]));
            acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx0), f[PITCH]));
            acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx1), f[PITCH + 1]));
          } else {
            // row past the end of the batch: zeros.  Query farther than 1e6 from the map: every corner
            // of every tap is outside every level -> 0, or NaN when a coordinate is not finite
            // (0 * weight, as the reference's arithmetic gives)
            acc = validq[j] ? __fadd_rn(__fadd_rn(cx[j], -cx[j]), __fadd_rn(cy[j], -cy[j])) : 0.0f;
          }
          if (k < NTAP - 1 || last_tap_valid)  // channel l*KL + 32*k + lane of the padded K order: K-block l*KL/32 + k
            *reinterpret_cast<float*>(arow + (l * (KL / 32) + k) * A_BLOCK_BYTES + (((lane >> 2) ^ sw) << 4) + (lane & 3) * 4) = acc;
        }
        __syncwarp();
      }
      fence_async_smem();  // generic-proxy writes of level l's K-blocks -> visible to the tensor core's async proxy
      __syncwarp();
      if (lane == 0) mbar_arrive(&a_ready[l]);
#pragma unroll
      for (int j = 0; j < QPW; ++j) {
        cur0[j] = nxt0[j];
        cur1[j] = nxt1[j];
      }
    }

    // ===================== epilogue (all lookup warps) =====================
    // TMEM lane quarter = warp % 4 (M = 64: accumulator row m lives in lane (m % 16) + 32 * (m / 16), lanes
    // 0-15 of each quarter); the N columns are split over the LOOKUP_WARPS / 4 warps that share a quarter
    {
      mbar_wait(acc_full, 0);
      tc_fence_after();
      const int quarter = warp & 3, part = warp >> 2;
      constexpr int PARTS = LOOKUP_WARPS / 4;
      const int m = 16 * quarter + lane;
      const long long q = (long long)q0 + m;
      const bool store = lane < 16 && q < p.Q;
      for (int n0 = 32 * part; n0 < p.N; n0 += 32 * PARTS) {
        float acc[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(32 * quarter) << 16) + (uint32_t)n0, acc);
        tmem_ld_wait();
        if (store) {
          float* orow = p.out + q * p.N + n0;
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            if (n0 + c < p.N) {
              const float4 b = *reinterpret_cast<const float4*>(smem_bias + n0 + c);
              float4 o = make_float4(acc[c] + b.x, acc[c + 1] + b.y, acc[c + 2] + b.z, acc[c + 3] + b.w);
              if (p.relu) {
                o.x = fmaxf(o.x, 0.0f);
                o.y = fmaxf(o.y, 0.0f);
                o.z = fmaxf(o.z, 0.0f);
                o.w = fmaxf(o.w, 0.0f);
              }
              *reinterpret_cast<float4*>(orow + c) = o;
            }
          }
        }
      }
    }
  } else if (warp == LOOKUP_WARPS) {
    // ===================== TMA producer: weight K-blocks =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < KB; ++kb) {
        mbar_wait(&b_empty[stage], phase ^ 1);
        mbar_expect_tx(&b_full[stage], (uint32_t)b_stage_bytes);
        tma_load_2d(&tmW, &b_full[stage], smem_b + stage * b_stage_bytes, kb * 32, 0);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc = make_idesc(TM, p.N);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < KB; ++kb) {
        if (kb % (KL / 32) == 0) {  // first K-block of a level: its taps must be in the A tile
          mbar_wait(&a_ready[kb / (KL / 32)], 0);
          tc_fence_after();
        }
        mbar_wait(&b_full[stage], phase);
        tc_fence_after();
        const uint32_t as = smem_u32(smem_a + kb * A_BLOCK_BYTES);
        const uint32_t bs = smem_u32(smem_b + stage * b_stage_bytes);
#pragma unroll
        for (int k8 = 0; k8 < 4; ++k8)  // 8 floats = 32 bytes of K per instruction
          tc_mma_tf32(tmem_base, umma_desc(as + 32 * k8), umma_desc(bs + 32 * k8), idesc, (kb | k8) != 0);
        tc_commit(&b_empty[stage]);  // the stage is free once these MMAs have read it
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      tc_commit(acc_full);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == LOOKUP_WARPS + 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols));
  }
}

static int padded_level_channels(int radius) {
  const int ss = (2 * radius + 1) * (2 * radius + 1);
  return (ss + 31) / 32 * 32;
}

}  // namespace lc
}  // namespace raftk

extern "C" int raftk_lookup_conv_weight_cols(int num_levels, int radius) {
  if (num_levels != 4 || (radius != 3 && radius != 4)) return 0;
  return 4 * raftk::lc::padded_level_channels(radius);
}

extern "C" int raftk_corr_lookup_conv1x1(const float* const* levels, int num_levels, const float* coords,
                                         const float* weight, const float* bias, float* out, int Q, int h, int w,
                                         int radius, int N, int relu, void* stream) {
  using namespace raftk;
  using namespace raftk::lc;
  if (Q < 0 || h <= 0 || w <= 0 || num_levels < 1 || num_levels > RAFTK_MAX_LEVELS) return RAFTK_ERR_SHAPE;
  if (num_levels != 4 || (radius != 3 && radius != 4) || N < 16 || N > 256 || N % 16) return RAFTK_ERR_UNSUPPORTED;
  if ((h >> (num_levels - 1)) < 1 || (w >> (num_levels - 1)) < 1) return RAFTK_ERR_TOO_SMALL;
  if (Q == 0) return RAFTK_OK;
  if (!levels || !coords || !weight || !bias || !out) return RAFTK_ERR_NULL;
  if (((uintptr_t)weight & 15) || ((uintptr_t)bias & 15) || ((uintptr_t)out & 15) || ((uintptr_t)coords & 7)) return RAFTK_ERR_ALIGN;
  Params p{};
  for (int l = 0; l < 4; ++l) {
    if (!levels[l]) return RAFTK_ERR_NULL;
    p.level[l] = levels[l];
    p.lh[l] = h >> l;
    p.lw[l] = w >> l;
    p.aligned[l] = (p.lw[l] % 4 == 0) && (((uintptr_t)levels[l] & 15) == 0);
  }
  p.coords = coords;
  p.bias = bias;
  p.out = out;
  p.Q = Q;
  p.N = N;
  p.relu = relu;
  const int KL = padded_level_channels(radius), K = 4 * KL, KB = K / 32;
  EncodeTiledFn encode = get_encode_fn();
  if (!encode) return RAFTK_ERR_UNSUPPORTED;
  CUtensorMap tmW;
  {
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)N};
    const cuuint64_t strides[1] = {(cuuint64_t)K * 4};
    const cuuint32_t box[2] = {32, (cuuint32_t)N};
    const cuuint32_t estr[2] = {1, 1};
    if (encode(&tmW, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(weight), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return RAFTK_ERR_UNSUPPORTED;
  }
  const int tile = (2 * radius + 3) * PITCH;
  const size_t smem = 1024 + (size_t)KB * A_BLOCK_BYTES + (size_t)STAGES * N * 128 + (size_t)LOOKUP_WARPS * tile * 4 + 128 + (size_t)N * 4;
  if (smem > (size_t)SMEM_LIMIT) return RAFTK_ERR_UNSUPPORTED;
  const int blocks = (Q + TM - 1) / TM;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e;
  if (radius == 4) {
    e = cudaFuncSetAttribute(lookup_conv_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    lookup_conv_kernel<4><<<blocks, NUM_THREADS, smem, s>>>(tmW, p);
  } else {
    e = cudaFuncSetAttribute(lookup_conv_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    lookup_conv_kernel<3><<<blocks, NUM_THREADS, smem, s>>>(tmW, p);
  }
  return launch_status();
}
