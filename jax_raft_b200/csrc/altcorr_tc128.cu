// The on-the-fly ("alt_corr") window lookup on the tensor cores at M = 128, with the A operand in TENSOR MEMORY.
// Same result contract, operands and arithmetic as altcorr_tc.cu (read its header first); what differs:
//
//   item       one 16 x 8 patch of query pixels (two neighbouring 8 x 8 patches), all levels in turn, UMMA M = 128.  The
//              union box of the patch has ~35 % fewer cells per query than two 8 x 8 boxes (fewer B bytes and fewer
//              stages per query: the stage pipeline, not the tensor pipe, paces these kernels — DESIGN.md §4.4), and
//              accumulator row m = TMEM lane m, so all 32 lanes of every tcgen05.ld of the extraction carry a query
//              (16 of 32 at M = 64).
//   A          128 queries x 2D bf16 = 128 KB does not fit shared memory beside the footprints — it lives in TMEM:
//              lane = query, 256 columns of packed bf16 pairs ([hi | mid], K = 2D <= 512), written by four loader warps
//              (TMEM lane quarter = warp & 3) straight from global memory with tcgen05.st.32x32b — a thread owns its
//              query's row — k-block by k-block as the last chunk of the previous item releases it.  The MMAs are
//              tcgen05.mma [d_tmem], [a_tmem], b_desc (TS form; profiles/scratch/ts_mma_probe.cu pins the layout), so
//              the MMAs read only B from shared memory.
//   B          a stage of the ring is ONE TMA box {64 channels, 8 ncb cells, rows} from a tensor map per (level, box
//              width ncb = 1 .. 4); rows = rows_for(ncb) is as tall as N = 8 ncb rows <= 128 allows (8, 8, 5, 4), so the
//              narrow boxes of the coarse levels need half the chunks of a fixed height and every stage is 8-16 KB;
//              accumulator column of box cell (r, x) = 8 ncb r + x.
//   D          two accumulators of 128 columns (columns 0-255; A: 256-511).
//   wide boxes when the union box of the patch is wider than 32 cells or taller than 32 rows, the left and right 8 x 8
//              halves get a box each and are multiplied in turn (a query extracts from its own half's box; the other
//              half's accumulator rows are ignored); only if a half does not fit either does the level take the scalar
//              float32 path — exactly where altcorr_tc.cu does.
//   roles      warps 0-3 extraction, 4-7 taps (32 queries each), 8 B producer, 9 MMA issuer, 10 item descriptions
//              (coordinates + boxes, four in flight; the extraction threads recompute their own window origins),
//              11-14 A loaders; 480 threads, 128 registers.
//
// Measured, bands of a 136 x 240 image (profiles/altcorr_time.py 136 240 rows), against altcorr_tc.cu: 136 rows 96 vs
// 114 us, 68 rows 59 vs 69, 34 rows 41 vs 45, 17 rows 34 vs 31 — raftk_altcorr_lookup_tc picks by problem size.

#include <cuda.h>
#include <cuda_bf16.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "altcorr_tc128.cuh"
#include "lookup_taps.cuh"
#include "raftk_common.cuh"
#include "tc_ptx.cuh"

namespace raftk {
namespace atc128 {

using namespace raftk::ptx;

constexpr int TQ = 128, PW = 16, PH = 8;        // queries per item: 16 x 8
// box rows per chunk = the y extent of the TMA box, by box width in column blocks: as tall as N = 8 * ncb * rows <= 128
// allows (N = 64, 128, 120, 128), so that the narrow boxes of the coarse levels need half as many chunks and stages
__host__ __device__ constexpr int rows_for(int ncb) { return ncb >= 4 ? 4 : ncb == 3 ? 5 : 8; }
constexpr int MAX_BOX_ROWS = 32;                // box height the tensor-core path takes
constexpr int MAX_CB = 4;                       // column blocks of 8 cells: box width <= 32, N = 32 * ncb <= 128
constexpr int RING_BYTES = 80 * 1024;           // B ring: stages of ncb * 4 KB
constexpr int NINFO = 4;                        // item descriptions in flight (the describing warp runs NINFO - 1 items ahead)
constexpr int NF = 2;                           // footprint buffers (levels in flight between extraction and taps)
constexpr int ROWBUF_PITCH = 36;
constexpr int MAX_STAGES = 16;
constexpr int ACC_COLS = 128;                   // per accumulator buffer
constexpr int A_COL = 2 * ACC_COLS;             // first TMEM column of the A operand
constexpr int EPI_WARPS = 8;
constexpr int PRODUCER_WARP = 8, MMA_WARP = 9, AUX_WARP = 10, LOADER_WARP0 = 11;
constexpr int NUM_THREADS = 15 * 32;
constexpr int SMEM_LIMIT = 232448;

#ifdef RAFTK_A128_BOUNDED_WAIT
// development aid: a wait that traps after ~seconds instead of hanging the GPU (hides ptx::mbar_wait in this namespace)
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  for (uint32_t it = 0; it < (1u << 22); ++it) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P1;\n\t"
        "}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
    if (ok) return;
  }
  printf("altcorr_tc128: wait timed out (block %d, warp %d, barrier +%d, parity %u)\n", (int)blockIdx.x, (int)(threadIdx.x >> 5),
         (int)(addr & 1023), parity);
  __trap();
}
#endif

struct Params {
  const float* f1;      // float32 sources, used by the scalar path only
  const float* f2[4];
  const __nv_bfloat16* f1_split;  // (B, nq, 2D): read directly by the A loaders
  int lh[4], lw[4];
  const float* coords;
  float* out;
  int B, nq, D, L;
  int rows_q, w_q;
  int npatch;
  int split_from;
  float inv_sqrt_d;
};

struct BMaps {
  CUtensorMap m[4 * MAX_CB];
};

struct QInfo {
  int bx, by, flags, pad;  // flags: bit 0 valid query, bit 1 the query has a window
};

// The (2r+1)^2 taps of level l for the 32 queries of one warp, as column tasks (altcorr_tc.cu: column_taps_level; lane k
// holds the coordinates of query k).
template <int R>
__device__ __forceinline__ void column_taps_level(const float* Fw, float cx, float cy, bool qvalid, int q, int l, float* out,
                                                  int C, int lane) {
  constexpr int S = 2 * R + 1, FP = 2 * R + 2, NCELL = FP * FP;
  const float sc = __int_as_float((127 - l) << 23);  // 2^-l, exact
#pragma unroll 1
  for (int base = 0; base < 32 * S; base += 32) {
    const int T = base + lane;
    const int k = T / S, i = T - k * S;
    const float ccx = __shfl_sync(0xffffffffu, cx, k), ccy = __shfl_sync(0xffffffffu, cy, k);
    const int ok = __shfl_sync(0xffffffffu, qvalid ? 1 : 0, k);
    const int qq = __shfl_sync(0xffffffffu, q, k);
    if (!ok) continue;
    const float cxl = ccx * sc, cyl = ccy * sc;
    const float fcx = floorf(cxl), fcy = floorf(cyl);
    const float dx = (float)(i - R);
    const float wx1 = __fadd_rn(__fadd_rn(cxl, dx), -__fadd_rn(fcx, dx));
    const float wx0 = __fadd_rn(1.0f, -wx1);
    const float* Fc = Fw + k * NCELL + i;
    float* orow = out + (long long)qq * C + l * S * S + i * S;
    float fa = Fc[0], fb = Fc[1];
#pragma unroll
    for (int j = 0; j < S; ++j) {
      const float dy = (float)(j - R);
      const float wy1 = __fadd_rn(__fadd_rn(cyl, dy), -__fadd_rn(fcy, dy));
      const float wy0 = __fadd_rn(1.0f, -wy1);
      const float na = Fc[(j + 1) * FP], nb = Fc[(j + 1) * FP + 1];
      float acc = __fmul_rn(__fmul_rn(wy0, wx0), fa);
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy0, wx1), fb));
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx0), na));
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(wy1, wx1), nb));
      orow[j] = acc;
      fa = na;
      fb = nb;
    }
  }
}

struct LevelBox {
  int X0, Y0, BW, BH;
  int tc;
  int ncb, nchunks, pad;
};

struct ItemInfo {
  // per level: nbox = 1: lbox[l][0] is the bounding box of all windows; nbox = 2: that box is too wide or tall and the
  // left (x < 8) and right halves of the patch have a box each, lbox[l][half] (the half's 64 queries extract, the other
  // half's accumulator rows are ignored); nbox = 0: not even those fit — scalar path
  LevelBox lbox[4][2];
  int nbox[4];
  float2 coord[TQ];
  int qidx[TQ];  // B * nq < 2^31 (checked on the host); -1: no such query (outside the image)
};

struct Item {
  int txi, tyi, b, l_begin, l_end;
};

__device__ __forceinline__ Item decode_item(const Params& p, long long i, int tiles_x, int tiles_y) {
  Item it;
  long long rem;
  if (i < p.split_from) {
    rem = i;
    it.l_begin = 0;
    it.l_end = p.L;
  } else {
    const long long k = i - p.split_from, ntail = p.npatch - p.split_from;
    it.l_begin = (int)(k / ntail);
    it.l_end = it.l_begin + 1;
    rem = p.split_from + k % ntail;
  }
  it.txi = (int)(rem % tiles_x);
  rem /= tiles_x;
  it.tyi = (int)(rem % tiles_y);
  it.b = (int)(rem / tiles_y);
  return it;
}

__device__ __forceinline__ int item_stage_bytes(const ItemInfo& nf, const Item& it) {
  int sb = 4 * 1024;
  for (int l = it.l_begin; l < it.l_end; ++l)
    for (int bi = 0; bi < nf.nbox[l]; ++bi) sb = max(sb, nf.lbox[l][bi].ncb * rows_for(nf.lbox[l][bi].ncb) * 1024);
  return sb;
}

// Four kind::f16 MMAs over one 64-element K block with A in TMEM: +8 columns of A and +32 B of B per K = 16 step.
__device__ __forceinline__ void tc_mma_f16_ts_x4(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t acc_first) {
  asm volatile(
      "{\n\t"
      ".reg .pred p, q;\n\t"
      ".reg .b32 a1, a2, a3;\n\t"
      ".reg .b64 b1, b2, b3;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.eq.b32 q, 0, 0;\n\t"
      "add.u32 a1, %1, 8;\n\t"
      "add.s64 b1, %2, 2;\n\t"
      "add.u32 a2, %1, 16;\n\t"
      "add.s64 b2, %2, 4;\n\t"
      "add.u32 a3, %1, 24;\n\t"
      "add.s64 b3, %2, 6;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [a1], b1, %3, q;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [a2], b2, %3, q;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [a3], b3, %3, q;\n\t"
      "}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(acc_first)
      : "memory");
}

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]),
      "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]),
      "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int R>
__global__ void __launch_bounds__(NUM_THREADS, 1)
altcorr_tc128_kernel(const __grid_constant__ BMaps tmB, const Params p) {
  constexpr int FP = 2 * R + 2, S = 2 * R + 1, NCELL = FP * FP;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int KB = p.D / 64;  // k-blocks per plane (hi, mid)
  uint8_t* smem_b = smem;                                    // ring of RING_BYTES
  float* F = reinterpret_cast<float*>(smem_b + RING_BYTES);  // [2][TQ][NCELL]: footprints of two consecutive levels
  float* rowbuf = F + NF * TQ * NCELL;                        // [4 warps][32 lanes][2 rows][ROWBUF_PITCH]
  uint64_t* bars = reinterpret_cast<uint64_t*>(rowbuf + 4 * 32 * 2 * ROWBUF_PITCH);
  uint64_t* a_full = bars;                      // [4]: one per k-block (hi and mid columns), 4 loader warps arrive
  uint64_t* a_empty = bars + 4;                 // [4]: the item's last MMAs are done with this k-block of A
  uint64_t* b_full = bars + 8;                  // [MAX_STAGES]
  uint64_t* b_empty = b_full + MAX_STAGES;      // [MAX_STAGES]
  uint64_t* acc_full = b_empty + MAX_STAGES;    // [2]
  uint64_t* acc_empty = acc_full + 2;           // [2]
  uint64_t* info_full = acc_empty + 2;          // [NINFO]
  uint64_t* info_empty = info_full + NINFO;     // [NINFO]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(info_empty + NINFO);
  ItemInfo* info = reinterpret_cast<ItemInfo*>(bars + 56);  // [NINFO]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_x = (p.w_q + PW - 1) / PW, tiles_y = (p.rows_q + PH - 1) / PH;
  const long long nitems = p.split_from + (long long)(p.npatch - p.split_from) * p.L;
  int my_levels = 0;
  for (long long i = blockIdx.x; i < nitems; i += gridDim.x) my_levels += i < p.split_from ? p.L : 1;

  if (warp == MMA_WARP && lane == 0) {
    for (int kb = 0; kb < 4; ++kb) {
      mbar_init(&a_full[kb], 4);  // one arrival per loader warp
      mbar_init(&a_empty[kb], 1);
    }
    for (int s = 0; s < MAX_STAGES; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 4);
    }
    for (int i = 0; i < NINFO; ++i) {
      mbar_init(&info_full[i], 1);
      mbar_init(&info_empty[i], 10);  // B producer, MMA issuer, 4 extraction warps, 4 tap warps
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == MMA_WARP) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int quarter = warp & 3;

  if (warp == AUX_WARP) {
    // ===================== item descriptions, one item ahead of everybody else =====================
    int j = 0;
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      ItemInfo& nf = info[j % NINFO];
      if (j >= NINFO) mbar_wait(&info_empty[j % NINFO], (uint32_t)(((j / NINFO) - 1) & 1));
      float cx[4], cy[4];
      bool qv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {  // lane t describes queries t, t + 32, t + 64, t + 96
        const int m = lane + 32 * u;
        const int qy = it.tyi * PH + (m >> 4), qx = it.txi * PW + (m & 15);
        qv[u] = qy < p.rows_q && qx < p.w_q;
        const long long q = (long long)it.b * p.nq + (long long)qy * p.w_q + qx;
        cx[u] = cy[u] = 0.0f;
        if (qv[u]) {
          const float2 c = __ldg(reinterpret_cast<const float2*>(p.coords) + q);
          cx[u] = c.x;
          cy[u] = c.y;
        }
        nf.coord[m] = make_float2(cx[u], cy[u]);
        nf.qidx[m] = qv[u] ? (int)q : -1;
      }
      for (int l = it.l_begin; l < it.l_end; ++l) {
        const int lh = l == 0 ? p.lh[0] : l == 1 ? p.lh[1] : l == 2 ? p.lh[2] : p.lh[3];
        const int lw = l == 0 ? p.lw[0] : l == 1 ? p.lw[1] : l == 2 ? p.lw[2] : p.lw[3];
        const float sc = __int_as_float((127 - l) << 23);
        int mnx = 0x7fffffff, mxx = -0x7fffffff, mny = 0x7fffffff, mxy = -0x7fffffff, any = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const bool normal = fabsf(cx[u]) <= 1.0e6f && fabsf(cy[u]) <= 1.0e6f;
          const int bx = clamp_to_int(floorf(cx[u] * sc)) - R, by = clamp_to_int(floorf(cy[u] * sc)) - R;
          const bool hw = qv[u] && normal && bx < lw && bx + FP > 0 && by < lh && by + FP > 0;
          if (hw) {
            mnx = min(mnx, bx);
            mxx = max(mxx, bx);
            mny = min(mny, by);
            mxy = max(mxy, by);
            any = 1;
          }
        }
        // lanes with bit 3 clear hold queries of the left half (x < 8), the others of the right half: reduce within the
        // halves first (offsets 16, 4, 2, 1), then across them
        auto make_box = [&](int any_, int mnx_, int mxx_, int mny_, int mxy_) {
          LevelBox lb;
          lb.X0 = any_ ? mnx_ : 0;
          lb.Y0 = any_ ? mny_ : 0;
          lb.BW = any_ ? mxx_ - mnx_ + FP : 0;
          lb.BH = any_ ? mxy_ - mny_ + FP : 0;
          lb.tc = (lb.BW <= 8 * MAX_CB && lb.BH <= MAX_BOX_ROWS) ? 1 : 0;
          lb.ncb = (lb.BW + 7) >> 3;
          lb.nchunks = (lb.BH + rows_for(lb.ncb) - 1) / rows_for(lb.ncb);
          lb.pad = 0;
          return lb;
        };
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
          if (off == 8) continue;
          mnx = min(mnx, __shfl_xor_sync(0xffffffffu, mnx, off));
          mxx = max(mxx, __shfl_xor_sync(0xffffffffu, mxx, off));
          mny = min(mny, __shfl_xor_sync(0xffffffffu, mny, off));
          mxy = max(mxy, __shfl_xor_sync(0xffffffffu, mxy, off));
          any |= __shfl_xor_sync(0xffffffffu, any, off);
        }
        const LevelBox mine = make_box(any, mnx, mxx, mny, mxy);  // lane 0: left half, lane 8: right half
        const int o_mnx = __shfl_xor_sync(0xffffffffu, mnx, 8), o_mxx = __shfl_xor_sync(0xffffffffu, mxx, 8);
        const int o_mny = __shfl_xor_sync(0xffffffffu, mny, 8), o_mxy = __shfl_xor_sync(0xffffffffu, mxy, 8);
        const int o_any = __shfl_xor_sync(0xffffffffu, any, 8);
        const LevelBox uni = make_box(any | o_any, min(mnx, o_mnx), max(mxx, o_mxx), min(mny, o_mny), max(mxy, o_mxy));
        const int other_fits = __shfl_xor_sync(0xffffffffu, mine.tc, 8);
        if (lane == 0 || lane == 8) {
          if (uni.tc) {
            if (lane == 0) {
              nf.lbox[l][0] = uni;
              nf.nbox[l] = 1;
            }
          } else if (mine.tc && other_fits) {
            nf.lbox[l][lane >> 3] = mine;
            if (lane == 0) nf.nbox[l] = 2;
          } else if (lane == 0) {
            nf.nbox[l] = 0;
          }
        }
      }
      __threadfence_block();
      __syncwarp();
      if (elect_one()) mbar_arrive(&info_full[j % NINFO]);
      __syncwarp();
    }
  } else if (warp >= LOADER_WARP0) {
    // ===================== A loaders: this thread's query row -> TMEM lane 32 * quarter + lane =====================
    // k-block kb = 64 channels of the hi plane (columns A_COL + 32 kb ..) and of the mid plane (A_COL + 32 (KB + kb) ..),
    // 128 bytes each, as soon as the previous item's last chunk is done with that k-block
    const int m = 32 * quarter + lane;
    const uint32_t tlane = tmem_base + ((uint32_t)(32 * quarter) << 16) + A_COL;
    int j = 0;
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      const int qy = it.tyi * PH + (m >> 4), qx = it.txi * PW + (m & 15);
      const bool valid = qy < p.rows_q && qx < p.w_q;
      const long long q = (long long)it.b * p.nq + (long long)qy * p.w_q + qx;
      const uint4* row = reinterpret_cast<const uint4*>(p.f1_split + (valid ? q : 0) * 2 * p.D);
      for (int kb = 0; kb < KB; ++kb) {
        uint32_t hi[32], mid[32];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const uint4 a = valid ? __ldg(row + kb * 8 + c) : make_uint4(0u, 0u, 0u, 0u);
          const uint4 b = valid ? __ldg(row + (KB + kb) * 8 + c) : make_uint4(0u, 0u, 0u, 0u);
          hi[4 * c] = a.x, hi[4 * c + 1] = a.y, hi[4 * c + 2] = a.z, hi[4 * c + 3] = a.w;
          mid[4 * c] = b.x, mid[4 * c + 1] = b.y, mid[4 * c + 2] = b.z, mid[4 * c + 3] = b.w;
        }
        if (j >= 1) {
          mbar_wait(&a_empty[kb], (uint32_t)((j - 1) & 1));
          tc_fence_after();
        }
        tmem_st_32x32(tlane + (uint32_t)(32 * kb), hi);
        tmem_st_32x32(tlane + (uint32_t)(32 * (KB + kb)), mid);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_full[kb]);
      }
    }
  } else if (warp == PRODUCER_WARP) {
    // ===================== B producer: the ring of (chunk, k-block, hi | mid) slabs =====================
    int cur = 0, j = 0, sbytes = 0, nst = 1;
    uint32_t used = 0, epar = 0;
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      const ItemInfo& nf = info[j % NINFO];
      mbar_wait(&info_full[j % NINFO], (uint32_t)((j / NINFO) & 1));
      {
        const int sb = item_stage_bytes(nf, it);
        if (sb != sbytes) {  // new geometry: every stage in flight must have been consumed before slots move
          for (int sl = 0; sl < MAX_STAGES; ++sl)
            if (used & (1u << sl)) {
              mbar_wait(&b_empty[sl], (epar >> sl) & 1u);
              epar ^= 1u << sl;
            }
          used = 0;
          cur = 0;
          sbytes = sb;
          nst = min(MAX_STAGES, RING_BYTES / sb);
        }
      }
      for (int l = it.l_begin; l < it.l_end; ++l)
      for (int bi = 0; bi < nf.nbox[l]; ++bi) {
        const LevelBox* lbp = &nf.lbox[l][bi];  // fields into registers one by one: a struct copy lands in local memory
        const int b_ncb = lbp->ncb, b_nchunks = lbp->nchunks, b_X0 = lbp->X0, b_Y0 = lbp->Y0;
        if (b_nchunks == 0) continue;
        const CUtensorMap* tmBl = &tmB.m[l * MAX_CB + b_ncb - 1];
        const int b_rows = rows_for(b_ncb);
        const int cb_bytes = b_rows * 1024;
        for (int c = 0; c < b_nchunks; ++c)
          for (int kb = 0; kb < KB; ++kb)
            for (int part = 0; part < 2; ++part) {
              const uint32_t bit = 1u << cur;
              if (used & bit) {
                mbar_wait(&b_empty[cur], (epar >> cur) & 1u);
                epar ^= bit;
              }
              used |= bit;
              if (elect_one()) {
                mbar_expect_tx(&b_full[cur], (uint32_t)(b_ncb * cb_bytes));
                tma_load_4d(tmBl, &b_full[cur], smem_b + cur * sbytes, part * p.D + kb * 64, b_X0, b_Y0 + b_rows * c, it.b);
              }
              __syncwarp();
              if (++cur == nst) cur = 0;
            }
      }
      if (elect_one()) mbar_arrive(&info_empty[j % NINFO]);
      __syncwarp();
    }
  } else if (warp == MMA_WARP) {
    // ===================== MMA issuer: the whole warp walks the loops, one elected lane issues =====================
    const uint64_t b_desc0 = umma_desc(smem_u32(smem_b));
    const uint32_t tmem_a0 = tmem_base + A_COL;
    int cur = 0, g = 0, j = 0, sbytes = 0, nst = 1;  // g: chunks issued so far (accumulator g & 1)
    uint32_t fpar = 0;
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      const ItemInfo& nf = info[j % NINFO];
      mbar_wait(&info_full[j % NINFO], (uint32_t)((j / NINFO) & 1));
      {
        const int sb = item_stage_bytes(nf, it);
        if (sb != sbytes) {
          cur = 0;
          sbytes = sb;
          nst = min(MAX_STAGES, RING_BYTES / sb);
        }
      }
      int last_l = -1, last_bi = 0;
      for (int l = it.l_begin; l < it.l_end; ++l)
        for (int bi = 0; bi < nf.nbox[l]; ++bi)
          if (nf.lbox[l][bi].nchunks > 0) last_l = l, last_bi = bi;
      const bool more = i + gridDim.x < nitems;
      bool first = true;
      for (int l = it.l_begin; l < it.l_end; ++l)
      for (int bi = 0; bi < nf.nbox[l]; ++bi) {
        const int b_ncb = nf.lbox[l][bi].ncb, b_nchunks = nf.lbox[l][bi].nchunks;
        if (b_nchunks == 0) continue;
        const uint32_t idesc = make_idesc_bf16(TQ, b_ncb * rows_for(b_ncb) * 8);
        for (int c = 0; c < b_nchunks; ++c, ++g) {
          const int buf = g & 1;
          const bool last_chunk = more && l == last_l && bi == last_bi && c == b_nchunks - 1;
          mbar_wait(&acc_empty[buf], ((g >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + (uint32_t)buf * ACC_COLS;
          for (int kb = 0; kb < KB; ++kb) {
            const uint32_t a_hi = tmem_a0 + (uint32_t)(32 * kb), a_mid = tmem_a0 + (uint32_t)(32 * (KB + kb));
            if (first) {
              mbar_wait(&a_full[kb], (uint32_t)(j & 1));  // first chunk: the loaders may still be writing this k-block
              tc_fence_after();
            }
            // stage holding B_hi[kb]: hi*hi and mid*hi
            mbar_wait(&b_full[cur], (fpar >> cur) & 1u);
            fpar ^= 1u << cur;
            tc_fence_after();
            uint64_t bs = b_desc0 + (uint64_t)(cur * (sbytes >> 4));
            if (elect_one()) {
              tc_mma_f16_ts_x4(tmem_d, a_hi, bs, idesc, kb != 0);
              tc_mma_f16_ts_x4(tmem_d, a_mid, bs, idesc, 1);
              tc_commit(&b_empty[cur]);
            }
            __syncwarp();
            if (++cur == nst) cur = 0;
            // stage holding B_mid[kb]: hi*mid
            mbar_wait(&b_full[cur], (fpar >> cur) & 1u);
            fpar ^= 1u << cur;
            tc_fence_after();
            bs = b_desc0 + (uint64_t)(cur * (sbytes >> 4));
            if (elect_one()) {
              tc_mma_f16_ts_x4(tmem_d, a_hi, bs, idesc, 1);
              tc_commit(&b_empty[cur]);
              if (last_chunk) tc_commit(&a_empty[kb]);  // the next item's k-block may be written
            }
            __syncwarp();
            if (++cur == nst) cur = 0;
          }
          first = false;
          if (elect_one()) tc_commit(&acc_full[buf]);
          __syncwarp();
        }
      }
      if (last_l < 0) {  // no tensor-core chunk in this item: A must still have landed before it is replaced
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(&a_full[kb], (uint32_t)(j & 1));
          if (more && elect_one()) mbar_arrive(&a_empty[kb]);
          __syncwarp();
        }
      }
      if (elect_one()) mbar_arrive(&info_empty[j % NINFO]);
      __syncwarp();
    }
  } else if (warp < 4) {
    // ===================== extraction warps 0-3: accumulators -> footprints, level by level =====================
    const int m = 32 * quarter + lane;  // the query whose accumulator row this thread's TMEM lane holds
    int g = 0, lv = 0, j = 0;
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      const ItemInfo& nf = info[j % NINFO];
      mbar_wait(&info_full[j % NINFO], (uint32_t)((j / NINFO) & 1));
      for (int l = it.l_begin; l < it.l_end; ++l, ++lv) {
        const int nbox = nf.nbox[l];
        QInfo qi;  // this query's window at level l (the arithmetic of the describing warp)
        {
          const float2 c = nf.coord[m];
          const bool qv = nf.qidx[m] >= 0;
          const int lh_ = l == 0 ? p.lh[0] : l == 1 ? p.lh[1] : l == 2 ? p.lh[2] : p.lh[3];
          const int lw_ = l == 0 ? p.lw[0] : l == 1 ? p.lw[1] : l == 2 ? p.lw[2] : p.lw[3];
          const float sc = __int_as_float((127 - l) << 23);
          const bool normal = fabsf(c.x) <= 1.0e6f && fabsf(c.y) <= 1.0e6f;
          qi.bx = clamp_to_int(floorf(c.x * sc)) - R;
          qi.by = clamp_to_int(floorf(c.y * sc)) - R;
          const bool hw = qv && normal && qi.bx < lw_ && qi.bx + FP > 0 && qi.by < lh_ && qi.by + FP > 0;
          qi.flags = (qv ? 1 : 0) | (hw ? 2 : 0);
          qi.pad = 0;
        }
        const int fb = lv & (NF - 1);
        float* Fl = F + fb * TQ * NCELL;
        float* Fq = Fl + m * NCELL;
        const bool qvalid = (qi.flags & 1) != 0, haswin = (qi.flags & 2) != 0;
        if (lv >= NF) asm volatile("bar.sync %0, 256;" ::"r"(4 + fb) : "memory");  // the taps of level lv - NF are done with it
        if (qvalid && !haswin)
          for (int e = 0; e < NCELL; ++e) Fq[e] = 0.0f;
        if (nbox > 0) {
         for (int bi = 0; bi < nbox; ++bi) {
          const LevelBox* lbp = &nf.lbox[l][bi];
          const int b_ncb = lbp->ncb, b_nchunks = lbp->nchunks, b_BH = lbp->BH, b_rows = rows_for(b_ncb);
          // with two boxes a query extracts from the box of its own half of the patch only
          const bool inbox = haswin && (nbox == 1 || ((lane >> 3) & 1) == bi);
          const int ox = qi.bx - lbp->X0, oy = qi.by - lbp->Y0;  // window origin inside the box
          float* rb = rowbuf + (warp * 32 + lane) * 2 * ROWBUF_PITCH;  // two box rows per thread
          for (int c = 0; c < b_nchunks; ++c, ++g) {
            const int buf = g & 1;
            mbar_wait(&acc_full[buf], (g >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(32 * quarter) << 16) + (uint32_t)buf * ACC_COLS;
#pragma unroll 1
            for (int r = 0; r < b_rows; r += 2) {  // two box rows per TMEM round trip
              const int row = c * b_rows + r;
              if (row >= b_BH) break;
              const bool two = r + 1 < b_rows && row + 1 < b_BH;
              float v[2][8 * MAX_CB];
#pragma unroll
              for (int cb = 0; cb < MAX_CB; ++cb)
                if (cb < b_ncb) {
                  tmem_ld_32x8(taddr + (uint32_t)((r * b_ncb + cb) * 8), v[0] + 8 * cb);
                  if (two) tmem_ld_32x8(taddr + (uint32_t)(((r + 1) * b_ncb + cb) * 8), v[1] + 8 * cb);
                }
              tmem_ld_wait();
#pragma unroll
              for (int rr = 0; rr < 2; ++rr)
#pragma unroll
                for (int cb = 0; cb < MAX_CB; ++cb)
                  if (cb < b_ncb && (rr == 0 || two)) {
                    float* d = rb + rr * ROWBUF_PITCH + 8 * cb;
                    *reinterpret_cast<float4*>(d) = make_float4(v[rr][8 * cb], v[rr][8 * cb + 1], v[rr][8 * cb + 2], v[rr][8 * cb + 3]);
                    *reinterpret_cast<float4*>(d + 4) = make_float4(v[rr][8 * cb + 4], v[rr][8 * cb + 5], v[rr][8 * cb + 6], v[rr][8 * cb + 7]);
                  }
#pragma unroll
              for (int rr = 0; rr < 2; ++rr) {
                const int wr = row + rr - oy;
                if (inbox && (rr == 0 || two) && (unsigned)wr < (unsigned)FP) {
                  float* Fr = Fq + wr * FP;
                  const float* src = rb + rr * ROWBUF_PITCH + ox;
#pragma unroll
                  for (int wc = 0; wc < FP; ++wc) Fr[wc] = src[wc];
                }
              }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[buf]);
          }
         }
        } else {
          // scalar path: the windows of this patch do not fit one box at this level (rough flow); one query per thread
          const bool n_ok = haswin;
          const long long qq = max(nf.qidx[m], 0);
          const int lh = l == 0 ? p.lh[0] : l == 1 ? p.lh[1] : l == 2 ? p.lh[2] : p.lh[3];
          const int lw = l == 0 ? p.lw[0] : l == 1 ? p.lw[1] : l == 2 ? p.lw[2] : p.lw[3];
          const float* f2l = l == 0 ? p.f2[0] : l == 1 ? p.f2[1] : l == 2 ? p.f2[2] : p.f2[3];
          const float* f1q = p.f1 + (n_ok ? qq : 0) * p.D;
          const float* f2b = f2l + (long long)it.b * lh * lw * p.D;
#pragma unroll 1
          for (int e = 0; e < NCELL; ++e) {
            const int ry = e / FP, rx = e - ry * FP;
            const int gy = qi.by + ry, gx = qi.bx + rx;
            float sacc = 0.0f;
            if (n_ok && (unsigned)gy < (unsigned)lh && (unsigned)gx < (unsigned)lw) {
              const float4* cp = reinterpret_cast<const float4*>(f2b + ((long long)gy * lw + gx) * p.D);
              const float4* ap = reinterpret_cast<const float4*>(f1q);
              for (int k = 0; k < p.D / 4; ++k) {
                const float4 vv = __ldg(cp + k), a = __ldg(ap + k);
                sacc = fmaf(a.x, vv.x, sacc);
                sacc = fmaf(a.y, vv.y, sacc);
                sacc = fmaf(a.z, vv.z, sacc);
                sacc = fmaf(a.w, vv.w, sacc);
              }
            }
            Fq[e] = sacc * p.inv_sqrt_d;
          }
        }
        __threadfence_block();
        asm volatile("bar.arrive %0, 256;" ::"r"(2 + fb) : "memory");  // every footprint of the patch at this level is complete
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&info_empty[j % NINFO]);
    }
  } else {
    // ===================== tap warps 4-7: warp 4 + k evaluates the 32 queries of lane quarter k =====================
    const int C = p.L * S * S;
    int lv = 0, j = 0;
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      const ItemInfo& nf = info[j % NINFO];
      mbar_wait(&info_full[j % NINFO], (uint32_t)((j / NINFO) & 1));
      const int m = 32 * quarter + lane;
      const float2 cc = nf.coord[m];
      const int q = nf.qidx[m];
      const bool qvalid = q >= 0;
      for (int l = it.l_begin; l < it.l_end; ++l, ++lv) {
        const int fb = lv & (NF - 1);
        asm volatile("bar.sync %0, 256;" ::"r"(2 + fb) : "memory");
        column_taps_level<R>(F + fb * TQ * NCELL + 32 * quarter * NCELL, cc.x, cc.y, qvalid, q, l, p.out, C, lane);
        if (lv + NF < my_levels) asm volatile("bar.arrive %0, 256;" ::"r"(4 + fb) : "memory");
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&info_empty[j % NINFO]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

static int encode_4d(EncodeTiledFn encode, CUtensorMap* tm, const void* base, int twoD, int w, int h, int B, int box_x, int box_y) {
  const cuuint64_t dims[4] = {(cuuint64_t)twoD, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)B};
  const cuuint64_t strides[3] = {(cuuint64_t)twoD * 2, (cuuint64_t)w * twoD * 2, (cuuint64_t)h * w * twoD * 2};
  const cuuint32_t box[4] = {64, (cuuint32_t)box_x, (cuuint32_t)box_y, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  return encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS
             ? RAFTK_OK
             : RAFTK_ERR_UNSUPPORTED;
}

}  // namespace atc128

// Arguments already validated by raftk_altcorr_lookup_tc (altcorr_tc.cu).
int altcorr_tc128_launch(const float* fmap1, const float* const* f2_levels, const void* f1_split, const void* const* f2_split,
                         int num_levels, const float* coords, float* out, int B, int nq, int h, int w, int D, int radius,
                         void* stream) {
  using namespace raftk::atc128;
  if ((long long)B * nq > 0x7fffffffLL) return RAFTK_ERR_UNSUPPORTED;
  EncodeTiledFn encode = get_encode_fn();
  if (!encode) return RAFTK_ERR_UNSUPPORTED;
  Params p{};
  BMaps tmB;
  memset(&tmB, 0, sizeof(tmB));
  const int rows_q = nq / w;
  for (int l = 0; l < 4; ++l) {
    const int ll = l < num_levels ? l : num_levels - 1;
    p.f2[l] = f2_levels[ll];
    p.lh[l] = h >> ll;
    p.lw[l] = w >> ll;
    for (int ncb = 1; ncb <= MAX_CB; ++ncb) {
      const int rc = encode_4d(encode, &tmB.m[l * MAX_CB + ncb - 1], f2_split[ll], 2 * D, w >> ll, h >> ll, B, 8 * ncb, rows_for(ncb));
      if (rc != RAFTK_OK) return rc;
    }
  }
  p.f1 = fmap1;
  p.f1_split = static_cast<const __nv_bfloat16*>(f1_split);
  p.coords = coords;
  p.out = out;
  p.B = B;
  p.nq = nq;
  p.D = D;
  p.L = num_levels;
  p.rows_q = rows_q;
  p.w_q = w;
  p.inv_sqrt_d = 1.0f / (float)sqrt((double)D);
  const int FP = 2 * radius + 2;
  const size_t smem = 1024 + (size_t)RING_BYTES + (size_t)NF * TQ * FP * FP * sizeof(float) +
                      (size_t)4 * 32 * 2 * ROWBUF_PITCH * sizeof(float) + 448 + NINFO * sizeof(ItemInfo);
  if (smem > (size_t)SMEM_LIMIT) return RAFTK_ERR_UNSUPPORTED;
  const long long nblk = (long long)B * ((rows_q + PH - 1) / PH) * ((w + PW - 1) / PW);
  if (nblk > 0x7fffffffLL) return RAFTK_ERR_SHAPE;
  {
    const long long S_ = sm_count();
    const long long full = nblk / S_, tail = nblk - full * S_;
    if (nblk < S_) p.split_from = 0;
    else if (tail == 0 || 4 * tail >= 3 * S_) p.split_from = (int)nblk;
    else p.split_from = (int)(full * S_);
  }
  p.npatch = (int)nblk;
  const long long ncta = p.split_from + (nblk - p.split_from) * num_levels;
  if (ncta > 0x7fffffffLL) return RAFTK_ERR_SHAPE;
  auto kernel = radius == 4 ? altcorr_tc128_kernel<4> : altcorr_tc128_kernel<3>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  const long long grid = ncta < (long long)sm_count() ? ncta : (long long)sm_count();
  kernel<<<(unsigned)grid, NUM_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(tmB, p);
  return launch_status();
}

}  // namespace raftk
