// Kernel family 2b on the tensor cores: the on-the-fly ("alt_corr") window lookup as small tcgen05 GEMMs.
// No reference counterpart; by linearity of the average pool
//   pool_l(f1 . f2^T)[q, Y, X] = f1[q] . pool_l(f2)[Y, X]
// the taps of level l (model.py:448-470) are bilinear samples of dot products against fmap2 pooled l times.
//
// One work item = one 8 x 8 patch of query pixels, all levels in turn (UMMA M = 64, cta_group::1); the CTAs are
// persistent and walk the items.  Neighbouring windows overlap, so per level the CTA computes the correlations of its
// 64 queries with every cell of the BOUNDING BOX of their (2r+2)^2 windows — a (64 x box cells x D) GEMM — and each
// query then picks its own window out of the box:
//
//   operands   bf16 [hi | mid] splits of fmap1 (scaled by 1/sqrt(D)) and of every pooled fmap2 level, written once
//              per pair by raftk_altcorr_split; products hi*hi + mid*hi + hi*mid in float32 (the scheme of the
//              volume GEMM, corr_volume_tc.cu: ~2e-5 relative).
//   A          the 64 queries x 2D tile, resident for all levels: 2 * D/64 TMA boxes {64 ch, 8 x, 8 y} (64 KB at D = 256).
//   B          the box, in CHUNKS of 4-6 box rows (per level, chosen on the host to pad the typical box least); a chunk
//              is up to four column blocks of 8 cells, each one TMA box {64 ch, 8 cells, rows}, so N = 8 * rows * (1..4)
//              <= 192 and a box may be up to 32 cells wide and 6 chunks tall (TMA zero-fills cells outside the level:
//              those correlations are 0, which is what the reference's zero-outside sampling makes of them).  One
//              (chunk, k-block, hi | mid) slab per stage of a ring of as many stages as fit 80 KB (4 ... 8).
//   D          float32 in TMEM, two accumulators of 192 columns: the epilogue of chunk c runs under the MMAs of
//              chunk c + 1, across level boundaries too.  M = 64: accumulator row m is TMEM lane (m % 16) + 32 * (m / 16).
//   epilogue   extraction warps 0-3 (TMEM lane quarter = warp: its 16 queries).  Two box rows per TMEM round trip:
//              tcgen05.ld.x8 per column block (static register indices), the rows go to the thread's row buffer in
//              shared memory, and the <= 2r+2 cells of the thread's own window are copied from there into its
//              footprint F[query][(2r+2)^2] — the window offset is a shared-memory address, never a register index.
//   taps       warps 4-7.  Footprints are double-buffered per level and handed over through named barriers: when the
//              last chunk of level l is in, warp 4 + k evaluates the taps of quarter k as column tasks (the schedule
//              and the float32 arithmetic of corr_lookup_tiled.cu: window indexing and tap values are those of the
//              volume lookup) while the extraction warps drain level l + 1 and the MMAs fill the accumulators.
//
// A cell's value depends only on the query's row of A, the cell's row of B and the fixed K order, not on the box or
// the chunk it was computed in, so row bands reproduce the full result bit for bit.  Queries that read nothing (far or
// non-finite coordinates, windows entirely outside the level) get a zero footprint and stay out of the box.  Boxes
// that do not fit (rough flow: wider than 32 cells or taller than 6 chunks) take a scalar fp32 path inside the same
// kernel (sequential fma chain over D, as altcorr.cu), slower and ~2e-5 away from the split product.
//
//   items      warp 10 describes item j + 1 (coordinates, window origins, per-level boxes: two info slots) and
//              fetches its A tile k-block by k-block as the last chunk of item j releases it (a_empty barriers), so
//              the B producer (warp 8) never stops streaming and the MMAs of item j + 1 follow those of item j;
//              B ring, accumulators, footprint buffers and barrier phases carry over item boundaries.  Whole waves
//              of patches are one item each; the patches of a last, partial wave are (patch, level) items.
//
// Measured, 136 x 240, r = 4, D = 256, smooth +-8 px flow (profiles/altcorr_time.py): 114 us per iteration against
// 560 us for the float32 FMA kernel (altcorr.cu).  History and phase elimination: DESIGN.md §4.4.  Two lessons:
// (1) the ISSUE path — with `if (lane == 0)` around the MMA / TMA loops the compiler wraps every tcgen05.mma in a
// vote-and-broadcast loop (~23 instructions of the one issuing thread per ~60-cycle MMA); the loops now run
// warp-uniformly and issue under elect.sync, four MMAs per asm statement (tc_ptx.cuh): 171 -> 137 us; (2) per-item
// fixed cost (set-up 5.7 us before the first accumulator, 4.6 us of tail after the last MMA, of ~32 us) — persistent
// CTAs with the next item prepared one ahead: 128 -> 114 us.  A second kernel with UMMA M = 128 and the A operand in
// tensor memory (altcorr_tc128.cu) is chosen for calls whose 16 x 8 patches fill the SMs (raftk_set_altcorr_mode);
// the phase elimination behind that split — three quarters of this kernel's time is the stage pipeline between the
// B producer and the MMA issuer, not the tensor pipe — is in DESIGN.md §4.4.

#include <cuda.h>
#include <cuda_bf16.h>

#include <cstdlib>
#include <cstring>

#include "altcorr_tc128.cuh"
#include "lookup_taps.cuh"
#include "raftk_common.cuh"
#include "tc_ptx.cuh"

namespace raftk {
namespace atc {

using namespace raftk::ptx;

constexpr int TQ = 64;                          // queries per CTA: 8 x 8
constexpr int MAX_ROWS = 6;                     // box rows per chunk: 4, 5 or 6, chosen per level on the host
constexpr int MAX_CB = 4;                       // column blocks of 8 cells: box width <= 32
constexpr int MAX_CHUNKS = 6;                   // box height <= 6 * rows per chunk
constexpr int RING_BYTES = 80 * 1024;           // B ring: as many stages as fit (3 of 24 KB ... 8 of 10 KB)
constexpr int ROWBUF_PITCH = 36;                // floats per lane of the row buffer (+4: 16-byte stores of 8 lanes hit 32 banks)
constexpr int MAX_STAGES = 8;
constexpr int A_BLOCK_BYTES = TQ * 128;
constexpr int ACC_COLS = MAX_CB * MAX_ROWS * 8;  // 192 accumulator columns per buffer
constexpr int EPI_WARPS = 8;                    // warps 0-3: window extraction (TMEM lane quarter = warp); warps 4-7: taps of quarter warp - 4
constexpr int PRODUCER_WARP = EPI_WARPS, MMA_WARP = EPI_WARPS + 1, AUX_WARP = EPI_WARPS + 2;
constexpr int NUM_THREADS = (EPI_WARPS + 3) * 32;  // + B producer, MMA issuer, item-description / A-tile producer
constexpr int SMEM_LIMIT = 232448;

struct Params {
  const float* f1;      // float32 sources, used by the scalar path only
  const float* f2[4];
  int lh[4], lw[4];
  int rows[4];          // box rows per chunk of each level (= the y extent of that level's TMA box)
  const float* coords;
  float* out;
  int B, nq, D, L;
  int rows_q, w_q;
  int npatch;           // 8 x 8 query patches
  int split_from;       // patches [0, split_from): one CTA per patch, all levels; the rest: one CTA per (patch, level)
  float inv_sqrt_d;
};

struct QInfo {
  int bx, by, flags, pad;  // flags: bit 0 valid query, bit 1 the query has a window (usable coordinates, inside the level)
};

// float32 rows -> [hi | mid] bf16 rows (x * scale ~ hi + mid to 2^-17)
__global__ void __launch_bounds__(256) split_rows_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                                         long long rows, int D, float scale) {
  const int d4 = D >> 2;
  const long long total = rows * d4;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long row = t / d4;
    const int c = (int)(t - row * d4) * 4;
    const float4 x = __ldg(reinterpret_cast<const float4*>(src + row * D + c));
    const float xs[4] = {x.x * scale, x.y * scale, x.z * scale, x.w * scale};
    __nv_bfloat16 hi[4], mid[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      hi[i] = __float2bfloat16_rn(xs[i]);
      mid[i] = __float2bfloat16_rn(xs[i] - __bfloat162float(hi[i]));
    }
    *reinterpret_cast<uint2*>(dst + row * 2 * D + c) = *reinterpret_cast<const uint2*>(hi);
    *reinterpret_cast<uint2*>(dst + row * 2 * D + D + c) = *reinterpret_cast<const uint2*>(mid);
  }
}

// The (2r+1)^2 taps of level l for the 16 queries of one warp, as COLUMN TASKS (the schedule of corr_lookup_tiled.cu):
// a lane owns (query k, x offset i) and walks the 2r+1 y offsets, reusing the two footprint values of row j + 1 as
// row j of the next tap; x weights once per task.  Arithmetic per tap as the reference evaluates it (jax
// map_coordinates order 1, float32, no contraction), anchored at the regular cell floor(c_l) + d:
//   s = c_l + d;  w1 = s - (floor(c_l) + d);  w0 = 1 - w1;  out = ((wy0 wx0) v00 + (wy0 wx1) v01) + (wy1 wx0) v10 + (wy1 wx1) v11
// (identical to i0 = floor(s) except when fl(c_l + d) rounds up to an integer, where the same products appear with the
// unit weight on the other corner: corr_lookup_tiled.cu, "Arithmetic").  Fw: the warp's footprints [16][(2r+2)^2],
// origin floor(c_l) - r; (cx, cy, qvalid, q) are those of the query THIS lane owns (lanes 0-15).
template <int R>
__device__ __forceinline__ void column_taps_level(const float* Fw, float cx, float cy, bool qvalid, long long q, int l,
                                                  float* out, int C, int lane) {
  constexpr int S = 2 * R + 1, FP = 2 * R + 2, NCELL = FP * FP;
  const float sc = __int_as_float((127 - l) << 23);  // 2^-l, exact
#pragma unroll 1
  for (int base = 0; base < 16 * S; base += 32) {
    const int T = base + lane;
    const bool active = T < 16 * S;
    const int k = active ? T / S : 0, i = T - k * S;
    const float ccx = __shfl_sync(0xffffffffu, cx, k), ccy = __shfl_sync(0xffffffffu, cy, k);
    const int ok = __shfl_sync(0xffffffffu, qvalid ? 1 : 0, k);
    const long long qq = __shfl_sync(0xffffffffu, q, k);
    if (!active || !ok) continue;
    const float cxl = ccx * sc, cyl = ccy * sc;
    const float fcx = floorf(cxl), fcy = floorf(cyl);
    const float dx = (float)(i - R);
    const float wx1 = __fadd_rn(__fadd_rn(cxl, dx), -__fadd_rn(fcx, dx));
    const float wx0 = __fadd_rn(1.0f, -wx1);
    const float* Fc = Fw + k * NCELL + i;
    float* orow = out + qq * C + l * S * S + i * S;
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

// what every role needs to know about one level of a patch
struct LevelBox {
  int X0, Y0, BW, BH;   // bounding box of the windows (BW = 0: no query of the patch has a window at this level)
  int tc;               // 1: tensor-core path (the box fits), 0: scalar path
  int ncb, nchunks, rows;
};

// One work item's description, written by the producer warp one item ahead of the other roles (two slots).
struct ItemInfo {
  LevelBox lbox[4];
  QInfo qinfo[4][TQ];
  float2 coord[TQ];
  long long qidx[TQ];
  int valid[TQ];
};

struct Item {
  int txi, tyi, b, l_begin, l_end;
};

// Work items [0, split_from): one patch, all levels; the items after them: one (patch, level), level 0 first — the
// short items fill what would be a last, partial wave of patches (and all of a small problem).
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

// Ring geometry of one item: a stage holds the widest chunk of the item's tensor-core levels; as many stages as fit.
__device__ __forceinline__ int item_stage_bytes(const ItemInfo& nf, const Item& it) {
  int sb = 8 * 1024;
  for (int l = it.l_begin; l < it.l_end; ++l)
    if (nf.lbox[l].tc) sb = max(sb, nf.lbox[l].ncb * nf.lbox[l].rows * 1024);
  return sb;
}

template <int R>
__global__ void __launch_bounds__(NUM_THREADS, 1)
altcorr_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB0,
                  const __grid_constant__ CUtensorMap tmB1, const __grid_constant__ CUtensorMap tmB2,
                  const __grid_constant__ CUtensorMap tmB3, const Params p) {
  constexpr int FP = 2 * R + 2, S = 2 * R + 1, NCELL = FP * FP;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int KB = p.D / 64;  // k-blocks per plane (hi, mid)
  uint8_t* smem_a = smem;                                   // [2 KB][A_BLOCK_BYTES]: hi blocks, then mid blocks
  uint8_t* smem_b = smem_a + 2 * KB * A_BLOCK_BYTES;        // ring of RING_BYTES
  float* F = reinterpret_cast<float*>(smem_b + RING_BYTES);  // [2][TQ][NCELL]: footprints of two consecutive levels
  float* rowbuf = F + 2 * TQ * NCELL;                        // [4 warps][16 owner lanes][2 rows][ROWBUF_PITCH]
  uint64_t* bars = reinterpret_cast<uint64_t*>(rowbuf + 4 * 16 * 2 * ROWBUF_PITCH);
  uint64_t* a_full = bars;                      // [4]: one per k-block (its hi and mid tiles)
  uint64_t* a_empty = bars + 4;                 // [4]: the item's last MMAs are done with this k-block of A
  uint64_t* b_full = bars + 8;                  // [MAX_STAGES]
  uint64_t* b_empty = b_full + MAX_STAGES;      // [MAX_STAGES]
  uint64_t* acc_full = b_empty + MAX_STAGES;    // [2]
  uint64_t* acc_empty = acc_full + 2;           // [2]
  uint64_t* info_full = acc_empty + 2;          // [2]: the producer has described the item in this slot
  uint64_t* info_empty = info_full + 2;         // [2]: MMA issuer + 4 extraction + 4 tap warps are done with it
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(info_empty + 2);
  ItemInfo* info = reinterpret_cast<ItemInfo*>(bars + 40);  // [2]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_x = (p.w_q + 7) >> 3, tiles_y = (p.rows_q + 7) >> 3;
  const long long nitems = p.split_from + (long long)(p.npatch - p.split_from) * p.L;
  // persistent: this CTA walks the items blockIdx.x, blockIdx.x + gridDim.x, ...; the roles below run one continuous
  // pipeline across items (B ring, accumulators, footprint buffers and barrier phases carry over), the producer
  // describes item j + 1 and fetches its A tile while item j is being multiplied
  int my_levels = 0;  // levels this CTA will process in total (the tap warps' hand-back of a footprint buffer needs it)
  for (long long i = blockIdx.x; i < nitems; i += gridDim.x) my_levels += i < p.split_from ? p.L : 1;

  if (warp == MMA_WARP && lane == 0) {
    for (int kb = 0; kb < 4; ++kb) {
      mbar_init(&a_full[kb], 1);
      mbar_init(&a_empty[kb], 1);
    }
    for (int s = 0; s < MAX_STAGES; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 4);  // one arrival per extraction warp
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
    // ===================== item descriptions and the A tile, one item ahead of everybody else =====================
    // (a warp of its own: the B producer below must not stop streaming while this one waits for global loads or for
    // the previous item to release the A tile; the whole warp walks the loops, one elected lane issues: tc_ptx.cuh)
    int j = 0;
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      ItemInfo& nf = info[j & 1];
      if (j >= 2) mbar_wait(&info_empty[j & 1], (uint32_t)(((j >> 1) - 1) & 1));
      // ---- the queries of this patch, their window origins, and per level the bounding box of the windows ----
      // haswin: finite, not absurdly far, and the window overlaps the level.  Queries without a window read nothing —
      // their footprint is all zeros whatever path a patch takes — and stay out of the bounding box, so that one
      // stray query does not change how its neighbours are computed.  Lane t describes queries t and t + 32.
      float cx[2], cy[2];
      bool qv[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int m = lane + 32 * u;
        const int qy = it.tyi * 8 + (m >> 3), qx = it.txi * 8 + (m & 7);
        qv[u] = qy < p.rows_q && qx < p.w_q;
        const long long q = (long long)it.b * p.nq + (long long)qy * p.w_q + qx;
        cx[u] = cy[u] = 0.0f;
        if (qv[u]) {
          cx[u] = __ldg(p.coords + 2 * q);
          cy[u] = __ldg(p.coords + 2 * q + 1);
        }
        nf.coord[m] = make_float2(cx[u], cy[u]);
        nf.qidx[m] = q;
        nf.valid[m] = qv[u] ? 1 : 0;
      }
      for (int l = it.l_begin; l < it.l_end; ++l) {
        const int lh = l == 0 ? p.lh[0] : l == 1 ? p.lh[1] : l == 2 ? p.lh[2] : p.lh[3];
        const int lw = l == 0 ? p.lw[0] : l == 1 ? p.lw[1] : l == 2 ? p.lw[2] : p.lw[3];
        const float sc = __int_as_float((127 - l) << 23);
        int mnx = 0x7fffffff, mxx = -0x7fffffff, mny = 0x7fffffff, mxy = -0x7fffffff, any = 0;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const bool normal = fabsf(cx[u]) <= 1.0e6f && fabsf(cy[u]) <= 1.0e6f;
          const int bx = clamp_to_int(floorf(cx[u] * sc)) - R, by = clamp_to_int(floorf(cy[u] * sc)) - R;
          const bool hw = qv[u] && normal && bx < lw && bx + FP > 0 && by < lh && by + FP > 0;
          nf.qinfo[l][lane + 32 * u] = QInfo{bx, by, (qv[u] ? 1 : 0) | (hw ? 2 : 0), 0};
          if (hw) {
            mnx = min(mnx, bx);
            mxx = max(mxx, bx);
            mny = min(mny, by);
            mxy = max(mxy, by);
            any = 1;
          }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
          mnx = min(mnx, __shfl_xor_sync(0xffffffffu, mnx, off));
          mxx = max(mxx, __shfl_xor_sync(0xffffffffu, mxx, off));
          mny = min(mny, __shfl_xor_sync(0xffffffffu, mny, off));
          mxy = max(mxy, __shfl_xor_sync(0xffffffffu, mxy, off));
          any |= __shfl_xor_sync(0xffffffffu, any, off);
        }
        if (lane == 0) {
          LevelBox lb;
          lb.X0 = any ? mnx : 0;
          lb.Y0 = any ? mny : 0;
          lb.BW = any ? mxx - mnx + FP : 0;  // width in cells
          lb.BH = any ? mxy - mny + FP : 0;  // height
          lb.rows = p.rows[l];
          lb.tc = (lb.BW <= 8 * MAX_CB && lb.BH <= lb.rows * MAX_CHUNKS) ? 1 : 0;
          lb.ncb = (lb.BW + 7) >> 3;
          lb.nchunks = (lb.BH + lb.rows - 1) / lb.rows;
          nf.lbox[l] = lb;
        }
      }
      __threadfence_block();
      __syncwarp();
      if (elect_one()) mbar_arrive(&info_full[j & 1]);
      __syncwarp();
      // ---- the A tile, one barrier per k-block; its buffer is free once the previous item's last MMAs have read it ----
      for (int kb = 0; kb < KB; ++kb) {
        if (j >= 1) mbar_wait(&a_empty[kb], (uint32_t)((j - 1) & 1));
        if (elect_one()) {
          mbar_expect_tx(&a_full[kb], (uint32_t)(2 * A_BLOCK_BYTES));  // rows outside the image are zero-filled and counted
          tma_load_4d(&tmA, &a_full[kb], smem_a + kb * A_BLOCK_BYTES, kb * 64, it.txi * 8, it.tyi * 8, it.b);              // hi
          tma_load_4d(&tmA, &a_full[kb], smem_a + (KB + kb) * A_BLOCK_BYTES, p.D + kb * 64, it.txi * 8, it.tyi * 8, it.b);  // mid
        }
        __syncwarp();
      }
    }
  } else if (warp == PRODUCER_WARP) {
    // ===================== B producer: the ring of (chunk, k-block, hi | mid) slabs =====================
    // Every slot has its own full / empty barrier; `used`: the slot holds a stage whose release has not been awaited
    // yet, `epar`: phase bit of its empty barrier.  Consecutive items almost always share the ring geometry; when it
    // changes the ring is drained first.
    int cur = 0, j = 0, sbytes = 0, nst = 1;
    uint32_t used = 0, epar = 0;
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      const ItemInfo& nf = info[j & 1];
      mbar_wait(&info_full[j & 1], (uint32_t)((j >> 1) & 1));
      // ---- the B ring: one (chunk, k-block, hi | mid) slab per stage ----
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
      for (int l = it.l_begin; l < it.l_end; ++l) {
        const LevelBox lb = nf.lbox[l];
        if (!lb.tc) continue;
        const CUtensorMap* tmB = l == 0 ? &tmB0 : l == 1 ? &tmB1 : l == 2 ? &tmB2 : &tmB3;
        const int cb_bytes = lb.rows * 1024;
        for (int c = 0; c < lb.nchunks; ++c)
          for (int kb = 0; kb < KB; ++kb)
            for (int part = 0; part < 2; ++part) {
              const uint32_t bit = 1u << cur;
              if (used & bit) {
                mbar_wait(&b_empty[cur], (epar >> cur) & 1u);
                epar ^= bit;
              }
              used |= bit;
              if (elect_one()) {
                mbar_expect_tx(&b_full[cur], (uint32_t)(lb.ncb * cb_bytes));
                for (int cb = 0; cb < lb.ncb; ++cb)
                  tma_load_4d(tmB, &b_full[cur], smem_b + cur * sbytes + cb * cb_bytes, part * p.D + kb * 64,
                              lb.X0 + 8 * cb, lb.Y0 + lb.rows * c, it.b);
              }
              __syncwarp();
              if (++cur == nst) cur = 0;
            }
      }
      if (elect_one()) mbar_arrive(&info_empty[j & 1]);
      __syncwarp();
    }
  } else if (warp == MMA_WARP) {
    // ===================== MMA issuer: the whole warp walks the loops, one elected lane issues =====================
    const uint64_t a_desc0 = umma_desc(smem_u32(smem_a)), b_desc0 = umma_desc(smem_u32(smem_b));
    int cur = 0, g = 0, j = 0, sbytes = 0, nst = 1;  // g: chunks issued so far (accumulator g & 1)
    uint32_t fpar = 0;                               // per ring slot: phase bit of its full barrier
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      const ItemInfo& nf = info[j & 1];
      mbar_wait(&info_full[j & 1], (uint32_t)((j >> 1) & 1));
      {
        const int sb = item_stage_bytes(nf, it);  // the producer's ring geometry for this item
        if (sb != sbytes) {
          cur = 0;
          sbytes = sb;
          nst = min(MAX_STAGES, RING_BYTES / sb);
        }
      }
      // the item's last tensor-core chunk hands the A tile back (none: plain arrivals below)
      int last_l = -1;
      for (int l = it.l_begin; l < it.l_end; ++l)
        if (nf.lbox[l].tc && nf.lbox[l].nchunks > 0) last_l = l;
      const bool more = i + gridDim.x < nitems;  // another item follows: its A tile waits for this one's release
      bool first = true;
      for (int l = it.l_begin; l < it.l_end; ++l) {
        const LevelBox lb = nf.lbox[l];
        if (!lb.tc) continue;
        const uint32_t idesc = make_idesc_bf16(TQ, lb.ncb * lb.rows * 8);
        for (int c = 0; c < lb.nchunks; ++c, ++g) {
          const int buf = g & 1;
          const bool last_chunk = more && l == last_l && c == lb.nchunks - 1;
          mbar_wait(&acc_empty[buf], ((g >> 1) & 1) ^ 1);  // the extraction has drained this accumulator
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + (uint32_t)buf * ACC_COLS;
          for (int kb = 0; kb < KB; ++kb) {
            // descriptors differ from their base only in the address field (bytes >> 4, no carry out of it below 256 KB)
            const uint64_t a_hi = a_desc0 + (uint64_t)(kb * (A_BLOCK_BYTES >> 4));
            const uint64_t a_mid = a_desc0 + (uint64_t)((KB + kb) * (A_BLOCK_BYTES >> 4));
            // stage holding B_hi[kb]: hi*hi and mid*hi
            if (first) mbar_wait(&a_full[kb], (uint32_t)(j & 1));  // first chunk: the A tile may still be arriving
            mbar_wait(&b_full[cur], (fpar >> cur) & 1u);
            fpar ^= 1u << cur;
            tc_fence_after();
            uint64_t bs = b_desc0 + (uint64_t)(cur * (sbytes >> 4));
            if (elect_one()) {
              tc_mma_f16_x4(tmem_d, a_hi, bs, idesc, kb != 0);
              tc_mma_f16_x4(tmem_d, a_mid, bs, idesc, 1);
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
              tc_mma_f16_x4(tmem_d, a_hi, bs, idesc, 1);
              tc_commit(&b_empty[cur]);
              if (last_chunk) tc_commit(&a_empty[kb]);  // the next item's A block may land here
            }
            __syncwarp();
            if (++cur == nst) cur = 0;
          }
          first = false;
          if (elect_one()) tc_commit(&acc_full[buf]);
          __syncwarp();
        }
      }
      if (last_l < 0) {  // no tensor-core chunk in this item: the A tile must still have landed before it is replaced
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(&a_full[kb], (uint32_t)(j & 1));
          if (more && elect_one()) mbar_arrive(&a_empty[kb]);
          __syncwarp();
        }
      }
      if (elect_one()) mbar_arrive(&info_empty[j & 1]);
      __syncwarp();
    }
  } else if (warp < 4) {
    // ===================== extraction warps 0-3: accumulators -> footprints, level by level =====================
    // F is double-buffered per level and handed to the tap warps through named barriers: 2 + buf "footprints of a
    // level complete" (extraction arrives, taps wait), 4 + buf "buffer free again" (taps arrive, extraction waits).
    const int m = 16 * quarter + (lane & 15);  // the query whose accumulator row this thread's TMEM lane holds
    const bool owner = lane < 16;
    int g = 0, lv = 0, j = 0;  // chunks, levels, items so far
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      const ItemInfo& nf = info[j & 1];
      mbar_wait(&info_full[j & 1], (uint32_t)((j >> 1) & 1));
      for (int l = it.l_begin; l < it.l_end; ++l, ++lv) {
        const LevelBox lb = nf.lbox[l];
        const QInfo qi = nf.qinfo[l][m];
        const int fb = lv & 1;
        float* Fl = F + fb * TQ * NCELL;
        float* Fq = Fl + m * NCELL;
        const bool qvalid = owner && (qi.flags & 1), haswin = owner && (qi.flags & 2);
        if (lv >= 2) asm volatile("bar.sync %0, 256;" ::"r"(4 + fb) : "memory");  // the taps of level lv - 2 are done with it
        if (qvalid && !haswin)
          for (int e = 0; e < NCELL; ++e) Fq[e] = 0.0f;
        if (lb.tc) {
          const int ox = qi.bx - lb.X0, oy = qi.by - lb.Y0;  // window origin inside the box
          float* rb = rowbuf + (warp * 16 + (lane & 15)) * 2 * ROWBUF_PITCH;  // two box rows per thread
          for (int c = 0; c < lb.nchunks; ++c, ++g) {
            const int buf = g & 1;
            mbar_wait(&acc_full[buf], (g >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(32 * quarter) << 16) + (uint32_t)buf * ACC_COLS;
#pragma unroll 1
            for (int r = 0; r < lb.rows; r += 2) {  // two box rows per TMEM round trip
              const int row = c * lb.rows + r;
              if (row >= lb.BH) break;
              const bool two = r + 1 < lb.rows && row + 1 < lb.BH;
              float v[2][8 * MAX_CB];
#pragma unroll
              for (int cb = 0; cb < MAX_CB; ++cb)
                if (cb < lb.ncb) {
                  tmem_ld_32x8(taddr + (uint32_t)(cb * lb.rows * 8 + r * 8), v[0] + 8 * cb);
                  if (two) tmem_ld_32x8(taddr + (uint32_t)(cb * lb.rows * 8 + (r + 1) * 8), v[1] + 8 * cb);
                }
              tmem_ld_wait();
              // the rows go to this thread's row buffer (16-byte stores, static register indices); the <= 2r+2 cells
              // of the thread's own window are then copied out at a dynamic shared-memory offset
#pragma unroll
              for (int rr = 0; rr < 2; ++rr)
#pragma unroll
                for (int cb = 0; cb < MAX_CB; ++cb)
                  if (cb < lb.ncb && owner && (rr == 0 || two)) {
                    float* d = rb + rr * ROWBUF_PITCH + 8 * cb;
                    *reinterpret_cast<float4*>(d) = make_float4(v[rr][8 * cb], v[rr][8 * cb + 1], v[rr][8 * cb + 2], v[rr][8 * cb + 3]);
                    *reinterpret_cast<float4*>(d + 4) = make_float4(v[rr][8 * cb + 4], v[rr][8 * cb + 5], v[rr][8 * cb + 6], v[rr][8 * cb + 7]);
                  }
#pragma unroll
              for (int rr = 0; rr < 2; ++rr) {
                const int wr = row + rr - oy;
                if (haswin && (rr == 0 || two) && (unsigned)wr < (unsigned)FP) {
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
        } else {
          // scalar path: the windows of this patch do not fit one box at this level (rough flow)
          const int t = threadIdx.x, qj = t >> 1, sub = t & 1;
          const QInfo qs = nf.qinfo[l][qj];
          const bool n_ok = (qs.flags & 2) != 0;
          const long long qq = nf.qidx[qj];
          const int lh = l == 0 ? p.lh[0] : l == 1 ? p.lh[1] : l == 2 ? p.lh[2] : p.lh[3];
          const int lw = l == 0 ? p.lw[0] : l == 1 ? p.lw[1] : l == 2 ? p.lw[2] : p.lw[3];
          const float* f2l = l == 0 ? p.f2[0] : l == 1 ? p.f2[1] : l == 2 ? p.f2[2] : p.f2[3];
          const float* f1q = p.f1 + (n_ok ? qq : 0) * p.D;
          const float* f2b = f2l + (long long)it.b * lh * lw * p.D;
#pragma unroll 1
          for (int e = sub; e < NCELL; e += 2) {
            const int ry = e / FP, rx = e - ry * FP;
            const int gy = qs.by + ry, gx = qs.bx + rx;
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
            Fl[qj * NCELL + e] = sacc * p.inv_sqrt_d;
          }
        }
        __threadfence_block();
        asm volatile("bar.arrive %0, 256;" ::"r"(2 + fb) : "memory");  // every footprint of the patch at this level is complete
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&info_empty[j & 1]);
    }
  } else {
    // ===================== tap warps 4-7: warp 4 + k evaluates the 16 queries of lane quarter k =====================
    // (under the extraction of the next level and the MMAs of the one after)
    const int C = p.L * S * S;
    int lv = 0, j = 0;
    for (long long i = blockIdx.x; i < nitems; i += gridDim.x, ++j) {
      const Item it = decode_item(p, i, tiles_x, tiles_y);
      const ItemInfo& nf = info[j & 1];
      mbar_wait(&info_full[j & 1], (uint32_t)((j >> 1) & 1));
      const int m = 16 * quarter + (lane & 15);
      const float2 cc = nf.coord[m];
      const bool qvalid = lane < 16 && nf.valid[m] != 0;
      const long long q = nf.qidx[m];
      for (int l = it.l_begin; l < it.l_end; ++l, ++lv) {
        const int fb = lv & 1;
        asm volatile("bar.sync %0, 256;" ::"r"(2 + fb) : "memory");
        column_taps_level<R>(F + fb * TQ * NCELL + 16 * quarter * NCELL, cc.x, cc.y, qvalid, q, l, p.out, C, lane);
        if (lv + 2 < my_levels) asm volatile("bar.arrive %0, 256;" ::"r"(4 + fb) : "memory");
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&info_empty[j & 1]);
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

}  // namespace atc
}  // namespace raftk

namespace raftk {
int& altcorr_mode_flag() {
  static int flag = [] {
    const char* e = getenv("RAFTK_ALTCORR_MODE");
    return (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 0;
  }();
  return flag;
}
}  // namespace raftk

extern "C" int raftk_set_altcorr_mode(int mode) {
  int& f = raftk::altcorr_mode_flag();
  const int prev = f;
  if (mode >= 0 && mode <= 2) f = mode;
  return prev;
}

extern "C" int raftk_altcorr_split(const float* src, void* dst, long long rows, int D, float scale, void* stream) {
  using namespace raftk;
  if (rows < 0 || D <= 0) return RAFTK_ERR_SHAPE;
  if (D % 4) return RAFTK_ERR_UNSUPPORTED;
  if (rows == 0) return RAFTK_OK;
  if (!src || !dst) return RAFTK_ERR_NULL;
  if (((uintptr_t)src & 15) || ((uintptr_t)dst & 15)) return RAFTK_ERR_ALIGN;
  const long long total = rows * (D / 4);
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)sm_count() * 8) blocks = (long long)sm_count() * 8;
  atc::split_rows_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, static_cast<__nv_bfloat16*>(dst), rows, D, scale);
  return launch_status();
}

extern "C" int raftk_altcorr_lookup_tc(const float* fmap1, const float* const* f2_levels, const void* f1_split,
                                       const void* const* f2_split, int num_levels, const float* coords, float* out,
                                       int B, int nq, int h, int w, int D, int radius, void* stream) {
  using namespace raftk;
  using namespace raftk::atc;
  if (num_levels < 1 || num_levels > RAFTK_MAX_LEVELS || B < 0 || nq < 0 || h <= 0 || w <= 0 || D <= 0 || radius < 0)
    return RAFTK_ERR_SHAPE;
  if ((h >> (num_levels - 1)) < 1 || (w >> (num_levels - 1)) < 1) return RAFTK_ERR_TOO_SMALL;
  if (num_levels > 4 || (radius != 3 && radius != 4) || D % 64 || D > 256 || nq % w) return RAFTK_ERR_UNSUPPORTED;
  if ((long long)B * nq == 0) return RAFTK_OK;
  if (!fmap1 || !f2_levels || !f1_split || !f2_split || !coords || !out) return RAFTK_ERR_NULL;
  if (((uintptr_t)fmap1 & 15) || ((uintptr_t)f1_split & 15)) return RAFTK_ERR_ALIGN;
  for (int l = 0; l < num_levels; ++l) {
    if (!f2_levels[l] || !f2_split[l]) return RAFTK_ERR_NULL;
    if (((uintptr_t)f2_levels[l] & 15) || ((uintptr_t)f2_split[l] & 15)) return RAFTK_ERR_ALIGN;
  }
  {
    // which tensor-core kernel: the M = 128 / A-in-TMEM form (altcorr_tc128.cu) wins once its 16 x 8 patches give the
    // SMs enough to do (bands of a 136 x 240 image, M = 128 against M = 64: 136 rows 96 vs 114 us, 68 rows 59 vs 69,
    // 34 rows 41 vs 45, 17 rows 34 vs 31); below that the 8 x 8 patches of this file's kernel spread the work wider
    const int mode = altcorr_mode_flag();
    const long long patches128 = (long long)B * ((nq / w + 7) / 8) * ((w + 15) / 16);
    const bool fits128 = (long long)B * nq <= 0x7fffffffLL;  // that kernel keeps query indices in 32 bits
    if (fits128 && (mode == 2 || (mode == 0 && 5 * patches128 >= 2 * (long long)sm_count())))
      return altcorr_tc128_launch(fmap1, f2_levels, f1_split, f2_split, num_levels, coords, out, B, nq, h, w, D, radius, stream);
  }
  EncodeTiledFn encode = get_encode_fn();
  if (!encode) return RAFTK_ERR_UNSUPPORTED;
  Params p{};
  CUtensorMap tmA, tmB[4];
  memset(tmB, 0, sizeof(tmB));
  const int rows_q = nq / w;
  int rc = encode_4d(encode, &tmA, f1_split, 2 * D, w, rows_q, B, 8, 8);
  if (rc != RAFTK_OK) return rc;
  for (int l = 0; l < 4; ++l) {
    const int ll = l < num_levels ? l : num_levels - 1;  // unused slots repeat the last level
    if (!f2_levels[ll] || !f2_split[ll]) return RAFTK_ERR_NULL;
    if (((uintptr_t)f2_levels[ll] & 15) || ((uintptr_t)f2_split[ll] & 15)) return RAFTK_ERR_ALIGN;
    p.f2[l] = f2_levels[ll];
    p.lh[l] = h >> ll;
    p.lw[l] = w >> ll;
    // box rows per chunk: the typical box of an 8 x 8 patch is FP + (7 >> l) + 2 rows tall; take the chunk height in
    // {4, 5, 6} that pads it least (ties: the taller chunk)
    {
      const int typ = 2 * radius + 2 + (7 >> ll) + 2;
      int best = 6, waste = (typ + 5) / 6 * 6 - typ;
      for (int rr = 5; rr >= 4; --rr)
        if ((typ + rr - 1) / rr * rr - typ < waste) {
          waste = (typ + rr - 1) / rr * rr - typ;
          best = rr;
        }
      p.rows[l] = best;
    }
    rc = encode_4d(encode, &tmB[l], f2_split[ll], 2 * D, w >> ll, h >> ll, B, 8, p.rows[l]);
    if (rc != RAFTK_OK) return rc;
  }
  p.f1 = fmap1;
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
  const size_t smem = 1024 + (size_t)2 * (D / 64) * A_BLOCK_BYTES + (size_t)RING_BYTES +
                      (size_t)2 * TQ * FP * FP * sizeof(float) + (size_t)4 * 16 * 2 * ROWBUF_PITCH * sizeof(float) + 320 +
                      2 * sizeof(ItemInfo);
  if (smem > (size_t)SMEM_LIMIT) return RAFTK_ERR_UNSUPPORTED;
  const long long nblk = (long long)B * ((rows_q + 7) / 8) * ((w + 7) / 8);
  if (nblk > 0x7fffffffLL) return RAFTK_ERR_SHAPE;
  // Work items.  One CTA per patch for whole waves of patches (the A tile and the set-up once for all levels); the
  // patches of a last, partial wave — and all of a problem smaller than two waves (a row band of a sharded image) —
  // as one CTA per (patch, level), level 0 first, so that the tail fills the SMs with short items.
  {
    const long long S_ = sm_count();
    const long long full = nblk / S_, tail = nblk - full * S_;
    if (nblk <= 2 * S_) p.split_from = 0;
    else if (tail == 0 || 4 * tail >= 3 * S_) p.split_from = (int)nblk;
    else p.split_from = (int)(full * S_);
  }
  p.npatch = (int)nblk;
  const long long ncta = p.split_from + (nblk - p.split_from) * num_levels;
  if (ncta > 0x7fffffffLL) return RAFTK_ERR_SHAPE;
  auto kernel = radius == 4 ? altcorr_tc_kernel<4> : altcorr_tc_kernel<3>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  // persistent: one CTA per SM at most, each walking the work items blockIdx.x, blockIdx.x + gridDim.x, ...
  const long long grid = ncta < (long long)sm_count() ? ncta : (long long)sm_count();
  kernel<<<(unsigned)grid, NUM_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(tmA, tmB[0], tmB[1], tmB[2], tmB[3], p);
  return launch_status();
}
