// Kernel family 1, product path: all-pairs correlation volume as a tcgen05/TMEM GEMM fed by
// TMA, with the 4-level average-pool pyramid fused into the epilogue.
// Replaces CorrBlock.build_pyramid (model.py:418-446; volume 472-481, pooling 443-445).
//
// Numerics.  The reference contracts float32.  Each operand is split into two bf16 terms,
// x = hi + mid (+ residual <= 2^-17 |x|), and three tensor-core products are accumulated in
// fp32:  hi*hi + mid*hi + hi*mid  (the dropped mid*mid term is <= 2^-16 relative).  Relative
// error ~2e-5: 50x inside the 1e-3 the north-star allows, for 1.5x the tensor work of one
// TF32 pass (single-pass TF32 would sit at ~4e-4 rms, ~2e-3 worst case: outside the bar).
//
// Work decomposition.
//   M tile  = 128 query pixels of one image pair          (TMEM lanes)
//   N tile  = an 8 (y) x 32 (x) patch of target pixels    (256 TMEM columns), fetched by a 4-D
//             TMA box over fmap2 (channels, x, y, pair); out-of-image rows/columns arrive as
//             zeros and are never stored.  Every 2^l x 2^l pooling block of levels 1..3 lies
//             inside one patch, and one epilogue thread owns one query's whole patch, so the
//             pooling is register-local: no cross-thread traffic, no re-read of level 0.
//   K       = 2D bf16 per row ([hi | mid]); the 128-query A tile (up to 128 KB) stays resident
//             in shared memory while target patches stream through a TMA ring (A-stationary:
//             halves L2->smem operand traffic, which at K = 256 is the second bottleneck after
//             the HBM stores).
//   CTAs    persistent, one per SM, launched as clusters of two (tcgen05 cta_group::2): a pair
//           owns a 256-query tile; each CTA keeps its own 128 query rows of A and loads HALF of
//           every target patch, the leader CTA issues M = 256 MMAs that read both halves.  That
//           halves per-CTA operand traffic and doubles the k-block stages in flight (4 x 16 KB).
//           The linearised tile list is cut into equal contiguous ranges, one per pair.
//
// Warp roles (384 threads): warp 0 TMA producer, warp 1 MMA issuer (one elected lane each; leader
// CTA only), warp 2 TMEM allocator, warps 4-11 epilogue (TMEM lane quarter = warp % 4, patch rows
// 0-3 / 4-7).  Accumulators are double-buffered in TMEM (2 x 256 columns) so the epilogue of tile
// i overlaps the MMAs of tile i+1.  Each 32-query x 32-float row block is staged in a swizzled
// shared-memory tile and leaves as ONE bulk tensor (TMA) store, which also clips rows, columns and
// queries at the level's edges; levels whose pitch TMA cannot address (w>>l not a multiple of 4)
// are written with plain stores from the same staging tile.  1/sqrt(D) rides on fmap1 in the split
// pre-pass, so the epilogue has no multiply.
//
// Roofline (B = 1, 55x128, D = 256): 25.4 GFLOP algorithmic (76 GFLOP issued), 275.7 MB
// (261.3 MB of stores): HBM-store floor ~42 us, three-product MMA floor ~46 us of bf16 pipe.
// Measured (ncu, B200): 77 us, tensor pipe 59 % active; mainloop alone (epilogue stubbed) runs at
// 85 %, so the register-resident pooling epilogue is the limiter (profiles/).

#include <cuda.h>
#include <cstring>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "raftk_common.cuh"
#include "pyramid_tiled.cuh"

namespace raftk {

int check_pyramid_args(const float*, const float*, float* const*, int, int, int, int, int, int);
int launch_pool_levels(float* const*, int, int, long long, int, int, cudaStream_t);

namespace tc {

constexpr int BM = 128;           // queries per tile
constexpr int PW = 32, PH = 8;    // target patch
constexpr int BN = PW * PH;       // 256 accumulator columns
constexpr int KBLK = 64;          // bf16 per 128-byte swizzle row
constexpr int A_BLOCK_BYTES = BM * 128;  // 16 KB: 128 rows x 64 bf16
constexpr int B_STAGE_BYTES = BN * 128;  // 32 KB: 256 rows x 64 bf16
constexpr int NUM_EPI_WARPS = 8;         // two per TMEM lane quarter: patch rows 0-3 and 4-7
constexpr int STAGING_PER_WARP = 4096;   // one 32-query x 32-float row block, reused by every level
constexpr int MAX_STAGES = 6;
constexpr int NUM_THREADS = 128 + 32 * NUM_EPI_WARPS;
constexpr int SMEM_LIMIT = 232448;  // 227 KB

struct Params {
  void* level[4];  // float32 (row-major or tiled) or, OUT_TILED_F16, float16 tiled slabs
  int lh[4], lw[4];
  int vec_ok[4];  // level row pitch and base allow 16-byte stores
  int ltx[4], lty[4];   // tiled layout: tiles per tile row / tile rows of each level
  long long lslab[4];   // tiled layout: floats per query slab
  int L;          // levels written by this kernel (<= 4)
  int B, nq, h, w, D;
  int m_tiles, px_tiles, n_tiles;
  long long total_tiles;
  int num_stages;
};

// ---- PTX wrappers ------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(const CUtensorMap* tm, uint64_t* bar, void* dst, int c0, int c1,
                                            int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::
          "r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(const CUtensorMap* tm, uint64_t* bar, void* dst, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 lanes x 32 consecutive columns -> 32 registers per thread (thread = lane, register = column)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
// 32 lanes x 16 consecutive columns -> 16 registers per thread
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- cluster / CTA-pair (cta_group::2) variants ----------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_shared(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  // default .release.cta: ordering of the TMEM reads is carried by tcgen05.fence, and a cluster-scope
  // release would make the warp wait for all of its outstanding global stores
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA loads of a CTA pair: data lands in the issuing CTA, completion bytes go to `bar_cluster`
// (the leader CTA's barrier).
__device__ __forceinline__ void tma_load_3d_pair(const CUtensorMap* tm, uint32_t bar_cluster, void* dst, int c0,
                                                 int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, "
      "%5}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d_pair(const CUtensorMap* tm, uint32_t bar_cluster, void* dst, int c0,
                                                 int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, "
      "%5, %6}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
// commit of the pair's MMAs: one arrival on the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void tc_mma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                 uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// One lane of a converged warp.  tcgen05 / TMA instructions issued under `if (elect_one())` inside warp-uniform
// control flow keep their descriptors and addresses in uniform registers; under `if (lane == 0)` the compiler treats
// them as divergent and wraps every instruction in a vote-and-broadcast loop (~13 extra instructions per MMA).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(pred));
  return pred != 0;
}
// Four kind::f16 MMAs over one 64-element K block (k = 0, 16, 32, 48: +32 B = +2 in the descriptors' address field)
// from ONE asm statement; acc_first: whether the first of the four accumulates (the other three always do).
template <int CG>
__device__ __forceinline__ void tc_mma_bf16_x4(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t acc_first) {
  if (CG == 1) {
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        ".reg .b64 a1, b1, a2, b2, a3, b3;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "setp.eq.b32 q, 0, 0;\n\t"
        "add.s64 a1, %1, 2;\n\t"
        "add.s64 b1, %2, 2;\n\t"
        "add.s64 a2, %1, 4;\n\t"
        "add.s64 b2, %2, 4;\n\t"
        "add.s64 a3, %1, 6;\n\t"
        "add.s64 b3, %2, 6;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, q;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], a2, b2, %3, q;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], a3, b3, %3, q;\n\t"
        "}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(acc_first)
        : "memory");
  } else {
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        ".reg .b64 a1, b1, a2, b2, a3, b3;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "setp.eq.b32 q, 0, 0;\n\t"
        "add.s64 a1, %1, 2;\n\t"
        "add.s64 b1, %2, 2;\n\t"
        "add.s64 a2, %1, 4;\n\t"
        "add.s64 b2, %2, 4;\n\t"
        "add.s64 a3, %1, 6;\n\t"
        "add.s64 b3, %2, 6;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], a1, b1, %3, q;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], a2, b2, %3, q;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], a3, b3, %3, q;\n\t"
        "}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(acc_first)
        : "memory");
  }
}

// K-major operand tile, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart.
// start address >> 4 | LBO (unused for swizzled K-major, canonical 1) | SBO = 1024 B | version 1
// (Blackwell) | layout type 2 (SWIZZLE_128B).  Advancing K by 16 bf16 = +32 B on the start address.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
         (2ull << 61);
}
// kind::f16 instruction descriptor: D = f32, A = B = bf16, both K-major, N = 256, M = 128.
__host__ __device__ constexpr uint32_t make_idesc(int m) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// ---- split kernel: float32 rows -> [hi | mid] bf16 rows ---------------------------------
__global__ void __launch_bounds__(256)
split_bf16_kernel(const float* __restrict__ f1, const float* __restrict__ f2, __nv_bfloat16* __restrict__ a_split,
                  __nv_bfloat16* __restrict__ b_split, long long rows1, long long rows2, int D, float scale1) {
  const int d4 = D >> 2;
  const long long total = (rows1 + rows2) * d4;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    long long row = t / d4;
    const int c = (int)(t - row * d4) * 4;
    const float* src;
    __nv_bfloat16* dst;
    const bool first = row < rows1;
    if (first) {
      src = f1 + row * D + c;
      dst = a_split + row * 2 * D + c;
    } else {
      row -= rows1;
      src = f2 + row * D + c;
      dst = b_split + row * 2 * D + c;
    }
    const float4 x = __ldg(reinterpret_cast<const float4*>(src));
    const float sc = first ? scale1 : 1.0f;  // 1/sqrt(D) rides on fmap1, so the epilogue has no multiply
    const float xs[4] = {x.x * sc, x.y * sc, x.z * sc, x.w * sc};
    __nv_bfloat16 hi[4], mid[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      hi[i] = __float2bfloat16_rn(xs[i]);
      mid[i] = __float2bfloat16_rn(xs[i] - __bfloat162float(hi[i]));
    }
    *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(hi);
    *reinterpret_cast<uint2*>(dst + D) = *reinterpret_cast<const uint2*>(mid);
  }
}

// ---- epilogue helpers ---------------------------------------------------------------------
// Plain-store fallback for a staged row block (32 queries x WIDTH floats) whose level TMA cannot
// address: `row` points at (first query of the block, row Y, column X0); `slab` = floats per
// query slab; `cols` = level columns left from X0.  Lanes sweep the block row-major, so a warp
// store covers whole row segments.
template <int WIDTH, bool SWIZZLED>
__device__ __forceinline__ void store_row_block(const float* S, float* row, long long slab, int cols, int q_valid,
                                                int lane) {
  for (int idx = lane; idx < 32 * WIDTH; idx += 32) {
    const int qq = idx / WIDTH, x = idx % WIDTH;
    if (qq < q_valid && x < cols) {
      const int c = x >> 2;
      const int pc = SWIZZLED ? (c ^ (qq & 7)) : c;
      row[qq * slab + x] = S[qq * WIDTH + pc * 4 + (x & 3)];
    }
  }
}

// TMA (bulk tensor) store plumbing for the epilogue.
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* tm, const void* src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(tm)),
               "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Where one pyramid level goes.
struct LevelDst {
  const CUtensorMap* tm;  // (x, y, query, pair) tensor map, box (WIDTH, 1, 32, 1)
  float* base;
  int lh, lw;
  bool tma;     // pitch and base allow TMA / 16-byte stores
};

// Stage one row block (lane = query, WIDTH floats each) in the warp's shared-memory tile and send
// it to level memory as row Y, columns X0.., of queries [m_first, m_first + 32) of pair b.
// Aligned levels go out as ONE bulk tensor store per row block: the TMA engine clips columns,
// rows and queries past the level / past nq, and keeps many stores in flight per SM — the LSU
// path (a few hundred outstanding 16-byte stores per SM) cannot sustain the 27 B/clk/SM that the
// HBM write roofline asks of this epilogue.  Unaligned levels fall back to LDS + STG.
template <int WIDTH, bool SWIZZLED>
__device__ __forceinline__ void emit_row(float* S, const float* v, const LevelDst& d, int nq, int b, int m_first, int Y,
                                         int X0, int lane) {
  // the previous row block may still be being read out of S by the TMA engine
  if (lane == 0) bulk_wait_read_all();
  __syncwarp();
#pragma unroll
  for (int c = 0; c < WIDTH / 4; ++c) {
    const int pc = SWIZZLED ? (c ^ (lane & 7)) : c;
    *reinterpret_cast<float4*>(S + lane * WIDTH + pc * 4) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
  }
  const int q_valid = min(32, nq - m_first);
  if (d.tma) {
    fence_async_smem();  // make the generic-proxy writes visible to the async proxy
    __syncwarp();
    if (lane == 0 && Y < d.lh && q_valid > 0) {
      tma_store_4d(d.tm, S, X0, Y, m_first, b);
      bulk_commit();
    }
  } else {
    __syncwarp();
    if (Y < d.lh && q_valid > 0) {
      const long long slab = (long long)d.lh * d.lw;
      store_row_block<WIDTH, SWIZZLED>(S, d.base + ((long long)b * nq + m_first) * slab + (long long)Y * d.lw + X0, slab,
                                       d.lw - X0, q_valid, lane);
    }
    __syncwarp();
  }
}

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int threads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// out[i] = ((a[2i] + a[2i+1]) + (b[2i] + b[2i+1])) / 4 — the reference's pooling of the level below
template <int N>
__device__ __forceinline__ void pool_rows(const float* a, const float* b, float* out) {
#pragma unroll
  for (int i = 0; i < N; ++i)
    out[i] = __fadd_rn(__fadd_rn(a[2 * i], a[2 * i + 1]), __fadd_rn(b[2 * i], b[2 * i + 1])) * 0.25f;
}


// ---- tiled-layout epilogue (pyramid_tiled.cuh) ----------------------------------------------
__device__ __forceinline__ void tma_store_5d(const CUtensorMap* tm, const void* src, int c0, int c1, int c2, int c3, int c4) {
  asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(tm)),
               "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
               : "memory");
}

// Stage 32 floats per lane (lane = query: two adjacent 4x4 tiles = 128 contiguous bytes of its slab) in the warp's
// swizzled staging tile and send them as ONE bulk tensor store: tile pair `pair`, tile row `trow` of queries
// [m_first, m_first + 32) of pair b.  The tensor map (32 floats, query, tile pairs, tile rows, pair) clips tile
// pairs / tile rows / queries that do not exist.
__device__ __forceinline__ void emit_pair(float* S, const float* v, const CUtensorMap* tm, int b, int m_first, int pair,
                                          int trow, int lane) {
  if (lane == 0) bulk_wait_read_all();  // the previous store may still be reading S
  __syncwarp();
#pragma unroll
  for (int c = 0; c < 8; ++c)
    *reinterpret_cast<float4*>(S + lane * 32 + ((c ^ (lane & 7)) * 4)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
  fence_async_smem();
  __syncwarp();
  if (lane == 0) {
    tma_store_5d(tm, S, 0, m_first, pair, trow, b);
    bulk_commit();
  }
}

// eight float32 -> eight float16 (round to nearest even), 16 bytes
__device__ __forceinline__ uint4 pack8_f16(const float* v) {
  const __half2 a = __floats2half2_rn(v[0], v[1]), b = __floats2half2_rn(v[2], v[3]);
  const __half2 c = __floats2half2_rn(v[4], v[5]), d = __floats2half2_rn(v[6], v[7]);
  return make_uint4(*reinterpret_cast<const unsigned*>(&a), *reinterpret_cast<const unsigned*>(&b),
                    *reinterpret_cast<const unsigned*>(&c), *reinterpret_cast<const unsigned*>(&d));
}

// float16 slabs: stage 64 values per lane (lane = query: four adjacent 4x4 tiles = 128 contiguous bytes of its slab,
// v0 = tiles 0-1, v1 = tiles 2-3) in the warp's swizzled staging tile and send them as ONE bulk tensor store: tile
// pairs `pair`, `pair` + 1 of tile row `trow`, queries [m_first, m_first + 32) of pair b.  The tensor map
// (32 halves = tile pair, query, tile pairs, tile rows, pair; box {32, 32, 2, 1, 1}) clips what does not exist.
__device__ __forceinline__ void emit_quad_f16(float* S, const float* v0, const float* v1, const CUtensorMap* tm, int b,
                                              int m_first, int pair, int trow, int lane) {
  if (lane == 0) bulk_wait_read_all();  // the previous store may still be reading S
  __syncwarp();
  // dense, unswizzled (with a 64-byte inner box the 128-byte swizzle does not map to a per-query XOR); the staging
  // stores of 8 lanes share banks 4 ways: ~100 cycles per emit against the ~2400 of its wait
#pragma unroll
  for (int c = 0; c < 8; ++c)  // box order: 32 halves, query, tile pair -> [pair][query][64 B]
    *reinterpret_cast<uint4*>(S + (c >> 2) * 512 + lane * 16 + (c & 3) * 4) = pack8_f16((c < 4 ? v0 : v1) + 8 * (c & 3));
  fence_async_smem();
  __syncwarp();
  if (lane == 0) {
    tma_store_5d(tm, S, 0, m_first, pair, trow, b);
    bulk_commit();
  }
}

// 2x2 pooling of a 4x4 block stored row-major with pitch 4 -> its 2x2 result (pitch `op`), the reference's
// order ((a+b)+(c+d))/4
__device__ __forceinline__ void pool_tile(const float* t, float* out, int op) {
#pragma unroll
  for (int py = 0; py < 2; ++py)
#pragma unroll
    for (int px = 0; px < 2; ++px) {
      const float* s = t + (2 * py) * 4 + 2 * px;
      out[py * op + px] = __fadd_rn(__fadd_rn(s[0], s[1]), __fadd_rn(s[4], s[5])) * 0.25f;
    }
}

// CG = 1: one CTA per tile.  CG = 2: a CTA pair (cluster of 2, tcgen05 cta_group::2) per
// 256-query tile: each CTA keeps its own 128 query rows of A and loads HALF of every target
// patch (4 of its 8 rows); the leader CTA issues M = 256 MMAs that read both halves and write
// each CTA's 128 accumulator lanes.  Per-CTA operand traffic and shared memory per stage halve,
// which doubles the number of k-block stages in flight.
// Tiled outputs (pyramid_tiled.cuh): the accumulator keeps the patch's row-major column order (pixel (y, x) at column
// y*32 + x); an epilogue warp reads its 16 patch columns of four patch rows (four tcgen05.ld.x16 = four 4x4 tiles per
// query), and the tile order is a compile-time permutation of those registers.
// OUT_TILED_F16: the tiled slabs hold float16 (round-to-nearest-even of the float32 values; every level is pooled in
// float32 from the unrounded level below and rounded once).  Level 0 leaves as 32 queries x 128 B (four tiles) per
// bulk tensor store — two per accumulator and warp instead of four — and the small levels as plain 16/64-byte stores.
enum { OUT_ROWMAJOR = 0, OUT_TILED = 1, OUT_TILED_F16 = 2 };

template <int CG, int OUT>
__global__ void __launch_bounds__(NUM_THREADS, 1)
corr_pyramid_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ CUtensorMap tmO0, const __grid_constant__ CUtensorMap tmO1,
                       const __grid_constant__ CUtensorMap tmO2, const __grid_constant__ CUtensorMap tmO3,
                       const Params p) {
  constexpr bool TILED = OUT != OUT_ROWMAJOR;
  constexpr int kBStage = B_STAGE_BYTES / CG;  // bytes of one B stage held by this CTA
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
  const bool leader = rank == 0;
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the 128B-swizzle tiles; plain offset arithmetic keeps the pointers
  // in the shared address space (LDS/STS instead of generic LD/ST)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int kblocks = p.D / KBLK;            // k-blocks per plane
  const int a_blocks = 2 * kblocks;          // hi blocks then mid blocks
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + a_blocks * A_BLOCK_BYTES;
  uint8_t* smem_stage = smem_b + p.num_stages * kBStage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_stage + NUM_EPI_WARPS * STAGING_PER_WARP);
  uint64_t* b_full = bars;                    // [MAX_STAGES]
  uint64_t* b_empty = bars + MAX_STAGES;      // [MAX_STAGES]
  uint64_t* a_full = bars + 2 * MAX_STAGES;   // [1]
  uint64_t* a_empty = a_full + 1;             // [1]
  uint64_t* acc_full = a_empty + 1;           // [2]
  uint64_t* acc_empty = acc_full + 2;         // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < MAX_STAGES; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], 1);
    }
    mbar_init(a_full, 1);
    mbar_init(a_empty, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], NUM_EPI_WARPS * CG);  // one arrival per epilogue warp (of both CTAs)
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    if (CG == 1) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    } else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    }
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // contiguous, balanced range of the linearised tile list (pair, m tile, n tile); with CG = 2
  // both CTAs of a pair walk the same range
  // Work split: every image pair's tile list (m tile major, n tile minor) is cut into equal
  // contiguous ranges, one per CTA (pair).  All CTAs therefore work on the same image pair at the
  // same time, which keeps that pair's split operands (7 MB) resident in L2 at any batch size.
  const long long unit = blockIdx.x / CG, units = gridDim.x / CG;
  const long long tiles_per_pair = (long long)p.m_tiles * p.n_tiles;
  const long long t_begin = tiles_per_pair * unit / units;
  const long long t_end = tiles_per_pair * (unit + 1) / units;
  // barriers the pair shares live in the leader CTA
  const uint32_t a_full_lead = CG == 2 ? mapa_shared(smem_u32(a_full), 0) : 0u;

  if (warp == 0) {
    // ===================== TMA producer: the whole warp walks the loops, one elected lane issues =====================
    {
      int stage = 0;
      uint32_t phase = 0, a_phase = 0;
      long long cur_bm = -1;
      for (int pair_b = 0; pair_b < p.B; ++pair_b)
      for (long long t = t_begin; t < t_end; ++t) {
        const long long bm = (long long)pair_b * p.m_tiles + t / p.n_tiles;
        const int nt = (int)(t % p.n_tiles);
        const int b = (int)(bm / p.m_tiles);
        if (bm != cur_bm) {
          cur_bm = bm;
          const int m0 = (int)(bm - (long long)b * p.m_tiles) * (BM * CG) + (int)rank * BM;
          mbar_wait(a_empty, a_phase ^ 1);  // MMAs of the previous A tile have drained
          if (elect_one()) {
            if (leader) mbar_expect_tx(a_full, CG * a_blocks * A_BLOCK_BYTES);
            for (int j = 0; j < a_blocks; ++j) {
              if (CG == 1)
                tma_load_3d(&tmA, a_full, smem_a + j * A_BLOCK_BYTES, j * KBLK, m0, b);
              else
                tma_load_3d_pair(&tmA, a_full_lead, smem_a + j * A_BLOCK_BYTES, j * KBLK, m0, b);
            }
          }
          __syncwarp();
          a_phase ^= 1;
        }
        const int x0 = (nt % p.px_tiles) * PW, y0 = (nt / p.px_tiles) * PH;
        for (int j = 0; j < a_blocks; ++j) {
          // order: hi[0], mid[0], hi[1], mid[1], ...  (mid block of k-block kb lives at K = D + kb*64)
          const int kb = j >> 1;
          const int kcoord = (j & 1) ? p.D + kb * KBLK : kb * KBLK;
          mbar_wait(&b_empty[stage], phase ^ 1);
          if (elect_one()) {
            if (leader) mbar_expect_tx(&b_full[stage], B_STAGE_BYTES);  // both CTAs' halves
            if (CG == 1)
              tma_load_4d(&tmB, &b_full[stage], smem_b + stage * kBStage, kcoord, x0, y0, b);
            else  // this CTA's 4 of the patch's 8 rows
              tma_load_4d_pair(&tmB, mapa_shared(smem_u32(&b_full[stage]), 0), smem_b + stage * kBStage, kcoord, x0,
                               y0 + (int)rank * (PH / 2), b);
          }
          __syncwarp();
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only): the whole warp walks the loops, one elected lane issues ====
    if (leader) {
      constexpr uint32_t kIdesc = make_idesc(BM * CG);
      auto commit = [](uint64_t* bar) {
        if (CG == 1) tc_commit(bar); else tc_commit_pair(bar);
      };
      // descriptors differ from their base only in the address field (bytes >> 4; no carry out of it below 256 KB)
      const uint64_t a_desc0 = umma_desc(smem_u32(smem_a)), b_desc0 = umma_desc(smem_u32(smem_b));
      int stage = 0, buf = 0;
      uint32_t phase = 0, a_phase = 0, acc_phase = 0;
      long long cur_bm = -1;
      bool have_a = false;
      for (int pair_b = 0; pair_b < p.B; ++pair_b)
      for (long long t = t_begin; t < t_end; ++t) {
        const long long bm = (long long)pair_b * p.m_tiles + t / p.n_tiles;
        if (bm != cur_bm) {
          if (have_a) {
            if (elect_one()) commit(a_empty);  // everything issued so far has finished reading A
            __syncwarp();
          }
          cur_bm = bm;
          have_a = true;
          mbar_wait(a_full, a_phase);
          a_phase ^= 1;
          tc_fence_after();
        }
        mbar_wait(&acc_empty[buf], acc_phase ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)buf * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          const uint64_t a_hi = a_desc0 + (uint64_t)(kb * (A_BLOCK_BYTES >> 4));
          const uint64_t a_mid = a_desc0 + (uint64_t)((kblocks + kb) * (A_BLOCK_BYTES >> 4));
          // stage holding B_hi[kb]: hi*hi and mid*hi
          mbar_wait(&b_full[stage], phase);
          tc_fence_after();
          uint64_t bs = b_desc0 + (uint64_t)(stage * (kBStage >> 4));
          if (elect_one()) {
            tc_mma_bf16_x4<CG>(tmem_d, a_hi, bs, kIdesc, kb != 0);
            tc_mma_bf16_x4<CG>(tmem_d, a_mid, bs, kIdesc, 1);
            commit(&b_empty[stage]);
          }
          __syncwarp();
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1;
          }
          // stage holding B_mid[kb]: hi*mid
          mbar_wait(&b_full[stage], phase);
          tc_fence_after();
          bs = b_desc0 + (uint64_t)(stage * (kBStage >> 4));
          if (elect_one()) {
            tc_mma_bf16_x4<CG>(tmem_d, a_hi, bs, kIdesc, 1);
            commit(&b_empty[stage]);
          }
          __syncwarp();
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (elect_one()) commit(&acc_full[buf]);
        __syncwarp();
        if (++buf == 2) {
          buf = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    // Eight warps: quarter = TMEM lane quarter this warp may read (hardware: warp id % 4),
    // half = which four patch rows (two level-1 rows, one level-2 row) it owns.
    const int quarter = warp & 3;
    const int half = (warp - 4) >> 2;
    float* S = reinterpret_cast<float*>(smem_stage + (warp - 4) * STAGING_PER_WARP);
    // The level-2 row of the upper half (rows 0-3) travels to the lower-half warp through the
    // last 1 KB of the LOWER warp's staging tile, which that warp only touches while staging
    // level-0 rows.  bar_free: "lower warp is past its level-0 rows, write now";
    // bar_ready: "upper row written".
    float* X = reinterpret_cast<float*>(smem_stage + (4 + quarter) * STAGING_PER_WARP) + 768;
    const int bar_ready = 1 + quarter, bar_free = 5 + quarter;
    const LevelDst dst[4] = {{&tmO0, static_cast<float*>(p.level[0]), p.lh[0], p.lw[0], p.vec_ok[0] != 0},
                             {&tmO1, static_cast<float*>(p.level[1]), p.lh[1], p.lw[1], p.vec_ok[1] != 0},
                             {&tmO2, static_cast<float*>(p.level[2]), p.lh[2], p.lw[2], p.vec_ok[2] != 0},
                             {&tmO3, static_cast<float*>(p.level[3]), p.lh[3], p.lw[3], p.vec_ok[3] != 0}};
    int buf = 0;
    uint32_t acc_phase = 0;
    for (int pair_b = 0; pair_b < p.B; ++pair_b)
    for (long long t = t_begin; t < t_end; ++t) {
      const long long bm = (long long)pair_b * p.m_tiles + t / p.n_tiles;
      const int nt = (int)(t % p.n_tiles);
      const int b = (int)(bm / p.m_tiles);
      const int m0 = (int)(bm - (long long)b * p.m_tiles) * (BM * CG) + (int)rank * BM + quarter * 32;  // first query of this warp
      const int x0 = (nt % p.px_tiles) * PW, y0 = (nt / p.px_tiles) * PH;

      mbar_wait(&acc_full[buf], acc_phase);
      tc_fence_after();
      if (OUT == OUT_TILED_F16) {
        // half = which four tile columns (16 target columns) of the patch this warp owns, both tile rows
        const uint32_t tcol = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)buf * BN + (uint32_t)(half * 16);
        const bool last_row = y0 + PH >= p.h;
        const int q = m0 + lane;
        float l1t[2][16];  // this warp's two level-1 tiles, [u][cy * 4 + cx]
#pragma unroll
        for (int ty = 0; ty < 2; ++ty) {
          float a[4][16];  // patch rows 4 ty + r, this warp's 16 columns
#pragma unroll
          for (int r = 0; r < 4; ++r) tmem_ld_32x16(tcol + (uint32_t)((4 * ty + r) * PW), a[r]);
          tmem_ld_wait();
          if (ty == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (CG == 1) mbar_arrive(&acc_empty[buf]);
              else mbar_arrive_cluster(mapa_shared(smem_u32(&acc_empty[buf]), 0));
            }
          }
          float v[2][32];  // tiles 0-1 and 2-3 of this warp's four, in slab order
#pragma unroll
          for (int u = 0; u < 2; ++u)
#pragma unroll
            for (int k = 0; k < 32; ++k) v[u][k] = a[(k >> 2) & 3][8 * u + 4 * (k >> 4) + (k & 3)];
          const int pair0 = (x0 >> 3) + 2 * half, trow = (y0 >> 2) + ty;
          if (pair0 * 2 < p.ltx[0] && trow < p.lty[0]) emit_quad_f16(S, v[0], v[1], &tmO0, b, m0, pair0, trow, lane);
          if (p.L > 1) {
#pragma unroll
            for (int u = 0; u < 2; ++u)
#pragma unroll
              for (int t = 0; t < 2; ++t) pool_tile(v[u] + t * 16, l1t[u] + (2 * ty) * 4 + 2 * t, 4);
          }
        }
        if (p.L > 1) {
          // level 1: rows (y0>>1) + 0..3, columns (x0>>1) + 8*half + 4u + 0..3 = tiles 2*half + u of tile row y0>>3:
          // 64 contiguous bytes per query; cells past the level are zero
          const int ry1 = p.lh[1] - (y0 >> 1), rx1 = p.lw[1] - ((x0 >> 1) + 8 * half);
          if (ry1 < 4 || rx1 < 8) {
#pragma unroll
            for (int u = 0; u < 2; ++u)
#pragma unroll
              for (int cy = 0; cy < 4; ++cy)
#pragma unroll
                for (int cx = 0; cx < 4; ++cx)
                  if (cy >= ry1 || 4 * u + cx >= rx1) l1t[u][cy * 4 + cx] = 0.0f;
          }
          if (q < p.nq && rx1 > 0 && ry1 > 0) {
            __half* t1 = static_cast<__half*>(p.level[1]) + ((long long)b * p.nq + q) * p.lslab[1] +
                         ((long long)(y0 >> 3) * p.ltx[1] + (x0 >> 3) + 2 * half) * 16;
#pragma unroll
            for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(t1 + 8 * c) = pack8_f16(&l1t[0][0] + 8 * c);
          }
        }
        if (p.L > 2) {
          // level 2: rows (y0>>2) + r, columns (x0>>2) + 4*half + c -> rows rt2, rt2 + 1 of one tile: 16 contiguous bytes
          float l2[2][4];
#pragma unroll
          for (int u = 0; u < 2; ++u) pool_tile(l1t[u], &l2[0][2 * u], 4);
          const int ry2 = p.lh[2] - (y0 >> 2), rx2 = p.lw[2] - ((x0 >> 2) + 4 * half);
#pragma unroll
          for (int r = 0; r < 2; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c)
              if (r >= ry2 || c >= rx2) l2[r][c] = 0.0f;
          const int tcol2 = (x0 >> 4) + half, trow2 = y0 >> 4, rt2 = (y0 >> 2) & 3;
          if (q < p.nq && rx2 > 0 && trow2 * 4 < p.lh[2]) {  // the tile holds valid cells
            __half* t2 = static_cast<__half*>(p.level[2]) + ((long long)b * p.nq + q) * p.lslab[2] +
                         ((long long)trow2 * p.ltx[2] + tcol2) * 16;
            *reinterpret_cast<uint4*>(t2 + rt2 * 4) = pack8_f16(&l2[0][0]);
            if (last_row && rt2 == 0) *reinterpret_cast<uint4*>(t2 + 8) = make_uint4(0u, 0u, 0u, 0u);  // padding rows
          }
          if (p.L > 3) {
            // level 3: row y0>>3, columns (x0>>3) + 2*half + c -> 4 bytes of a tile row
            const int ry3 = p.lh[3] - (y0 >> 3), rx3 = p.lw[3] - ((x0 >> 3) + 2 * half);
            float l3[2];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              l3[c] = __fadd_rn(__fadd_rn(l2[0][2 * c], l2[0][2 * c + 1]), __fadd_rn(l2[1][2 * c], l2[1][2 * c + 1])) * 0.25f;
              if (ry3 < 1 || c >= rx3) l3[c] = 0.0f;
            }
            const int tcol3 = x0 >> 5, trow3 = y0 >> 5, rt3 = (y0 >> 3) & 3;
            if (q < p.nq && tcol3 * 4 < p.lw[3] && trow3 * 4 < p.lh[3]) {  // the tile holds valid cells
              __half* t3 = static_cast<__half*>(p.level[3]) + ((long long)b * p.nq + q) * p.lslab[3] +
                           ((long long)trow3 * p.ltx[3] + tcol3) * 16 + 2 * half;
              *reinterpret_cast<__half2*>(t3 + rt3 * 4) = __floats2half2_rn(l3[0], l3[1]);
              if (last_row)  // no patch below: the rest of the tile is padding
                for (int r = rt3 + 1; r < 4; ++r) *reinterpret_cast<__half2*>(t3 + r * 4) = __floats2half2_rn(0.0f, 0.0f);
            }
          }
        }
        if (++buf == 2) {
          buf = 0;
          acc_phase ^= 1;
        }
        continue;
      }
      if (TILED) {
        // half = which 16 target columns (four tile columns) of the patch this warp owns, both tile rows
        const uint32_t tcol = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)buf * BN + (uint32_t)(half * 16);
        const bool last_row = y0 + PH >= p.h;
        float l1t[2][16];  // this warp's two level-1 tiles, [u][cy * 4 + cx]
#pragma unroll
        for (int ty = 0; ty < 2; ++ty) {
          float a[4][16];  // patch rows 4 ty + r, this warp's 16 columns
#pragma unroll
          for (int r = 0; r < 4; ++r) tmem_ld_32x16(tcol + (uint32_t)((4 * ty + r) * PW), a[r]);
          tmem_ld_wait();
          if (ty == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (CG == 1) mbar_arrive(&acc_empty[buf]);
              else mbar_arrive_cluster(mapa_shared(smem_u32(&acc_empty[buf]), 0));
            }
          }
          const int trow = (y0 >> 2) + ty;
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            float v[32];  // tiles 2u, 2u + 1 of this warp's four, in slab order
#pragma unroll
            for (int k = 0; k < 32; ++k) v[k] = a[(k >> 2) & 3][8 * u + 4 * (k >> 4) + (k & 3)];
            const int pair0 = (x0 >> 3) + 2 * half + u;
            if (pair0 * 2 < p.ltx[0] && trow < p.lty[0]) emit_pair(S, v, &tmO0, b, m0, pair0, trow, lane);
            if (p.L > 1) {
#pragma unroll
              for (int t = 0; t < 2; ++t) pool_tile(v + t * 16, l1t[u] + (2 * ty) * 4 + 2 * t, 4);
            }
          }
        }
        if (p.L > 1) {
          // level 1: rows (y0>>1) + 0..3, columns (x0>>1) + 8*half + 4u + 0..3; cells past the level are zero
          const int ry1 = p.lh[1] - (y0 >> 1), rx1 = p.lw[1] - ((x0 >> 1) + 8 * half);
          if (ry1 < 4 || rx1 < 8) {
#pragma unroll
            for (int u = 0; u < 2; ++u)
#pragma unroll
              for (int cy = 0; cy < 4; ++cy)
#pragma unroll
                for (int cx = 0; cx < 4; ++cx)
                  if (cy >= ry1 || 4 * u + cx >= rx1) l1t[u][cy * 4 + cx] = 0.0f;
          }
          if (rx1 > 0 && ry1 > 0) emit_pair(S, &l1t[0][0], &tmO1, b, m0, (x0 >> 4) + half, y0 >> 3, lane);
        }
        if (p.L > 2) {
          // level 2: rows (y0>>2) + r, columns (x0>>2) + 4*half + c -> half a tile (2 rows x 4 columns)
          float l2[2][4];
#pragma unroll
          for (int u = 0; u < 2; ++u) pool_tile(l1t[u], &l2[0][2 * u], 4);
          const int ry2 = p.lh[2] - (y0 >> 2), rx2 = p.lw[2] - ((x0 >> 2) + 4 * half);
#pragma unroll
          for (int r = 0; r < 2; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c)
              if (r >= ry2 || c >= rx2) l2[r][c] = 0.0f;
          const int q = m0 + lane;
          const int tcol2 = (x0 >> 4) + half, trow2 = y0 >> 4, rt2 = (y0 >> 2) & 3;
          if (q < p.nq && rx2 > 0 && trow2 * 4 < p.lh[2]) {  // the tile holds valid cells
            float* t2 = static_cast<float*>(p.level[2]) + ((long long)b * p.nq + q) * p.lslab[2] + ((long long)trow2 * p.ltx[2] + tcol2) * 16;
            *reinterpret_cast<float4*>(t2 + rt2 * 4) = make_float4(l2[0][0], l2[0][1], l2[0][2], l2[0][3]);
            *reinterpret_cast<float4*>(t2 + rt2 * 4 + 4) = make_float4(l2[1][0], l2[1][1], l2[1][2], l2[1][3]);
            if (last_row && rt2 == 0) {  // no patch below: the tile's lower half is padding
              *reinterpret_cast<float4*>(t2 + 8) = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
              *reinterpret_cast<float4*>(t2 + 12) = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            }
          }
          if (p.L > 3) {
            // level 3: row y0>>3, columns (x0>>3) + 2*half + c -> 8 bytes of a tile row
            const int ry3 = p.lh[3] - (y0 >> 3), rx3 = p.lw[3] - ((x0 >> 3) + 2 * half);
            float l3[2];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              l3[c] = __fadd_rn(__fadd_rn(l2[0][2 * c], l2[0][2 * c + 1]), __fadd_rn(l2[1][2 * c], l2[1][2 * c + 1])) * 0.25f;
              if (ry3 < 1 || c >= rx3) l3[c] = 0.0f;
            }
            const int tcol3 = x0 >> 5, trow3 = y0 >> 5, rt3 = (y0 >> 3) & 3;
            if (q < p.nq && tcol3 * 4 < p.lw[3] && trow3 * 4 < p.lh[3]) {  // the tile holds valid cells
              float* t3 = static_cast<float*>(p.level[3]) + ((long long)b * p.nq + q) * p.lslab[3] + ((long long)trow3 * p.ltx[3] + tcol3) * 16 +
                          2 * half;
              *reinterpret_cast<float2*>(t3 + rt3 * 4) = make_float2(l3[0], l3[1]);
              if (last_row)  // no patch below: the rest of the tile is padding
                for (int r = rt3 + 1; r < 4; ++r) *reinterpret_cast<float2*>(t3 + r * 4) = make_float2(0.0f, 0.0f);
            }
          }
        }
        if (++buf == 2) {
          buf = 0;
          acc_phase ^= 1;
        }
        continue;
      }
      const uint32_t taddr =
          tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)buf * BN + (uint32_t)(half * 4 * PW);

      float l1[2][16];
#pragma unroll
      for (int rp = 0; rp < 2; ++rp) {
        float r[2][32];
        tmem_ld_32x32(taddr + (2 * rp) * PW, r[0]);
        tmem_ld_32x32(taddr + (2 * rp + 1) * PW, r[1]);
        tmem_ld_wait();
        if (rp == 1) {
          // this warp's share of the accumulator is in registers: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (CG == 1) mbar_arrive(&acc_empty[buf]);
            else mbar_arrive_cluster(mapa_shared(smem_u32(&acc_empty[buf]), 0));  // the leader's barrier
          }
        }
        // ---- level 0: two rows ----
        emit_row<32, true>(S, r[0], dst[0], p.nq, b, m0, y0 + 4 * half + 2 * rp, x0, lane);
        emit_row<32, true>(S, r[1], dst[0], p.nq, b, m0, y0 + 4 * half + 2 * rp + 1, x0, lane);
        // ---- level 1 row ----
        if (p.L > 1) {
          pool_rows<16>(r[0], r[1], l1[rp]);
          emit_row<16, false>(S, l1[rp], dst[1], p.nq, b, m0, (y0 >> 1) + 2 * half + rp, x0 >> 1, lane);
        }
      }
      if (p.L > 2) {
        // ---- level 2 row ----
        float l2[8];
        pool_rows<8>(l1[0], l1[1], l2);
        emit_row<8, false>(S, l2, dst[2], p.nq, b, m0, (y0 >> 2) + half, x0 >> 2, lane);
        if (p.L > 3) {
          // ---- level 3 row: needs the level-2 rows of both halves ----
          if (half == 0) {
            named_bar_sync(bar_free, 64);
            *reinterpret_cast<float4*>(X + lane * 8) = make_float4(l2[0], l2[1], l2[2], l2[3]);
            *reinterpret_cast<float4*>(X + lane * 8 + 4) = make_float4(l2[4], l2[5], l2[6], l2[7]);
            named_bar_arrive(bar_ready, 64);
          } else {
            named_bar_arrive(bar_free, 64);
            named_bar_sync(bar_ready, 64);
            float up[8], l3[4];
            const float4 u0 = *reinterpret_cast<const float4*>(X + lane * 8);
            const float4 u1 = *reinterpret_cast<const float4*>(X + lane * 8 + 4);
            up[0] = u0.x; up[1] = u0.y; up[2] = u0.z; up[3] = u0.w;
            up[4] = u1.x; up[5] = u1.y; up[6] = u1.z; up[7] = u1.w;
            __syncwarp();
            pool_rows<4>(up, l2, l3);
            emit_row<4, false>(S, l3, dst[3], p.nq, b, m0, y0 >> 3, x0 >> 3, lane);
          }
        }
      }
      if (++buf == 2) {
        buf = 0;
        acc_phase ^= 1;
      }
    }
    if (lane == 0) bulk_wait_all();  // every bulk store has left shared memory and is globally performed
  }

  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();  // the peer may still signal our barriers until here
  if (warp == 2) {
    tc_fence_after();
    if (CG == 1)
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    else
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

// ---- host side ----------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

static size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

static bool supported(int D) { return D % KBLK == 0 && D >= KBLK && D <= 256; }

static size_t scratch_bytes(int B, int nq, int h, int w, int D) {
  if (!supported(D)) return 0;
  return align256((size_t)B * nq * 2 * D * 2) + align256((size_t)B * h * w * 2 * D * 2);
}

}  // namespace tc
}  // namespace raftk

extern "C" size_t raftk_corr_pyramid_scratch_bytes(int B, int nq, int h, int w, int D) {
  if (B <= 0 || nq <= 0 || h <= 0 || w <= 0 || D <= 0) return 0;
  return raftk::tc::scratch_bytes(B, nq, h, w, D);
}

static int corr_pyramid_impl(const float* fmap1, const float* fmap2, void* const* levels, int num_levels, int B, int nq,
                             int h, int w, int D, void* scratch, size_t scratch_bytes, void* stream, int out_mode) {
  using namespace raftk;
  using namespace raftk::tc;
  const bool tiled = out_mode != OUT_ROWMAJOR;
  int rc = check_pyramid_args(fmap1, fmap2, reinterpret_cast<float* const*>(levels), num_levels, B, nq, h, w, D);
  if (rc != RAFTK_OK) return rc;
  if (B == 0 || nq == 0) return RAFTK_OK;
  if (tiled) {
    // the tiled epilogue covers what the 8x32 patch reaches (4 levels) on the tensor-core tiling; other
    // configurations: build row-major levels and convert (raftk_pyramid_tile)
    if (!supported(D) || num_levels > 4) return RAFTK_ERR_UNSUPPORTED;
    for (int l = 0; l < num_levels; ++l)
      if ((uintptr_t)levels[l] & 127) return RAFTK_ERR_ALIGN;
  }
  // Feature widths the tensor-core tiling does not cover (D not a multiple of 64, or an A tile
  // larger than shared memory) take the fp32 FMA kernels: same contract, no scratch.
  if (!supported(D))
    return raftk_corr_pyramid_simt(fmap1, fmap2, reinterpret_cast<float* const*>(levels), num_levels, B, nq, h, w, D, stream);
  const size_t need = tc::scratch_bytes(B, nq, h, w, D);
  if (!scratch || scratch_bytes < need) return RAFTK_ERR_SCRATCH;
  if (((uintptr_t)scratch & 255) || ((uintptr_t)fmap1 & 15) || ((uintptr_t)fmap2 & 15)) return RAFTK_ERR_ALIGN;
  EncodeTiledFn encode = get_encode_fn();
  if (!encode) return RAFTK_ERR_UNSUPPORTED;
  cudaStream_t s = static_cast<cudaStream_t>(stream);

  __nv_bfloat16* a_split = reinterpret_cast<__nv_bfloat16*>(scratch);
  __nv_bfloat16* b_split =
      reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<uint8_t*>(scratch) + align256((size_t)B * nq * 2 * D * 2));
  const long long rows1 = (long long)B * nq, rows2 = (long long)B * h * w;
  {
    const long long total = (rows1 + rows2) * (D / 4);
    long long blocks = (total + 255) / 256;
    if (blocks > (long long)RAFTK_SM_COUNT * 8) blocks = (long long)RAFTK_SM_COUNT * 8;
    split_bf16_kernel<<<(unsigned)blocks, 256, 0, s>>>(fmap1, fmap2, a_split, b_split, rows1, rows2, D,
                                                      1.0f / (float)sqrt((double)D));
    rc = launch_status();
    if (rc != RAFTK_OK) return rc;
  }

  // CTA pairs whenever a pair's 256-query tile is not mostly padding
  const int CG = nq > BM ? 2 : 1;
  CUtensorMap tmA, tmB;
  {
    const cuuint64_t dims[3] = {(cuuint64_t)(2 * D), (cuuint64_t)nq, (cuuint64_t)B};
    const cuuint64_t strides[2] = {(cuuint64_t)(2 * D) * 2, (cuuint64_t)nq * 2 * D * 2};
    const cuuint32_t box[3] = {KBLK, BM, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    if (encode(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, a_split, dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return RAFTK_ERR_UNSUPPORTED;
  }
  {
    const cuuint64_t dims[4] = {(cuuint64_t)(2 * D), (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)B};
    const cuuint64_t strides[3] = {(cuuint64_t)(2 * D) * 2, (cuuint64_t)w * 2 * D * 2, (cuuint64_t)h * w * 2 * D * 2};
    // a CTA (pair) loads (half) a patch per stage
    const cuuint32_t box[4] = {KBLK, (cuuint32_t)PW, (cuuint32_t)(PH / CG), 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    if (encode(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, b_split, dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return RAFTK_ERR_UNSUPPORTED;
  }

  Params p{};
  p.L = num_levels < 4 ? num_levels : 4;
  for (int l = 0; l < p.L; ++l) {
    p.level[l] = levels[l];
    p.lh[l] = h >> l;
    p.lw[l] = w >> l;
    p.vec_ok[l] = ((w >> l) % 4 == 0) && (((uintptr_t)levels[l] & 15) == 0);
  }
  // output tensor maps (x, y, query, pair) for the levels whose pitch and base TMA accepts; the
  // others are written with plain stores
  CUtensorMap tmO[4];
  memset(tmO, 0, sizeof(tmO));
  for (int l = 0; l < p.L; ++l) {
    p.ltx[l] = tiled_tiles_x(w >> l);
    p.lty[l] = tiled_tiles_y(h >> l);
    p.lslab[l] = tiled_slab_floats(h >> l, w >> l);
  }
  if (out_mode == OUT_TILED_F16) {
    // level 0 leaves through bulk tensor stores: (32 halves = tile pair, pairs per tile row, tile rows, query, pair),
    // two pairs (128 B per query) per store
    const cuuint64_t dims[5] = {32, (cuuint64_t)nq, (cuuint64_t)p.ltx[0] / 2, (cuuint64_t)p.lty[0], (cuuint64_t)B};
    const cuuint64_t strides[4] = {(cuuint64_t)p.lslab[0] * 2, 64, (cuuint64_t)p.ltx[0] * 32, (cuuint64_t)nq * p.lslab[0] * 2};
    const cuuint32_t box[5] = {32, 32, 2, 1, 1};
    const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    if (encode(&tmO[0], CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 5, levels[0], dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return RAFTK_ERR_UNSUPPORTED;
  } else if (tiled) {
    // levels 0 and 1 leave through bulk tensor stores: (32 floats = tile pair, pairs per tile row, tile rows, query, pair)
    // (the query is the SECOND dimension: profiles/layout_bench.cu measures this order 20 % faster per store than
    // query-last with one staging tile per warp)
    for (int l = 0; l < p.L && l < 2; ++l) {
      const cuuint64_t dims[5] = {32, (cuuint64_t)nq, (cuuint64_t)p.ltx[l] / 2, (cuuint64_t)p.lty[l], (cuuint64_t)B};
      const cuuint64_t strides[4] = {(cuuint64_t)p.lslab[l] * 4, 128, (cuuint64_t)p.ltx[l] * 64,
                                     (cuuint64_t)nq * p.lslab[l] * 4};
      const cuuint32_t box[5] = {32, 32, 1, 1, 1};
      const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
      if (encode(&tmO[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 5, levels[l], dims, strides, box, estr,
                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return RAFTK_ERR_UNSUPPORTED;
    }
  }
  for (int l = 0; l < p.L && !tiled; ++l) {
    if (!p.vec_ok[l]) continue;
    const cuuint64_t lw = (cuuint64_t)(w >> l), lh = (cuuint64_t)(h >> l);
    const cuuint64_t dims[4] = {lw, lh, (cuuint64_t)nq, (cuuint64_t)B};
    const cuuint64_t strides[3] = {lw * 4, lh * lw * 4, (cuuint64_t)nq * lh * lw * 4};
    const cuuint32_t box[4] = {(cuuint32_t)(PW >> l), 1, 32, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    if (encode(&tmO[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, levels[l], dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, l == 0 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
               CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      p.vec_ok[l] = 0;
  }
  p.B = B;
  p.nq = nq;
  p.h = h;
  p.w = w;
  p.D = D;
  p.m_tiles = (nq + BM * CG - 1) / (BM * CG);
  p.px_tiles = (w + PW - 1) / PW;
  p.n_tiles = p.px_tiles * ((h + PH - 1) / PH);
  p.total_tiles = (long long)B * p.m_tiles * p.n_tiles;
  const int a_bytes = 2 * (D / KBLK) * A_BLOCK_BYTES;
  const int fixed = 1024 + a_bytes + NUM_EPI_WARPS * STAGING_PER_WARP + 256;
  const int b_stage = B_STAGE_BYTES / CG;
  int stages = (SMEM_LIMIT - fixed) / b_stage;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 2) return RAFTK_ERR_UNSUPPORTED;
  p.num_stages = stages;
  const int smem_bytes = fixed + stages * b_stage;
  const long long max_units = RAFTK_SM_COUNT / CG;
  const long long tiles_per_pair = (long long)p.m_tiles * p.n_tiles;
  const int units = (int)(tiles_per_pair < max_units ? tiles_per_pair : max_units);
  auto kernel = out_mode == OUT_TILED_F16
                    ? (CG == 2 ? corr_pyramid_tc_kernel<2, OUT_TILED_F16> : corr_pyramid_tc_kernel<1, OUT_TILED_F16>)
                    : tiled ? (CG == 2 ? corr_pyramid_tc_kernel<2, OUT_TILED> : corr_pyramid_tc_kernel<1, OUT_TILED>)
                            : (CG == 2 ? corr_pyramid_tc_kernel<2, OUT_ROWMAJOR> : corr_pyramid_tc_kernel<1, OUT_ROWMAJOR>);
  // per-device attribute; cheap enough to set on every call (keeps the library stateless)
  if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT) != cudaSuccess)
    return launch_status();
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(units * CG);
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (cudaLaunchKernelEx(&cfg, kernel, tmA, tmB, tmO[0], tmO[1], tmO[2], tmO[3], p) != cudaSuccess)
    return launch_status();
  rc = launch_status();
  if (rc != RAFTK_OK) return rc;
  // levels beyond the 8x32 patch's reach (only if a caller asks for more than 4)
  if (num_levels > 4)
    return launch_pool_levels(reinterpret_cast<float* const*>(levels), 4, num_levels, (long long)B * nq, h, w, s);
  return RAFTK_OK;
}

extern "C" int raftk_corr_pyramid(const float* fmap1, const float* fmap2, float* const* levels, int num_levels,
                                  int B, int nq, int h, int w, int D, void* scratch, size_t scratch_bytes,
                                  void* stream) {
  return corr_pyramid_impl(fmap1, fmap2, reinterpret_cast<void* const*>(levels), num_levels, B, nq, h, w, D, scratch,
                           scratch_bytes, stream, raftk::tc::OUT_ROWMAJOR);
}

extern "C" int raftk_corr_pyramid_tiled(const float* fmap1, const float* fmap2, float* const* levels, int num_levels,
                                        int B, int nq, int h, int w, int D, void* scratch, size_t scratch_bytes,
                                        void* stream) {
  return corr_pyramid_impl(fmap1, fmap2, reinterpret_cast<void* const*>(levels), num_levels, B, nq, h, w, D, scratch,
                           scratch_bytes, stream, raftk::tc::OUT_TILED);
}

extern "C" int raftk_corr_pyramid_tiled_f16(const float* fmap1, const float* fmap2, void* const* levels, int num_levels,
                                            int B, int nq, int h, int w, int D, void* scratch, size_t scratch_bytes,
                                            void* stream) {
  return corr_pyramid_impl(fmap1, fmap2, levels, num_levels, B, nq, h, w, D, scratch, scratch_bytes, stream,
                           raftk::tc::OUT_TILED_F16);
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of raftk_corr_pyramid (no counterpart in the reference).
//
// raftk_corr_pyramid needs a caller-provided scratch buffer for the bf16 hi/mid split of the two
// feature maps.  Some hosts of this library have no way to provide one at the point of the call:
// a custom-call handler running inside an XLA program that was compiled without a scratch
// operand for it, or a process whose framework pool (XLA's preallocating allocator, torch's
// caching allocator) already owns the device memory, so that a raw cudaMalloc from here would
// stall the device or simply fail.  For them the scratch lives in the library image itself: a
// fixed array in device memory that every GPU loading the library carries until the process
// exits.  Its size bounds the problem: RAFTK_STATIC_SCRATCH_BYTES (16 MiB) holds one 440x1024
// pair at D = 256 (14.4 MB); anything larger is refused with RAFTK_ERR_STATIC_LIMIT.
//
// Host code cannot use the array's name as a pointer — on the host that symbol is only a handle
// — so the device address is looked up with cudaGetSymbolAddress once per device, with that
// device current, the first time the entry point runs there, and kept.  That first call can load
// the module; make it before any stream capture begins.
//
// Every call on a device uses the same array and leaves its split operands behind.  Nothing in it
// has to start at zero: the split kernel of a call rewrites every byte that call's GEMM reads
// (rows past the end of a tile are zero-filled by TMA, never fetched), so stale contents are
// harmless.  Two calls that overlap on one GPU (different streams) would overwrite each other's
// operands: callers must order them.
// ---------------------------------------------------------------------------------------------
#include <mutex>

static __device__ __align__(256) unsigned char g_raftk_static_scratch[RAFTK_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
void* g_static_scratch_addr[kMaxDevices] = {};
std::mutex g_static_scratch_mutex;

// Device address of the static scratch on the current device (looked up once per device).
int static_scratch_address(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return static_cast<int>(e);
  if (dev < 0 || dev >= kMaxDevices) return RAFTK_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
  if (!g_static_scratch_addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_raftk_static_scratch);
    if (e != cudaSuccess) return static_cast<int>(e);
    g_static_scratch_addr[dev] = p;
  }
  *out = g_static_scratch_addr[dev];
  return RAFTK_OK;
}
}  // namespace

extern "C" int raftk_corr_pyramid_static(const float* fmap1, const float* fmap2, float* const* levels,
                                         int num_levels, int B, int nq, int h, int w, int D, void* stream) {
  using namespace raftk;
  int rc = check_pyramid_args(fmap1, fmap2, levels, num_levels, B, nq, h, w, D);
  if (rc != RAFTK_OK) return rc;
  if (B == 0 || nq == 0) return RAFTK_OK;
  const size_t need = raftk_corr_pyramid_scratch_bytes(B, nq, h, w, D);
  if (need > (size_t)RAFTK_STATIC_SCRATCH_BYTES) return RAFTK_ERR_STATIC_LIMIT;
  void* scratch = nullptr;
  rc = static_scratch_address(&scratch);
  if (rc != RAFTK_OK) return rc;
  return raftk_corr_pyramid(fmap1, fmap2, levels, num_levels, B, nq, h, w, D, scratch,
                            (size_t)RAFTK_STATIC_SCRATCH_BYTES, stream);
}
