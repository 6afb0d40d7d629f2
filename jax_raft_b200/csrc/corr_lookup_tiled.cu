// Kernel family 2, product path of round 2: radius-r bilinear window lookup over the TILED correlation
// pyramid (pyramid_tiled.cuh).  Replaces CorrBlock.index_pyramid (model.py:448-470) + grid_sample
// (model.py:24-34); same results, bit for bit, as corr_lookup.cu on the row-major pyramid.
//
// Why a second kernel.  ncu on the round-1 kernel at batch 16: DRAM 83 % busy on 1.77x the algorithmic
// bytes, issue slots 73 % busy (717 instructions per query), L1/shared pipe 73 % busy (190 wavefronts per
// query, two-way bank conflicts included).  All three had to come down together:
//
//   bytes         4 x 4-float tiles = one 64-byte L2 fetch granule; a footprint is fetched as 32-byte half tiles
//                 (LDG.256: rows 2k, 2k+1 of a tile), 3 loads per lane per query instead of 5 LDG.128, and
//                 only the half tiles that hold footprint rows — 2557 B instead of 4073 B of DRAM reads
//                 per query (profiles/layout_bench.cu).
//   instructions  COLUMN TASKS.  A lane owns one column of taps (query, level, x offset i) and walks its
//                 2r+1 y offsets: the two footprint values of row j+1 are loaded once and reused as row j of
//                 the next tap, the x weights are computed once per task, the y weights of a (query, level)
//                 once per warp (a shared-memory table, read as broadcasts).  11 float operations + 2.2
//                 shared loads per tap instead of ~30 instructions.
//   stores        a task's 2r+1 outputs are CONSECUTIVE channels (channel = l*(2r+1)^2 + i*(2r+1) + j), and
//                 tasks are consecutive too, so the 32 tasks of a pass fill (2r+1)*128 contiguous bytes of
//                 the output: staged in shared memory (stride 2r+1 words: conflict-free) and sent as ONE
//                 bulk copy (cp.async.bulk shared -> global) per pass — no LSU store wavefronts at all.
//   latency       a warp works on G = 2 queries at a time (8 (query, level) footprints, 72 column tasks: two full
//                 passes of 32 lanes, and the last 8 columns split four ways so the third pass costs a third),
//                 and while it evaluates one group the half-tile loads of its next group (registers) and the
//                 coordinates of the one after (2 registers) are in flight.
//
// Measured on B200 (profiles/lookup_time.py, back to back, 8 coordinate sets): batch 16: 78 us = 4.2 TB/s of
// algorithmic bytes (0.64 of the measured HBM peak), batch 64: 299 us (0.67); the row-major kernel: 100 us (0.50).
// DRAM moves 2557 + 1296 B per query (ncu), i.e. 1.33x the algorithmic 2904 B: with the previous launch's output
// still being written back that is 435 MB per launch = 74 us at the 5.9 TB/s this access pattern reaches
// (profiles/layout_bench.cu), so the kernel runs within ~5 % of what the layout allows.
// Tried and dropped, with ncu in hand (all bit-exact, none faster): one footprint tile per shared-memory
// instruction (27 lanes x 3 taps, pitch 14: conflict-free reads, but every lane recomputes its y weights —
// 77 us); cp.async staging straight into double-buffered tiles (88 us: LDGSTS issue rate); no software pipeline
// with 5 blocks per SM (104 us); one query per warp (113 us).
//
// Arithmetic.  Per tap and per axis the reference evaluates (jax map_coordinates, order 1)
//     s = c_l + d;  i0 = floor(s);  w1 = s - i0;  w0 = 1 - w1          (c_l = c / 2^l, exact; d = offset)
// and sums (wy*wx)*v over the corners (y0,x0), (y0,x1), (y1,x0), (y1,x1) in float32 without contraction.
// fl(c_l + d) can round UP to the integer floor(c_l)+d+1; then i0 is one more than the regular
// floor(c_l)+d and w1 = 0, w0 = 1.  This kernel always anchors the tap at the regular cell floor(c_l)+d and
// uses  w1' = s - (floor(c_l) + d)  (exact: both are multiples of ulp(s)),  w0' = 1 - w1':  identical in the
// regular case; in the round-up case w1' = 1, w0' = 0, which moves the two unit-weight corners from
// "(i0, i0+1) with weights (1, 0)" to "(i0-1, i0) with weights (0, 1)" — the same two products summed in the
// same order, plus zeros.  Results equal the float32 oracle bit for bit (a zero may differ in sign).
// Queries with non-finite or |c| > 1e6 coordinates read nothing: their footprints are zeros and the weight
// arithmetic above yields the reference's 0 (finite c) or NaN (non-finite c).

#include <cuda_fp16.h>

#include "pyramid_tiled.cuh"
#include "lookup_taps.cuh"

namespace raftk {
namespace lt {

struct Params {
  const void* level[4];  // float32 slabs, or (HALF kernels) float16 slabs of the same tile geometry
  long long slab[4];     // ELEMENTS per query slab
  int lh[4], lw[4], tx[4];
  const float* coords;
  float* out;
  int Q;
  // fused coordinate update (raftk_corr_lookup_tiled_update): coords_rw += delta + dbias first
  const float* delta;
  int ld_delta;
  const float* dbias;
  const float* coords0;
  float* coords_rw;
  float* flow;
  float* flow_pitched;  // may be NULL
  int ld_flow;
};

struct QLRec {  // 32 bytes, one per (query, level) of a group
  float cxl, cyl, fcx, fcy;  // c / 2^l and its floor
  int bxa, by, xs, far;      // footprint origin: x rounded down to a multiple of 4 (+ xs), y exact; far: read nothing
};

template <int R, int G>
struct Geo {
  static constexpr int S1 = 2 * R + 1;           // taps per axis
  static constexpr int FP = 2 * R + 2;           // footprint side
  static constexpr int KP = FP / 2 + 1;          // row pairs that can hold footprint rows (even-aligned)
  static constexpr int NC = (FP + 2) / 4 + 1;    // 16-byte column chunks (4-aligned)
  static constexpr int SLOTS = KP * NC;          // half-tile loads per (query, level)
  static constexpr int P = 4 * NC + 4;           // tile row pitch in floats (R = 4: 20, R = 3: 16)
  // floats per (query, level) tile: the smallest size >= FP * P that is 12 (mod 32), so that the column reads of
  // consecutive tiles (9 lanes each, shifted by xs <= 3) start 12 banks apart
  static constexpr int TILE = FP * P + ((12 - (FP * P) % 32) + 32) % 32;
  static constexpr int NQL = 4 * G;
  static constexpr int LROUNDS = (NQL * SLOTS + 31) / 32;
  static constexpr int TASKS = NQL * S1;
  static constexpr int PASSES = (TASKS + 31) / 32;
  // the last, partial pass: its REM column tasks are split over SPL lanes each (JP y offsets per lane), so that it
  // costs JP / S1 of a full pass instead of a whole one
  static constexpr int FULL = TASKS / 32, REM = TASKS % 32;
  static constexpr int SPL = REM ? (32 / REM < S1 ? 32 / REM : S1) : 1;
  static constexpr int JP = (S1 + SPL - 1) / SPL;
  static constexpr int C = 4 * S1 * S1;
  static constexpr int OSTAGE = 32 * S1;  // floats per staged pass
  // per-warp shared memory, in floats
  static constexpr int F_OFF = 0;
  static constexpr int REC_OFF = F_OFF + NQL * TILE;
  static constexpr int WY_OFF = REC_OFF + 2 * NQL * 8;
  static constexpr int OS_OFF = WY_OFF + ((TASKS + 3) / 4) * 4;
  static constexpr int WARP_FLOATS = OS_OFF + 2 * OSTAGE;
};

__device__ __forceinline__ void ldg256_hint(const float* p, float* v, uint64_t pol) {
  asm volatile("ld.global.nc.L2::cache_hint.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8], %9;"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
               : "l"(p), "l"(pol));
}
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_store(float* dst, const float* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_addr(src_smem)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ float load_as_float(const float* p) { return __ldg(p); }
__device__ __forceinline__ float load_as_float(const __half* p) { return __half2float(__ldg(p)); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }


// NJ consecutive taps (y offsets j0 .. j0 + NJ - 1, those < S1) of one column task: Fp = the column's first footprint
// value of row j0, wyp = w1 of y offset j0, o = the task's output slot of y offset j0.  All shared-memory reads
// first, so that their latency overlaps instead of stalling each tap.
template <int NJ, int P>
__device__ __forceinline__ void column_taps(const float* Fp, const float* wyp, float wx0, float wx1, float* o, int nj) {
  float fa[NJ + 1], fb[NJ + 1], w1[NJ];
#pragma unroll
  for (int r = 0; r <= NJ; ++r) {
    fa[r] = Fp[r * P];
    fb[r] = Fp[r * P + 1];
  }
#pragma unroll
  for (int j = 0; j < NJ; ++j) w1[j] = wyp[j];
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const float w0 = __fadd_rn(1.0f, -w1[j]);
    float acc = __fmul_rn(__fmul_rn(w0, wx0), fa[j]);
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(w0, wx1), fb[j]));
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(w1[j], wx0), fa[j + 1]));
    acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(w1[j], wx1), fb[j + 1]));
    if (j < nj) o[j] = acc;
  }
}

// HALF: the slabs hold float16 (pyramid_tiled.cuh, "float16 storage"): a 4 x 4 tile is 32 bytes, so one LDG.256 brings
// a whole tile (two per lane and query instead of three half tiles); values are widened to float32 while they are
// staged, and everything after the staging — weights, taps, stores — is the float32 code path unchanged.
template <int R, int G, int WARPS, int MINB, bool PIPE, bool UPDATE, bool HALF>
__global__ void __launch_bounds__(WARPS * 32, MINB) lookup_tiled_kernel(const Params p) {
  using X = Geo<R, G>;
  constexpr int S1 = X::S1, FP = X::FP, P = X::P, TILE = X::TILE, NQL = X::NQL;
  extern __shared__ __align__(16) float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* W = smem + warp * X::WARP_FLOATS;
  float* F = W + X::F_OFF;
  QLRec* rec = reinterpret_cast<QLRec*>(W + X::REC_OFF);
  float* wy = W + X::WY_OFF;  // w1 of (query-level, y offset); w0 = 1 - w1
  float* ostage = W + X::OS_OFF;

  // ---- per-lane constants of the load phase: a lane always serves the same level and the same 16-byte column
  // chunk; lanes 8L..8L+7 fetch level L: chunk c = lane & 3, row pairs k = 2*hi + j (j = 0, 1), 4 + hi (j = 2)
  const int lvl = lane >> 3, c4 = 4 * (lane & 3), hi = (lane >> 2) & 1;
  // third round: rows only 2 pairs apart, so the hi = 1 lanes take the chunks two columns over (different banks)
  const int c4x = c4 ^ (hi << 3);
  const char* const lbase = static_cast<const char*>(p.level[lvl]);
  const long long lslab = p.slab[lvl] * (HALF ? 2 : 4);  // bytes per query slab
  const int llh = p.lh[lvl], llw = p.lw[lvl], ltx = p.tx[lvl];
  // ---- per-lane constants of the tap phase: task T = lane + 32 * pass -> (query-level ql, x offset i)
  // (full passes: one column per lane; remainder pass: column REM_T0 + lane / SPL, y offsets JP * (lane % SPL) ..)
  int tql[X::PASSES], ti[X::PASSES];
#pragma unroll
  for (int ps = 0; ps < X::PASSES; ++ps) {
    const int T = ps < X::FULL ? lane + 32 * ps : 32 * X::FULL + lane / X::SPL;
    tql[ps] = T / S1;
    ti[ps] = T - tql[ps] * S1;
  }
  const int rj0 = X::JP * (lane % X::SPL);  // remainder pass: first y offset of this lane
  const bool ractive = lane / X::SPL < X::REM && rj0 < S1;
  pdl_wait();  // the coordinates come from the kernel before this one

  const uint64_t pol = l2_policy_evict_last();
  const int ngroups = (p.Q + G - 1) / G;
  const int gstride = gridDim.x * WARPS;
  constexpr unsigned kRecMask = NQL >= 32 ? 0xffffffffu : ((1u << NQL) - 1u);

  // ---- phase A: the (query, level) records of group g, from coordinates fetched one iteration earlier ----
  // fetch_coords(g) leaves the coordinates of lane's query of group g in (nx, ny); with the fused update these
  // are the operands of UpdateCell's tail instead (model.py:504-506) and make_records does the arithmetic
  float nx = 0.0f, ny = 0.0f;
  float2 n_d = make_float2(0.0f, 0.0f), n_c0 = make_float2(0.0f, 0.0f);
  auto fetch_coords = [&](int g) {
    if (lane < NQL) {
      const int q = g * G + (lane >> 2);
      if (g < ngroups && q < p.Q) {
        if (!UPDATE) {
          const float2 c = __ldg(reinterpret_cast<const float2*>(p.coords) + q);
          nx = c.x;
          ny = c.y;
        } else if ((lane & 3) == 0) {
          const float2 c1 = *reinterpret_cast<const float2*>(p.coords_rw + 2 * (size_t)q);
          nx = c1.x;
          ny = c1.y;
          n_d = *reinterpret_cast<const float2*>(p.delta + (size_t)q * p.ld_delta);
          n_c0 = __ldg(reinterpret_cast<const float2*>(p.coords0) + q);
        }
      }
    }
  };
  float2 dbias = make_float2(0.0f, 0.0f);
  if (UPDATE) dbias = __ldg(reinterpret_cast<const float2*>(p.dbias));
  auto make_records = [&](int g, QLRec* rq) {
    if (lane < NQL) {
      const int qi = lane >> 2, l = lane & 3;
      const int q = g * G + qi;
      const bool valid = q < p.Q;
      float x = valid ? nx : 0.0f, y = valid ? ny : 0.0f;
      if (UPDATE) {
        // done by the level-0 lane of the query alone (one reader-writer per query), then handed to the other three
        if (valid && l == 0) {
          x = __fadd_rn(nx, __fadd_rn(n_d.x, dbias.x));
          y = __fadd_rn(ny, __fadd_rn(n_d.y, dbias.y));
          const float2 f = make_float2(__fadd_rn(x, -n_c0.x), __fadd_rn(y, -n_c0.y));
          *reinterpret_cast<float2*>(p.coords_rw + 2 * (size_t)q) = make_float2(x, y);
          *reinterpret_cast<float2*>(p.flow + 2 * (size_t)q) = f;
          if (p.flow_pitched) *reinterpret_cast<float2*>(p.flow_pitched + (size_t)q * p.ld_flow) = f;
        }
        x = __shfl_sync(kRecMask, x, lane & ~3);
        y = __shfl_sync(kRecMask, y, lane & ~3);
      }
      const bool far = !valid || !(fabsf(x) <= 1.0e6f && fabsf(y) <= 1.0e6f);
      const float sc = __int_as_float((127 - l) << 23);  // 2^-l
      QLRec r;
      r.cxl = x * sc;
      r.cyl = y * sc;
      r.fcx = floorf(r.cxl);
      r.fcy = floorf(r.cyl);
      const int bx = far ? 0 : (int)r.fcx - R;
      r.by = far ? 0 : (int)r.fcy - R;
      r.xs = bx & 3;
      r.bxa = bx - r.xs;
      r.far = far ? 1 : 0;
      rq[lane] = r;
    }
    __syncwarp();
  };

  // ---- phase B: half-tile loads of group g into registers: round (qi, j) = this lane's row pair k(j) of
  // footprint (qi, lvl); state: bit 2m = the slot holds footprint rows, bit 2m+1 = ... and was loaded ----
  float v[(HALF ? 2 : 3) * G][8];
  unsigned state = 0;
  auto prefetch = [&](int g, const QLRec* rq) {
    state = 0;
#pragma unroll
    for (int qi = 0; qi < G; ++qi) {
      const int4 o = *reinterpret_cast<const int4*>(&rq[qi * 4 + lvl].bxa);  // bxa, by, xs, far
      const char* slab = lbase + (size_t)(g * G + qi) * lslab;
      if (HALF) {
        // slot (qi, j) = whole tile: tile column c4 / 4, tile row 2 * j + hi of the footprint's 4 x 4 tile block
        const int ys = o.y & 3;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int m = qi * 2 + j;
          const int tyr = 2 * j + hi;
          const int gx = o.x + c4;
          const bool xneed = c4 <= o.z + FP - 1;
          const bool xin = !o.w && (unsigned)gx < (unsigned)llw;
          const int gy = o.y - ys + 4 * tyr;
          const bool need = xneed && 4 * tyr <= FP - 1 + ys;
          const bool ld = need && xin && (unsigned)gy < (unsigned)llh;
          if (need) state |= 1u << (2 * m);
          if (ld) {
            state |= 2u << (2 * m);
            ldg256_hint(reinterpret_cast<const float*>(slab + ((size_t)((gy >> 2) * ltx + (gx >> 2))) * 32), v[m], pol);
          }
        }
        continue;
      }
      const int ys = o.y & 1;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int m = qi * 3 + j;
        const int k = j < 2 ? 2 * hi + j : 4 + hi;
        const int cc = j < 2 ? c4 : c4x;
        const int gx = o.x + cc;
        const bool xneed = cc <= o.z + FP - 1;
        const bool xin = !o.w && (unsigned)gx < (unsigned)llw;
        const int gy = o.y - ys + 2 * k;
        const bool need = xneed && 2 * k <= FP - 1 + ys;
        const bool ld = need && xin && (unsigned)gy < (unsigned)llh;
        if (need) state |= 1u << (2 * m);
        if (ld) {
          state |= 2u << (2 * m);
          ldg256_hint(reinterpret_cast<const float*>(slab + ((size_t)(((gy >> 2) * ltx + (gx >> 2)) * 16 + (gy & 2) * 4)) * 4), v[m],
                      pol);
        }
      }
    }
  };

  // ---- phase C: registers -> footprint tiles F[ql][FP rows][P] (row 0 = footprint row by, column 0 = bxa);
  // a needed slot that was not loaded (outside the level, far query) stores zeros ----
  auto stage = [&](const QLRec* rq) {
#pragma unroll
    for (int qi = 0; qi < G; ++qi) {
      float* tile = F + (qi * 4 + lvl) * TILE;
      if (HALF) {
        const int ys = rq[qi * 4 + lvl].by & 3;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int m = qi * 2 + j;
          const int r0 = 4 * (2 * j + hi) - ys;
          const bool need = (state >> (2 * m)) & 1u, ld = (state >> (2 * m + 1)) & 1u;
          float* t = tile + r0 * P + c4;
#pragma unroll
          for (int rr = 0; rr < 4; ++rr) {  // tile row rr = 4 halves = words 2 rr, 2 rr + 1
            // a slot that was not loaded (outside the level) is zero BEFORE the widening: the select then sits on
            // the load's own predicate (zero-initialised registers), not on every widened value
            const float w0 = ld ? v[m][2 * rr] : 0.0f, w1 = ld ? v[m][2 * rr + 1] : 0.0f;
            const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&w0));
            const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&w1));
            if (need && (unsigned)(r0 + rr) < (unsigned)FP) *reinterpret_cast<float4*>(t + rr * P) = make_float4(a.x, a.y, b.x, b.y);
          }
        }
        continue;
      }
      const int ys = rq[qi * 4 + lvl].by & 1;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int m = qi * 3 + j;
        const int k = j < 2 ? 2 * hi + j : 4 + hi;
        const int r0 = 2 * k - ys;
        const bool need = (state >> (2 * m)) & 1u, ld = (state >> (2 * m + 1)) & 1u;
        float* t = tile + r0 * P + (j < 2 ? c4 : c4x);
        const float4 lo = ld ? make_float4(v[m][0], v[m][1], v[m][2], v[m][3]) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        const float4 up = ld ? make_float4(v[m][4], v[m][5], v[m][6], v[m][7]) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (need && (unsigned)r0 < (unsigned)FP) *reinterpret_cast<float4*>(t) = lo;
        if (need && (unsigned)(r0 + 1) < (unsigned)FP) *reinterpret_cast<float4*>(t + P) = up;
      }
    }
  };

  int obuf = 0;
  // ---- phase D: y-weight table, then the column tasks ----
  auto taps = [&](int g, const QLRec* rq) {
    // y weights of (ql, j): one table entry per lane and round (there are as many entries as column tasks)
#pragma unroll
    for (int ps = 0; ps < X::PASSES; ++ps) {
      const int T = lane + 32 * ps;
      if (T < X::TASKS) {
        const int ql = T / S1, j = T - ql * S1;
        const float cyl = rq[ql].cyl, fcy = rq[ql].fcy;
        const float d = (float)(j - R);
        const float s = __fadd_rn(cyl, d), t = __fadd_rn(fcy, d);
        wy[T] = __fadd_rn(s, -t);
      }
    }
    __syncwarp();  // tiles and weight table visible to the whole warp
    const int nvalid = min(G, p.Q - g * G);
    const int tasks_valid = nvalid * 4 * S1;
    float* obase = p.out + (size_t)g * G * X::C;
#pragma unroll
    for (int ps = 0; ps < X::PASSES; ++ps) {
      float* og = ostage + obuf * X::OSTAGE;
      if (lane == 0) bulk_wait_read<1>();  // the bulk copy that last read this staging buffer has drained it
      __syncwarp();
      const bool full = ps < X::FULL;
      if (full || ractive) {
        const int ql = tql[ps], i = ti[ps];
        const float cxl = rq[ql].cxl, fcx = rq[ql].fcx;
        const int xs = rq[ql].xs;
        const float d = (float)(i - R);
        const float sx = __fadd_rn(cxl, d), tx = __fadd_rn(fcx, d);
        const float wx1 = __fadd_rn(sx, -tx);
        const float wx0 = __fadd_rn(1.0f, -wx1);
        const float* Fp = F + ql * TILE + xs + i;
        const float* wyp = wy + ql * S1;
        if (full) {
          column_taps<S1, P>(Fp, wyp, wx0, wx1, og + lane * S1, S1);
        } else {
          column_taps<X::JP, P>(Fp + rj0 * P, wyp + rj0, wx0, wx1, og + (lane / X::SPL) * S1 + rj0, S1 - rj0);
        }
      }
      fence_async_smem();  // generic-proxy writes -> visible to the bulk copy engine
      __syncwarp();
      if (lane == 0) {
        const int n = min(32, tasks_valid - 32 * ps);
        if (n > 0) bulk_store(obase + (size_t)32 * ps * S1, og, (uint32_t)(n * S1 * 4));
        bulk_commit();
      }
      obuf ^= 1;
    }
  };

  int g = blockIdx.x * WARPS + warp;
  if (PIPE) {
    // software pipeline: during the taps of group g, the half-tile loads of this warp's next group and the
    // coordinates of the one after it are in flight
    int rb = 0;
    if (g < ngroups) {
      fetch_coords(g);
      make_records(g, rec);
      prefetch(g, rec);
      fetch_coords(g + gstride);
    }
    while (g < ngroups) {
      const QLRec* rq = rec + rb * NQL;
      stage(rq);
      const int gn = g + gstride;
      if (gn < ngroups) {
        make_records(gn, rec + (rb ^ 1) * NQL);
        prefetch(gn, rec + (rb ^ 1) * NQL);
        fetch_coords(gn + gstride);
      }
      taps(g, rq);
      __syncwarp();  // every lane is done with the tiles before the next group overwrites them
      g = gn;
      rb ^= 1;
    }
  } else {
    // no registers held across the taps: other warps of the SM cover the load latency
    fetch_coords(g);
    for (; g < ngroups; g += gstride) {
      make_records(g, rec);
      prefetch(g, rec);
      fetch_coords(g + gstride);
      stage(rec);
      taps(g, rec);
      __syncwarp();
    }
  }
  if (lane == 0) bulk_wait_read<0>();  // shared memory must outlive the last bulk copies
}


// ---- generic path over tiled slabs (any radius <= 7, <= RAFTK_MAX_LEVELS levels): one warp per query, scalar
// staged footprints, the tap code of corr_lookup.cu ----
struct GenericParams {
  const void* level[RAFTK_MAX_LEVELS];
  long long slab[RAFTK_MAX_LEVELS];
  int lh[RAFTK_MAX_LEVELS], lw[RAFTK_MAX_LEVELS], tx[RAFTK_MAX_LEVELS];
  const float* coords;
  float* out;
  int Q, L, r;
};

template <typename T>
__global__ void __launch_bounds__(256) lookup_tiled_generic_kernel(const GenericParams p) {
  extern __shared__ float smem[];
  const int r = p.r, S = 2 * r + 1, FP = 2 * r + 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = blockIdx.x * 8 + warp;
  if (q >= p.Q) return;
  float* F = smem + warp * (p.L * FP * FP);
  const float cx = __ldg(p.coords + 2 * (size_t)q), cy = __ldg(p.coords + 2 * (size_t)q + 1);
  float sc = 1.0f;
  for (int l = 0; l < p.L; ++l, sc *= 0.5f) {
    const int lh = p.lh[l], lw = p.lw[l], tx = p.tx[l];
    const int bx = clamp_to_int(floorf(cx * sc)) - r, by = clamp_to_int(floorf(cy * sc)) - r;
    const T* slab = static_cast<const T*>(p.level[l]) + (size_t)q * p.slab[l];
    float* Fl = F + l * FP * FP;
    for (int e = lane; e < FP * FP; e += 32) {
      const int ry = e / FP, rx = e - ry * FP;
      const int gy = by + ry, gx = bx + rx;
      float val = 0.0f;
      if ((unsigned)gy < (unsigned)lh && (unsigned)gx < (unsigned)lw) val = load_as_float(slab + tiled_offset(gy, gx, tx));
      Fl[e] = val;
    }
  }
  __syncwarp();
  lookup_taps<0>(F, cx, cy, p.L, r, p.out + (size_t)q * (p.L * S * S), lane);
}

}  // namespace lt
}  // namespace raftk

struct raftk_tiled_update_args {
  const float* delta;
  int ld_delta;
  const float* bias;
  const float* coords0;
  float* coords1;
  float* flow;
  float* flow_pitched;
  int ld_flow;
};

namespace {

template <int R, int G, int WARPS, int MINB, bool PIPE, bool UPDATE, bool HALF>
int launch_fast(const raftk::lt::Params& p, cudaStream_t s) {
  using namespace raftk;
  using X = lt::Geo<R, G>;
  const size_t smem = (size_t)(WARPS * X::WARP_FLOATS) * sizeof(float);
  auto kernel = lt::lookup_tiled_kernel<R, G, WARPS, MINB, PIPE, UPDATE, HALF>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return static_cast<int>(e);
  }
  const int ngroups = (p.Q + G - 1) / G;
  int blocks = (ngroups + WARPS - 1) / WARPS;
  const int cap = sm_count() * MINB;
  if (blocks > cap) blocks = cap;
  return launch_pdl(kernel, blocks, WARPS * 32, smem, s, p);
}

// two queries per warp, 4 warps per block, 4 blocks per SM (128 registers), software-pipelined; the schedules this
// was chosen against are in the file header
template <int R, bool UPDATE, bool HALF>
int launch_fast_r(const raftk::lt::Params& p, cudaStream_t s) {
  return launch_fast<R, 2, 4, 4, true, UPDATE, HALF>(p, s);
}

template <bool UPDATE, bool HALF>
int launch_fast_rh(const raftk::lt::Params& p, int radius, cudaStream_t s) {
  return radius == 4 ? launch_fast_r<4, UPDATE, HALF>(p, s) : launch_fast_r<3, UPDATE, HALF>(p, s);
}

int lookup_tiled_launch(const void* const* levels, int num_levels, const float* coords, float* out, int Q, int h, int w,
                        int radius, const raftk_tiled_update_args* upd, void* stream, bool half) {
  using namespace raftk;
  if (Q < 0 || h <= 0 || w <= 0 || radius < 0 || num_levels < 1 || num_levels > RAFTK_MAX_LEVELS) return RAFTK_ERR_SHAPE;
  if ((h >> (num_levels - 1)) < 1 || (w >> (num_levels - 1)) < 1) return RAFTK_ERR_TOO_SMALL;
  if (radius > 7) return RAFTK_ERR_UNSUPPORTED;
  if (Q == 0) return RAFTK_OK;
  if (!levels || !coords || !out) return RAFTK_ERR_NULL;
  for (int l = 0; l < num_levels; ++l) {
    if (!levels[l]) return RAFTK_ERR_NULL;
    if ((uintptr_t)levels[l] & 127) return RAFTK_ERR_ALIGN;
  }
  if (((uintptr_t)out & 15) || ((uintptr_t)coords & 7)) return RAFTK_ERR_ALIGN;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool fast = (radius == 3 || radius == 4) && num_levels == 4;
  if (!fast) {
    if (upd) return RAFTK_ERR_UNSUPPORTED;
    lt::GenericParams g{};
    for (int l = 0; l < num_levels; ++l) {
      g.level[l] = levels[l];
      g.lh[l] = h >> l;
      g.lw[l] = w >> l;
      g.tx[l] = tiled_tiles_x(w >> l);
      g.slab[l] = tiled_slab_floats(h >> l, w >> l);
    }
    g.coords = coords;
    g.out = out;
    g.Q = Q;
    g.L = num_levels;
    g.r = radius;
    const int FP = 2 * radius + 2;
    const size_t smem = (size_t)8 * num_levels * FP * FP * sizeof(float);
    if (half)
      lt::lookup_tiled_generic_kernel<__half><<<(Q + 7) / 8, 256, smem, s>>>(g);
    else
      lt::lookup_tiled_generic_kernel<float><<<(Q + 7) / 8, 256, smem, s>>>(g);
    return launch_status();
  }
  lt::Params p{};
  for (int l = 0; l < 4; ++l) {
    p.level[l] = levels[l];
    p.lh[l] = h >> l;
    p.lw[l] = w >> l;
    p.tx[l] = tiled_tiles_x(w >> l);
    p.slab[l] = tiled_slab_floats(h >> l, w >> l);
  }
  p.coords = coords;
  p.out = out;
  p.Q = Q;
  if (upd) {
    p.delta = upd->delta;
    p.ld_delta = upd->ld_delta;
    p.dbias = upd->bias;
    p.coords0 = upd->coords0;
    p.coords_rw = upd->coords1;
    p.flow = upd->flow;
    p.flow_pitched = upd->flow_pitched;
    p.ld_flow = upd->ld_flow;
    return half ? launch_fast_rh<true, true>(p, radius, s) : launch_fast_rh<true, false>(p, radius, s);
  }
  return half ? launch_fast_rh<false, true>(p, radius, s) : launch_fast_rh<false, false>(p, radius, s);
}

int check_update_args(const float* delta_pre, int ld_delta, const float* bias, const float* coords0, float* coords1,
                      float* flow, float* flow_pitched, int ld_flow, int Q) {
  if (Q > 0) {
    if (!delta_pre || !bias || !coords0 || !coords1 || !flow) return RAFTK_ERR_NULL;
    if (ld_delta < 2 || ld_delta % 2 || (flow_pitched && (ld_flow < 2 || ld_flow % 2))) return RAFTK_ERR_SHAPE;
    if (((uintptr_t)delta_pre | (uintptr_t)bias | (uintptr_t)coords0 | (uintptr_t)coords1 | (uintptr_t)flow |
         (uintptr_t)flow_pitched) & 7)
      return RAFTK_ERR_ALIGN;
  }
  return RAFTK_OK;
}

}  // namespace

extern "C" int raftk_corr_lookup_tiled(const float* const* levels, int num_levels, const float* coords, float* out, int Q,
                                       int h, int w, int radius, void* stream) {
  return lookup_tiled_launch(reinterpret_cast<const void* const*>(levels), num_levels, coords, out, Q, h, w, radius, nullptr,
                             stream, false);
}

extern "C" int raftk_corr_lookup_tiled_update(const float* const* levels, int num_levels, const float* delta_pre,
                                              int ld_delta, const float* bias, const float* coords0, float* coords1,
                                              float* flow, float* flow_pitched, int ld_flow, float* out, int Q, int h,
                                              int w, int radius, void* stream) {
  const int rc = check_update_args(delta_pre, ld_delta, bias, coords0, coords1, flow, flow_pitched, ld_flow, Q);
  if (rc != RAFTK_OK) return rc;
  const raftk_tiled_update_args upd{delta_pre, ld_delta, bias, coords0, coords1, flow, flow_pitched, ld_flow};
  return lookup_tiled_launch(reinterpret_cast<const void* const*>(levels), num_levels, coords1, out, Q, h, w, radius, &upd,
                             stream, false);
}

extern "C" int raftk_corr_lookup_tiled_f16(const void* const* levels, int num_levels, const float* coords, float* out, int Q,
                                           int h, int w, int radius, void* stream) {
  return lookup_tiled_launch(levels, num_levels, coords, out, Q, h, w, radius, nullptr, stream, true);
}

extern "C" int raftk_corr_lookup_tiled_f16_update(const void* const* levels, int num_levels, const float* delta_pre,
                                                  int ld_delta, const float* bias, const float* coords0, float* coords1,
                                                  float* flow, float* flow_pitched, int ld_flow, float* out, int Q, int h,
                                                  int w, int radius, void* stream) {
  const int rc = check_update_args(delta_pre, ld_delta, bias, coords0, coords1, flow, flow_pitched, ld_flow, Q);
  if (rc != RAFTK_OK) return rc;
  const raftk_tiled_update_args upd{delta_pre, ld_delta, bias, coords0, coords1, flow, flow_pitched, ld_flow};
  return lookup_tiled_launch(levels, num_levels, coords1, out, Q, h, w, radius, &upd, stream, true);
}
