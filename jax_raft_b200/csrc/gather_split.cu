// Query-row sharding (BASELINE config 4): the one exchange of the correlation path — every rank needs
// ALL of fmap2 — fused with the first thing done to the gathered map.  No reference counterpart (the reference is
// single-device).
//
// The baseline is an NCCL all-gather into a local buffer followed by the bf16 hi/mid split of level 0
// (raftk_altcorr_split).  Here every rank keeps its band of fmap2 rows in peer-accessible memory (symmetric memory:
// the host passes one device pointer per rank) and ONE kernel reads each band where it lives — its own from HBM, the
// others over NVLink with plain 16-byte loads, every element exactly once — and writes both results locally: the full
// float32 map (the input of the pooled levels and of the scalar paths) and the [hi | mid] bf16 operands of the
// tensor-core lookup.  The gathered map is never re-read for the split, and no collective library call sits between
// the two steps.  Loads are ordinary ld.global (no TMA, no multicast): a remote band crosses NVLink once per reader.
// Image rows are walked round-robin over the bands, starting at the next rank's, so NVLink reads (the slower side)
// are in flight from the first wave to the last, under the local reads and all the writes, and the ranks do not all
// pull from the same peer at once.
//
// Synchronisation, two ways:
//  * sync_flags == NULL: the caller's (a device-side barrier before the launch — every band is written — and after
//    it — nobody overwrites a band that is still being read);
//  * sync_flags given (one 256-byte SyncBlock per rank, next to its band in the same peer-mapped allocation, zeroed
//    once): the kernel does both itself.  Call e of every rank: block 0 stores e into arrive[rank] of every peer
//    (st.release.sys — the band was written by earlier kernels of this stream); every block waits until its own
//    arrive[g] >= e for all g before it reads; the last block to finish (local counter) stores e into done[rank] of
//    every peer and waits for its own done[g] >= e, so when the kernel completes nobody is still reading this rank's
//    band.  e is a device-side counter (SyncBlock::epoch), not an argument: a captured launch replays correctly.
//    Waits are bounded (a few seconds) and set SyncBlock::timeout instead of hanging the GPU.

#include <cuda_bf16.h>

#include "raftk_common.cuh"

namespace raftk {
namespace gs {

constexpr int kMaxRanks = 8;

struct Params {
  const float* band[kMaxRanks];  // rank g's rows [y0[g], y0[g + 1]) of fmap2: (B, rows_g, w, D)
  int y0[kMaxRanks + 1];
  int world;
  float* full;            // (B, h, w, D)
  __nv_bfloat16* split;   // (B, h, w, 2D) or NULL
  int B, h, w, D;
  int max_rows;            // the largest band
};

struct SyncBlock {       // 256 bytes per rank, zero-initialised once; peers write arrive[] / done[], the rest is local
  unsigned arrive[kMaxRanks];  // [g] = e: rank g's band of call e is complete
  unsigned done[kMaxRanks];    // [g] = e: rank g has finished reading THIS rank's band in call e
  unsigned epoch;              // calls completed by this rank
  unsigned blocks_done;        // blocks of the running call that have finished
  unsigned timeout;            // sticky: a wait gave up
  unsigned pad[64 - 2 * kMaxRanks - 3];
};
static_assert(sizeof(SyncBlock) == 256, "SyncBlock is part of the ABI (RAFTK_GATHER_SYNC_BYTES)");

struct SyncParams {
  SyncBlock* flags[kMaxRanks];  // [rank] local, the others peer-mapped; flags[0] == NULL: no in-kernel synchronisation
  int rank;
};

// system-scope accesses: bands and flags are written by other GPUs
__device__ __forceinline__ float4 ld_sys_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Bounded wait for *p >= e (wrap-safe); false after ~4 s.
__device__ __forceinline__ bool wait_at_least(const unsigned* p, unsigned e) {
  for (unsigned spins = 0; spins < (1u << 25); ++spins) {
    if ((int)(ld_acquire_sys(p) - e) >= 0) return true;
    __nanosleep(spins < 64 ? 20 : 128);
  }
  return false;
}

constexpr int kThreads = 256;
constexpr int kUnroll = 4;                       // 16-byte loads in flight per thread: NVLink latency is several times HBM's
constexpr int kSegment = kThreads * kUnroll;     // float4 per work item
constexpr int kBlocksPerSM = 4;                  // 64 KB of loads in flight per SM

// Work item = (row slot, segment of kSegment float4 of that image row); row slots enumerate (b, band, row in band)
// with the band fastest.  The loads of an item are all issued before its first store; a software pipeline over items
// (next item's loads before this item's stores) changed nothing — the remote side runs at what plain loads pull over
// NVLink (~480 GB/s measured on 2 GPUs, ~530 GB/s per rank on 8), not at a latency the loop exposes; 32-byte loads
// (LDG.256, three per thread) gave the same 36 us, so the request count is not what bounds it either.
__global__ void __launch_bounds__(kThreads, kBlocksPerSM) gather_split_kernel(const Params p, const SyncParams sp) {
  __shared__ int s_last;
  __shared__ const float* s_band[kMaxRanks];  // dynamically indexed: shared memory, not a local copy of the parameters
  __shared__ int s_y0[kMaxRanks + 1];
  if (threadIdx.x <= kMaxRanks) s_y0[threadIdx.x] = p.y0[threadIdx.x];
  if (threadIdx.x < kMaxRanks) s_band[threadIdx.x] = p.band[threadIdx.x];
  __syncthreads();
  SyncBlock* const mine = sp.flags[0] ? sp.flags[sp.rank] : nullptr;
  unsigned e = 0;
  if (mine) {
    e = *reinterpret_cast<volatile unsigned*>(&mine->epoch) + 1;
    const int t = threadIdx.x;
    if (t < p.world && t != sp.rank) {
      if (blockIdx.x == 0) st_release_sys(&sp.flags[t]->arrive[sp.rank], e);
      if (!wait_at_least(&mine->arrive[t], e)) mine->timeout = 1;
    }
    __syncthreads();
  }
  const int d4 = p.D >> 2;
  const int row_elems = p.w * d4;  // float4 per image row
  const int segs = (row_elems + kSegment - 1) / kSegment;
  const int slots_per_image = p.world * p.max_rows;
  const long long items = (long long)p.B * slots_per_image * segs;
  for (long long item = blockIdx.x; item < items; item += gridDim.x) {
    const int slot = (int)(item / segs);
    const int seg = (int)(item - (long long)slot * segs);
    const int b = slot / slots_per_image, q = slot - b * slots_per_image;
    int g = q % p.world + sp.rank + 1;
    if (g >= p.world) g -= p.world;
    const int j = q / p.world;
    const int rows_g = s_y0[g + 1] - s_y0[g];
    if (j >= rows_g) continue;  // an empty slot: band g is shorter than the largest band
    const int img_row = b * p.h + s_y0[g] + j;
    const float* src = s_band[g] + ((long long)b * rows_g + j) * row_elems * 4;
    float* dst_full = p.full + (long long)img_row * row_elems * 4;
    float4 x[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int el = seg * kSegment + u * kThreads + threadIdx.x;
      if (el < row_elems) x[u] = ld_sys_f4(src + (long long)el * 4);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int el = seg * kSegment + u * kThreads + threadIdx.x;
      if (el >= row_elems) continue;
      *reinterpret_cast<float4*>(dst_full + (long long)el * 4) = x[u];
      if (p.split) {
        const int px = el / d4, c = (el - px * d4) * 4;
        const float xs[4] = {x[u].x, x[u].y, x[u].z, x[u].w};
        __nv_bfloat16 hi[4], mid[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          hi[i] = __float2bfloat16_rn(xs[i]);
          mid[i] = __float2bfloat16_rn(xs[i] - __bfloat162float(hi[i]));
        }
        __nv_bfloat16* dst = p.split + ((long long)img_row * p.w + px) * 2 * p.D + c;
        *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(hi);
        *reinterpret_cast<uint2*>(dst + p.D) = *reinterpret_cast<const uint2*>(mid);
      }
    }
  }
  if (mine) {
    // every load of this block has returned (its data went into the stores above)
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      s_last = atomicAdd(&mine->blocks_done, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
      const int t = threadIdx.x;
      if (t < p.world && t != sp.rank) {
        st_release_sys(&sp.flags[t]->done[sp.rank], e);
        if (!wait_at_least(&mine->done[t], e)) mine->timeout = 1;
      }
      __syncthreads();
      if (t == 0) {
        mine->blocks_done = 0;
        __threadfence();
        *reinterpret_cast<volatile unsigned*>(&mine->epoch) = e;
      }
    }
  }
}

}  // namespace gs
}  // namespace raftk

extern "C" int raftk_gather_rows_split(const float* const* bands, void* const* sync_flags, int rank, const int* row_start,
                                       int world, float* full, void* split_bf16, int B, int h, int w, int D, void* stream) {
  using namespace raftk;
  if (world < 1 || world > gs::kMaxRanks || B < 0 || h <= 0 || w <= 0 || D <= 0) return RAFTK_ERR_SHAPE;
  if (sync_flags && (rank < 0 || rank >= world)) return RAFTK_ERR_SHAPE;
  if (D % 4) return RAFTK_ERR_UNSUPPORTED;
  if (B == 0) return RAFTK_OK;  // nothing to exchange: no rank launches, the epochs stay in step
  if (!bands || !row_start || !full) return RAFTK_ERR_NULL;
  if (((uintptr_t)full & 15) || ((uintptr_t)split_bf16 & 15)) return RAFTK_ERR_ALIGN;
  gs::Params p{};
  gs::SyncParams sp{};
  if (row_start[0] != 0 || row_start[world] != h) return RAFTK_ERR_SHAPE;
  for (int g = 0; g < world; ++g) {
    if (row_start[g + 1] < row_start[g]) return RAFTK_ERR_SHAPE;
    if (!bands[g] && row_start[g + 1] > row_start[g]) return RAFTK_ERR_NULL;
    if ((uintptr_t)bands[g] & 15) return RAFTK_ERR_ALIGN;
    if (sync_flags) {
      if (!sync_flags[g]) return RAFTK_ERR_NULL;
      if ((uintptr_t)sync_flags[g] & 15) return RAFTK_ERR_ALIGN;
      sp.flags[g] = static_cast<gs::SyncBlock*>(sync_flags[g]);
    }
    p.band[g] = bands[g];
    p.y0[g] = row_start[g];
    if (row_start[g + 1] - row_start[g] > p.max_rows) p.max_rows = row_start[g + 1] - row_start[g];
  }
  for (int g = world; g <= gs::kMaxRanks; ++g) p.y0[g] = h;
  sp.rank = sync_flags ? rank : (rank >= 0 && rank < world ? rank : 0);  // without flags: only the walking order
  p.world = world;
  p.full = full;
  p.split = static_cast<__nv_bfloat16*>(split_bf16);
  p.B = B;
  p.h = h;
  p.w = w;
  p.D = D;
  const int row_elems = w * (D / 4);
  const long long items = (long long)B * world * p.max_rows * ((row_elems + gs::kSegment - 1) / gs::kSegment);
  long long blocks = items;
  if (blocks > (long long)sm_count() * gs::kBlocksPerSM) blocks = (long long)sm_count() * gs::kBlocksPerSM;
  gs::gather_split_kernel<<<(unsigned)blocks, gs::kThreads, 0, static_cast<cudaStream_t>(stream)>>>(p, sp);
  return launch_status();
}
