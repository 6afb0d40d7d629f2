// The tiled pyramid layout (round 2): what raftk_corr_pyramid_tiled writes and raftk_corr_lookup_tiled reads.
//
// The reference hands the pyramid from build_pyramid to index_pyramid as an opaque value (model.py:530-539,
// stored at :491/:601), so its memory layout is this library's choice.  Row-major slabs make the lookup pay
// for DRAM it does not use: the L2 fetches 64-byte granules, and a 40-byte footprint row that starts anywhere
// in a row-major slab costs ~100 bytes (measured: 4075 B read per query for 1608 B of footprint).  Here level
// l of one query pixel is stored as TILES of 4 x 4 float32 = exactly one 64-byte granule:
//
//   slab(q, l)  = tiles_y x tiles_x tiles, tile (ty, tx) at float offset (ty * tiles_x + tx) * 16
//   tile        = rows iy = 0..3, row iy at float offset iy * 4, columns ix = 0..3 contiguous
//   cell (y, x) = slab + ((y >> 2) * tiles_x + (x >> 2)) * 16 + (y & 3) * 4 + (x & 3)
//   tiles_y     = ceil(lh / 4),  tiles_x = 2 * ceil(lw / 8)   (a tile row is whole 128-byte lines)
//   padding     = cells with y >= lh or x >= lw inside a tile that holds at least one valid cell are ZERO
//                 (so a reader masks whole 16-byte tile rows only); tiles_x may add one all-padding tile
//                 per tile row, which nobody reads.
//
// A (2r+2)^2 = 10 x 10 footprint then touches (1 + 9/4)^2 = 10.6 granules = 676 B per level instead of
// 15.6 granules = 1000 B (profiles/layout_bench.cu: DRAM read 2557 B instead of 4073 B per query).
#pragma once

#include "raftk_common.cuh"

namespace raftk {

__host__ __device__ __forceinline__ int tiled_tiles_x(int lw) { return 2 * ((lw + 7) >> 3); }
__host__ __device__ __forceinline__ int tiled_tiles_y(int lh) { return (lh + 3) >> 2; }
__host__ __device__ __forceinline__ long long tiled_slab_floats(int lh, int lw) {
  return (long long)tiled_tiles_y(lh) * tiled_tiles_x(lw) * 16;
}
__host__ __device__ __forceinline__ long long tiled_offset(int y, int x, int tiles_x) {
  return ((long long)(y >> 2) * tiles_x + (x >> 2)) * 16 + (y & 3) * 4 + (x & 3);
}

}  // namespace raftk
