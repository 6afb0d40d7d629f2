// Tiled pyramid layout (pyramid_tiled.cuh): sizes, and converters between the reference's row-major levels
// (B*nq, h>>l, w>>l) and the tiled slabs.  The converters serve callers that bring their own row-major pyramid
// to the tiled lookup, and the tests (which compare tiled levels with the oracle's row-major ones); the product
// path never converts — raftk_corr_pyramid_tiled writes tiles directly.
#include <cuda_fp16.h>

#include "pyramid_tiled.cuh"

namespace raftk {

// one thread per 16-byte tile row: tiled[q][ty][tx][iy][0..3] <- row-major (zeros outside the level)
__device__ __forceinline__ void store_tile_row(float* dst, float4 v) { *reinterpret_cast<float4*>(dst) = v; }
__device__ __forceinline__ void store_tile_row(__half* dst, float4 v) {  // round to nearest even, as the GEMM epilogue does
  const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
  *reinterpret_cast<uint2*>(dst) = make_uint2(*reinterpret_cast<const unsigned*>(&a), *reinterpret_cast<const unsigned*>(&b));
}
__device__ __forceinline__ float load_cell(const float* p) { return __ldg(p); }
__device__ __forceinline__ float load_cell(const __half* p) { return __half2float(__ldg(p)); }

template <typename T>
__global__ void __launch_bounds__(256) tile_level_kernel(const float* __restrict__ src, T* __restrict__ dst, long long Q,
                                                         int lh, int lw, int tiles_x, int tiles_y) {
  const long long rows_per_slab = (long long)tiles_y * tiles_x * 4;
  const long long total = Q * rows_per_slab;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long q = t / rows_per_slab;
    const int rem = (int)(t - q * rows_per_slab);
    const int iy = rem & 3, tile = rem >> 2;
    const int ty = tile / tiles_x, tx = tile - ty * tiles_x;
    const int y = ty * 4 + iy, x0 = tx * 4;
    float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (y < lh) {
      const float* row = src + (q * lh + y) * lw;
      if (x0 < lw) v.x = __ldg(row + x0);
      if (x0 + 1 < lw) v.y = __ldg(row + x0 + 1);
      if (x0 + 2 < lw) v.z = __ldg(row + x0 + 2);
      if (x0 + 3 < lw) v.w = __ldg(row + x0 + 3);
    }
    store_tile_row(dst + t * 4, v);
  }
}

// one thread per row-major element
template <typename T>
__global__ void __launch_bounds__(256) untile_level_kernel(const T* __restrict__ src, float* __restrict__ dst, long long Q,
                                                           int lh, int lw, int tiles_x, long long slab) {
  const long long per = (long long)lh * lw, total = Q * per;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long q = t / per;
    const int rem = (int)(t - q * per);
    const int y = rem / lw, x = rem - y * lw;
    dst[t] = load_cell(src + q * slab + tiled_offset(y, x, tiles_x));
  }
}

// T = element type of the tiled side; the row-major side is always float32
template <typename T>
static int convert(const void* const* src, void* const* dst, int num_levels, long long Q, int h, int w, bool to_tiled,
                   void* stream) {
  if (Q < 0 || h <= 0 || w <= 0 || num_levels < 1 || num_levels > RAFTK_MAX_LEVELS) return RAFTK_ERR_SHAPE;
  if ((h >> (num_levels - 1)) < 1 || (w >> (num_levels - 1)) < 1) return RAFTK_ERR_TOO_SMALL;
  if (Q == 0) return RAFTK_OK;
  if (!src || !dst) return RAFTK_ERR_NULL;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  for (int l = 0; l < num_levels; ++l) {
    if (!src[l] || !dst[l]) return RAFTK_ERR_NULL;
    const void* tiled = to_tiled ? dst[l] : src[l];
    if ((uintptr_t)tiled & 127) return RAFTK_ERR_ALIGN;
    const int lh = h >> l, lw = w >> l, tx = tiled_tiles_x(lw), ty = tiled_tiles_y(lh);
    const long long work = to_tiled ? Q * ty * tx * 4 : Q * lh * lw;
    long long blocks = (work + 255) / 256;
    if (blocks > (long long)sm_count() * 32) blocks = (long long)sm_count() * 32;
    if (to_tiled)
      tile_level_kernel<T><<<(unsigned)blocks, 256, 0, s>>>(static_cast<const float*>(src[l]), static_cast<T*>(dst[l]), Q, lh, lw,
                                                             tx, ty);
    else
      untile_level_kernel<T><<<(unsigned)blocks, 256, 0, s>>>(static_cast<const T*>(src[l]), static_cast<float*>(dst[l]), Q, lh,
                                                               lw, tx, tiled_slab_floats(lh, lw));
    const int rc = launch_status();
    if (rc != RAFTK_OK) return rc;
  }
  return RAFTK_OK;
}

}  // namespace raftk

extern "C" size_t raftk_pyramid_tiled_slab_floats(int h, int w, int level) {
  if (h <= 0 || w <= 0 || level < 0 || level >= RAFTK_MAX_LEVELS || (h >> level) < 1 || (w >> level) < 1) return 0;
  return (size_t)raftk::tiled_slab_floats(h >> level, w >> level);
}

extern "C" int raftk_pyramid_tile(const float* const* rowmajor, float* const* tiled, int num_levels, long long Q, int h,
                                  int w, void* stream) {
  return raftk::convert<float>(reinterpret_cast<const void* const*>(rowmajor), reinterpret_cast<void* const*>(tiled), num_levels,
                               Q, h, w, true, stream);
}

extern "C" int raftk_pyramid_untile(const float* const* tiled, float* const* rowmajor, int num_levels, long long Q, int h,
                                    int w, void* stream) {
  return raftk::convert<float>(reinterpret_cast<const void* const*>(tiled), reinterpret_cast<void* const*>(rowmajor), num_levels,
                               Q, h, w, false, stream);
}

extern "C" int raftk_pyramid_tile_f16(const float* const* rowmajor, void* const* tiled, int num_levels, long long Q, int h,
                                      int w, void* stream) {
  return raftk::convert<__half>(reinterpret_cast<const void* const*>(rowmajor), tiled, num_levels, Q, h, w, true, stream);
}

extern "C" int raftk_pyramid_untile_f16(const void* const* tiled, float* const* rowmajor, int num_levels, long long Q, int h,
                                        int w, void* stream) {
  return raftk::convert<__half>(tiled, reinterpret_cast<void* const*>(rowmajor), num_levels, Q, h, w, false, stream);
}
