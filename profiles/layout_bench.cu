// Micro-benchmarks behind the round-2 pyramid layout: memory skeletons only.
//
//  (a) store side — how fast can 148 SMs push level 0 of one 440x1024 pair (7040 query slabs) when a
//      warp's store instruction carries 32 queries x {128, 256, 512} contiguous bytes, for
//        rowmajor : slab = 55 x 128 floats, one 128-byte row segment per query        (round-1 layout)
//        tiled    : slab = 14 x 32 tiles of 4x4 floats (64 B), tiles of a tile row contiguous, so a
//                   query's share of an 8x32 patch is two runs of 512 B
//  (b) fetch side — the lookup's memory traffic without its arithmetic: per query four 10x10 footprints
//      (one per level) at random positions, read as aligned 16-byte chunks (row-major slabs) or as
//      32-byte half tiles (tiled slabs), plus the 1296-byte output row.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/_build/layout_bench profiles/layout_bench.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// ---------------------------------------------------------------------------------------------------------
// (a) stores.  One "unit" = 32 queries x one 8x32 patch of level 0 = 32 KB, written by ONE warp as
// 32768 / STORE_BYTES bulk tensor stores from an NBUF-deep staging ring.
//   MODE 0  row-major slabs, box {32 floats, 1 row, 32 queries}                       4 KB per store
//   MODE 1  tiled slabs,     box {32 floats (tile pair), 1 pair, 1 tile row, 32 queries}  4 KB per store
//   MODE 2  tiled slabs,     box {32 floats (tile pair), 32 queries, NP pairs}        NP*4 KB per store
// ---------------------------------------------------------------------------------------------------------
template <int MODE, int NP, int NBUF>
__global__ void __launch_bounds__(1024) store_kernel(const __grid_constant__ CUtensorMap tm, int units, int nq_blocks,
                                                     int patches_x, int patches_y) {
  extern __shared__ __align__(1024) unsigned char smem[];
  constexpr int STORE_BYTES = MODE == 2 ? NP * 4096 : 4096;
  constexpr int STORES = 32768 / STORE_BYTES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  unsigned char* S = smem + (size_t)warp * NBUF * STORE_BYTES;
  int it = 0;
  for (int u = blockIdx.x * nwarps + warp; u < units; u += gridDim.x * nwarps) {
    // unit -> (query block, patch y, patch x); consecutive units of a CTA walk patches of the same query block
    const int px = u % patches_x, py = (u / patches_x) % patches_y, qb = u / (patches_x * patches_y);
    if (qb >= nq_blocks) break;
    for (int s = 0; s < STORES; ++s, ++it) {
      float* buf = reinterpret_cast<float*>(S + (it % NBUF) * STORE_BYTES);
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(NBUF - 1) : "memory");
      __syncwarp();
      // lane = query writes its share of the store: STORE_BYTES / 32 bytes, as 128-byte rows with the 128B swizzle
#pragma unroll
      for (int j = 0; j < STORE_BYTES / 4096; ++j)
#pragma unroll
        for (int c = 0; c < 8; ++c)
          *reinterpret_cast<float4*>(buf + j * 1024 + lane * 32 + ((c ^ (lane & 7)) * 4)) = make_float4(u, s, lane, c);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        const unsigned long long tmp = reinterpret_cast<unsigned long long>(&tm);
        if (MODE == 0) {  // store s = patch row s
          asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(tmp),
                       "r"(smem_u32(buf)), "r"(px * 32), "r"(py * 8 + s), "r"(qb * 32)
                       : "memory");
        } else if (MODE == 1) {  // store s: tile row 2*py + (s & 1), tiles 8*px + 2*(s >> 1) ..+1
          asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(tmp),
                       "r"(smem_u32(buf)), "r"(0), "r"(px * 4 + (s >> 1)), "r"(py * 2 + (s & 1)), "r"(qb * 32)
                       : "memory");
        } else {  // store s: NP tile pairs; pairs per tile row in a patch = 4
          constexpr int PER_ROW = 4 / NP;  // stores per tile row of the patch
          const int trow = py * 2 + s / PER_ROW, pair0 = px * 4 + (s % PER_ROW) * NP;
          asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(tmp),
                       "r"(smem_u32(buf)), "r"(0), "r"(qb * 32), "r"(pair0), "r"(trow)
                       : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// LSU variant of the tiled pattern: the warp writes, per query, 512 contiguous bytes with one STG.128 per lane
// (what a shared-memory transpose in the epilogue would allow)
__global__ void __launch_bounds__(1024) lsu_tiled_kernel(float* out, int units, int nq_blocks, int patches_x, int patches_y,
                                                         long long slab, int tiles_x) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int u = blockIdx.x * nwarps + warp; u < units; u += gridDim.x * nwarps) {
    const int px = u % patches_x, py = (u / patches_x) % patches_y, qb = u / (patches_x * patches_y);
    if (qb >= nq_blocks) break;
    for (int trow = 0; trow < 2; ++trow)
      for (int q = 0; q < 32; ++q) {
        float* p = out + (long long)(qb * 32 + q) * slab + ((long long)(py * 2 + trow) * tiles_x + px * 8) * 16 + lane * 4;
        asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"((float)u), "f"((float)q),
                     "f"((float)lane), "f"((float)warp)
                     : "memory");
      }
  }
}

// ---------------------------------------------------------------------------------------------------------
// (b) fetch skeleton
// ---------------------------------------------------------------------------------------------------------
struct FetchParams {
  const float* level[4];
  int lh[4], lw[4], tx[4];      // level size, tiles per tile row (tiled)
  long long slab[4];            // floats per query slab
  const float* coords;
  float* out;
  int Q;
};

__device__ __forceinline__ float4 ldg128(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void ldg256(const float* p, float* v) {
  asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
               : "l"(p));
}

// MODE 0: row-major slabs, 40 chunk slots per level (10 rows x 4 aligned 16-byte chunks), 5 LDG.128 per lane
// MODE 1: tiled slabs, same 40 chunk slots, 5 LDG.128 per lane
// MODE 2: tiled slabs, 24 half-tile slots per level (6 row pairs x 4 tile columns), 3 LDG.256 per lane
template <int MODE>
__global__ void __launch_bounds__(256) fetch_kernel(const FetchParams p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = gridDim.x * 8;
  for (int q = blockIdx.x * 8 + warp; q < p.Q; q += nwarps) {
    const float cx = __ldg(p.coords + 2 * (size_t)q), cy = __ldg(p.coords + 2 * (size_t)q + 1);
    float acc = 0.0f;
    constexpr int SLOTS = MODE == 2 ? 24 : 40;
    constexpr int ROUNDS = SLOTS * 4 / 32;
#pragma unroll
    for (int m = 0; m < ROUNDS; ++m) {
      const int slot = lane + 32 * m;
      const int l = slot / SLOTS, rem = slot % SLOTS;
      const float sc = __int_as_float((127 - l) << 23);
      const int bx = (int)floorf(cx * sc) - 4, by = (int)floorf(cy * sc) - 4;
      const int lh = p.lh[l], lw = p.lw[l];
      const float* slab = p.level[l] + (size_t)q * p.slab[l];
      if (MODE == 2) {
        const int k = rem >> 2, c = rem & 3;
        const int gy = (by & ~1) + 2 * k, gx = (bx & ~3) + 4 * c;
        if (gx <= bx + 9 && gy <= by + 9 && (unsigned)gy < (unsigned)lh && (unsigned)gx < (unsigned)lw) {
          float v[8];
          ldg256(slab + ((size_t)(gy >> 2) * p.tx[l] + (gx >> 2)) * 16 + (gy & 2) * 4, v);
          acc += v[0] + v[1] + v[2] + v[3] + v[4] + v[5] + v[6] + v[7];
        }
      } else {
        const int r = rem >> 2, c = rem & 3;
        const int gy = by + r, gx = (bx & ~3) + 4 * c;
        if (gx <= bx + 9 && (unsigned)gy < (unsigned)lh && (unsigned)gx < (unsigned)lw) {
          const float* a = MODE == 0 ? slab + (size_t)gy * lw + gx
                                     : slab + ((size_t)(gy >> 2) * p.tx[l] + (gx >> 2)) * 16 + (gy & 3) * 4;
          const float4 v = ldg128(a);
          acc += v.x + v.y + v.z + v.w;
        }
      }
    }
    // 1296-byte output row: 10 full warp stores + 4 floats
    float* o = p.out + (size_t)q * 324;
#pragma unroll
    for (int k = 0; k < 10; ++k) o[lane + 32 * k] = acc + (float)k;
    if (lane < 4) o[320 + lane] = acc;
  }
}

int main(int argc, char** argv) {
  const int h = 55, w = 128, nq = 7040;
  // argv[1]: batch of the fetch skeleton; argv[2]: store test to run (-1: all, -2: none)
  const int sel = argc > 2 ? atoi(argv[2]) : -1;
  // argv[3]: cudaLimitMaxL2FetchGranularity to request (32, 64 or 128; 0 = leave the default)
  {
    size_t g0 = 0;
    CK(cudaFree(0));
    CK(cudaDeviceGetLimit(&g0, cudaLimitMaxL2FetchGranularity));
    const int want = argc > 3 ? atoi(argv[3]) : 0;
    if (want) {
      cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)want);
      size_t g1 = 0;
      cudaDeviceGetLimit(&g1, cudaLimitMaxL2FetchGranularity);
      printf("L2 fetch granularity: default %zu, requested %d -> %s, now %zu\n", g0, want, cudaGetErrorString(e), g1);
    } else {
      printf("L2 fetch granularity: default %zu\n", g0);
    }
  }
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &qr));
  EncodeFn encode = (EncodeFn)fp;
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  float* flush;
  const size_t flush_bytes = 512u << 20;
  CK(cudaMalloc(&flush, flush_bytes));

  // ---------------- (a) stores ----------------
  if (sel != -2) {
    const int tiles_x = 32, tiles_y = 14;
    const long long slab_rm = (long long)h * w, slab_t = (long long)tiles_x * tiles_y * 16;
    float *out_rm, *out_t;
    CK(cudaMalloc(&out_rm, (size_t)nq * slab_rm * 4));
    CK(cudaMalloc(&out_t, (size_t)nq * slab_t * 4));
    CUtensorMap tm0, tm1, tm2[3];
    bool tm2_ok[3] = {true, true, true};
    {
      const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)nq};
      const cuuint64_t strides[2] = {(cuuint64_t)w * 4, (cuuint64_t)slab_rm * 4};
      const cuuint32_t box[3] = {32, 1, 32}, es[3] = {1, 1, 1};
      if (encode(&tm0, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, out_rm, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        printf("encode 0 failed\n");
        return 1;
      }
    }
    {
      const cuuint64_t dims[4] = {32, (cuuint64_t)tiles_x / 2, (cuuint64_t)tiles_y, (cuuint64_t)nq};
      const cuuint64_t strides[3] = {128, (cuuint64_t)tiles_x * 64, (cuuint64_t)slab_t * 4};
      const cuuint32_t box[4] = {32, 1, 1, 32}, es[4] = {1, 1, 1, 1};
      if (encode(&tm1, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out_t, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        printf("encode 1 failed\n");
        return 1;
      }
    }
    for (int i = 0; i < 3; ++i) {
      const int np = 1 << i;
      // (32 floats of a tile pair; query; pair within the tile row; tile row): shared-memory order [pair][query][32 floats]
      const cuuint64_t dims[4] = {32, (cuuint64_t)nq, (cuuint64_t)tiles_x / 2, (cuuint64_t)tiles_y};
      const cuuint64_t strides[3] = {(cuuint64_t)slab_t * 4, 128, (cuuint64_t)tiles_x * 64};
      const cuuint32_t box[4] = {32, 32, (cuuint32_t)np, 1}, es[4] = {1, 1, 1, 1};
      if (encode(&tm2[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out_t, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        printf("encode 2.%d failed\n", i);
        tm2_ok[i] = false;
      }
    }
    const int patches_x = 4, patches_y = 7, nq_blocks = nq / 32;
    const int units = nq_blocks * patches_x * patches_y;  // 6160 units x 32 KB = 202 MB (row-major: row 55 clipped)
    const double bytes_t = (double)units * 32768.0, bytes_rm = (double)nq * slab_rm * 4;
    int test_id = 0;
    auto time_it = [&](const char* name, double bytes, auto launch) {
      if (sel >= 0 && test_id++ != sel) return;
      float best = 1e9f;
      for (int rep = 0; rep < 5; ++rep) {
        CK(cudaMemsetAsync(flush, rep, flush_bytes));
        cudaEventRecord(a);
        launch();
        cudaEventRecord(b);
        CK(cudaEventSynchronize(b));
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
      }
      CK(cudaGetLastError());
      printf("%-58s %7.1f us  %6.2f TB/s\n", name, best * 1e3, bytes / best / 1e9);
    };
#define RUN_STORE(MODE, NP, NBUF, WARPS, TM, BYTES, LABEL)                                                            \
  {                                                                                                                  \
    constexpr int sb = (MODE == 2 ? NP * 4096 : 4096);                                                               \
    const int smem = WARPS * NBUF * sb;                                                                              \
    if (smem <= 224 * 1024) {                                                                                        \
      CK(cudaFuncSetAttribute(store_kernel<MODE, NP, NBUF>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));     \
      char nm[128];                                                                                                  \
      snprintf(nm, sizeof nm, "%s warps/SM %d, ring %d x %d KB", LABEL, WARPS, NBUF, sb / 1024);                     \
      time_it(nm, BYTES, [&] { store_kernel<MODE, NP, NBUF><<<148, WARPS * 32, smem>>>(TM, units, nq_blocks, patches_x, patches_y); }); \
    }                                                                                                                \
  }
    RUN_STORE(0, 1, 1, 8, tm0, bytes_rm, "TMA row-major 128 B rows          ")
    RUN_STORE(0, 1, 2, 8, tm0, bytes_rm, "TMA row-major 128 B rows          ")
    RUN_STORE(1, 1, 1, 8, tm1, bytes_t, "TMA tiled box{32,1,1,32q}         ")
    RUN_STORE(1, 1, 2, 8, tm1, bytes_t, "TMA tiled box{32,1,1,32q}         ")
    RUN_STORE(1, 1, 4, 8, tm1, bytes_t, "TMA tiled box{32,1,1,32q}         ")
    if (tm2_ok[0]) RUN_STORE(2, 1, 1, 8, tm2[0], bytes_t, "TMA tiled box{32,32q,1 pair}      ")
    if (tm2_ok[0]) RUN_STORE(2, 1, 2, 8, tm2[0], bytes_t, "TMA tiled box{32,32q,1 pair}      ")
    if (tm2_ok[1]) RUN_STORE(2, 2, 1, 8, tm2[1], bytes_t, "TMA tiled box{32,32q,2 pairs}     ")
    if (tm2_ok[1]) RUN_STORE(2, 2, 2, 8, tm2[1], bytes_t, "TMA tiled box{32,32q,2 pairs}     ")
    if (tm2_ok[2]) RUN_STORE(2, 4, 1, 8, tm2[2], bytes_t, "TMA tiled box{32,32q,4 pairs}     ")
    if (tm2_ok[2]) RUN_STORE(2, 4, 1, 4, tm2[2], bytes_t, "TMA tiled box{32,32q,4 pairs}     ")
    if (tm2_ok[2]) RUN_STORE(2, 4, 2, 4, tm2[2], bytes_t, "TMA tiled box{32,32q,4 pairs}     ")
    if (tm2_ok[0]) RUN_STORE(2, 1, 2, 16, tm2[0], bytes_t, "TMA tiled box{32,32q,1 pair}      ")
    for (int warps : {8, 16, 32}) {
      char nm[128];
      snprintf(nm, sizeof nm, "LSU tiled, 512 B per query per STG.128      warps/SM %d", warps);
      time_it(nm, bytes_t, [&] { lsu_tiled_kernel<<<148, warps * 32>>>(out_t, units, nq_blocks, patches_x, patches_y, slab_t, tiles_x); });
    }
    CK(cudaFree(out_rm));
    CK(cudaFree(out_t));
  }

  // ---------------- (b) fetch ----------------
  if (sel < 0) {
    const int B = argc > 1 ? atoi(argv[1]) : 16;
    const int Q = B * nq;
    FetchParams rm{}, tl{};
    float *lv_rm[4], *lv_t[4];
    for (int l = 0; l < 4; ++l) {
      const int lh = h >> l, lw = w >> l, tx = (lw + 3) / 4, ty = (lh + 3) / 4;
      rm.lh[l] = tl.lh[l] = lh;
      rm.lw[l] = tl.lw[l] = lw;
      rm.tx[l] = tl.tx[l] = tx;
      rm.slab[l] = (long long)lh * lw;
      tl.slab[l] = (long long)tx * ty * 16;
      CK(cudaMalloc(&lv_rm[l], (size_t)Q * rm.slab[l] * 4));
      CK(cudaMalloc(&lv_t[l], (size_t)Q * tl.slab[l] * 4));
      CK(cudaMemset(lv_rm[l], 0, (size_t)Q * rm.slab[l] * 4));
      CK(cudaMemset(lv_t[l], 0, (size_t)Q * tl.slab[l] * 4));
      rm.level[l] = lv_rm[l];
      tl.level[l] = lv_t[l];
    }
    std::vector<float> hc((size_t)Q * 2);
    srand(1);
    for (int q = 0; q < Q; ++q) {
      const int pix = q % nq;
      hc[2 * q] = (float)(pix % w) + 16.0f * ((float)rand() / RAND_MAX - 0.5f);
      hc[2 * q + 1] = (float)(pix / w) + 16.0f * ((float)rand() / RAND_MAX - 0.5f);
    }
    float *coords, *out;
    CK(cudaMalloc(&coords, (size_t)Q * 8));
    CK(cudaMalloc(&out, (size_t)Q * 1296));
    CK(cudaMemcpy(coords, hc.data(), (size_t)Q * 8, cudaMemcpyHostToDevice));
    rm.coords = tl.coords = coords;
    rm.out = tl.out = out;
    rm.Q = tl.Q = Q;
    const double alg = (double)Q * 2904.0;
    auto time_fetch = [&](const char* name, auto launch) {
      float best = 1e9f;
      for (int rep = 0; rep < 5; ++rep) {
        CK(cudaMemsetAsync(flush, rep, flush_bytes));
        cudaEventRecord(a);
        launch();
        cudaEventRecord(b);
        CK(cudaEventSynchronize(b));
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
      }
      CK(cudaGetLastError());
      printf("fetch B=%d %-44s %7.1f us  %6.2f TB/s algorithmic (2904 B/query)\n", B, name, best * 1e3, alg / best / 1e9);
    };
    for (int bps : {4, 8}) {
      const int blocks = 148 * bps;
      char nm[96];
      snprintf(nm, sizeof nm, "row-major, LDG.128 chunks, %d blocks/SM", bps);
      time_fetch(nm, [&] { fetch_kernel<0><<<blocks, 256>>>(rm); });
      snprintf(nm, sizeof nm, "tiled 4x4, LDG.128 chunks, %d blocks/SM", bps);
      time_fetch(nm, [&] { fetch_kernel<1><<<blocks, 256>>>(tl); });
      snprintf(nm, sizeof nm, "tiled 4x4, LDG.256 half tiles, %d blocks/SM", bps);
      time_fetch(nm, [&] { fetch_kernel<2><<<blocks, 256>>>(tl); });
    }
  }
  return 0;
}
