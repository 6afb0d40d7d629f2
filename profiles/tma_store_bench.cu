// Micro-benchmark behind DESIGN.md §4.1: how fast can one SM push the volume epilogue's store
// pattern (32 queries x 128-byte row segments, one per query slab) through (a) TMA bulk tensor
// stores and (b) LSU 16-byte stores, as a function of warps and staging tiles in flight?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/_build/tma_store_bench profiles/tma_store_bench.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

template <int NBUF>
__global__ void __launch_bounds__(1024) tma_store_kernel(const __grid_constant__ CUtensorMap tm, int h, int w, int nq, int rows_per_warp) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  float* S = reinterpret_cast<float*>(smem + (size_t)warp * NBUF * 4096);
  // each warp writes rows_per_warp row blocks; block id and warp pick a (query group, row, x0)
  int it = 0;
  for (int r = 0; r < rows_per_warp; ++r, ++it) {
    float* buf = S + (it % NBUF) * 1024;
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(NBUF - 1) : "memory");
    __syncwarp();
#pragma unroll
    for (int c = 0; c < 8; ++c)
      *reinterpret_cast<float4*>(buf + lane * 32 + ((c ^ (lane & 7)) * 4)) = make_float4(r, c, lane, warp);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) {
      const long long blk = ((long long)blockIdx.x * nwarps + warp) * rows_per_warp + r;
      const int x0 = (int)(blk % (w / 32)) * 32;
      const int y = (int)((blk / (w / 32)) % h);
      const int q0 = (int)((blk / ((long long)(w / 32) * h)) * 32 % nq);
      asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                       reinterpret_cast<unsigned long long>(&tm)),
                   "r"(smem_u32(buf)), "r"(x0), "r"(y), "r"(q0)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// Patch-major variant: a query's 8x32 patch is 1 KB contiguous, so a 32-query block is 32 runs of
// 1 KB (box {256 floats, 32 queries}, 32 KB staged per store; no swizzle).
__global__ void __launch_bounds__(1024) tma_store_patch_kernel(const __grid_constant__ CUtensorMap tm, int patches, int nq,
                                                             int stores_per_warp) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  float* buf = reinterpret_cast<float*>(smem + (size_t)warp * 32768);
  for (int r = 0; r < stores_per_warp; ++r) {
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    __syncwarp();
    for (int c = 0; c < 64; ++c) *reinterpret_cast<float4*>(buf + lane * 256 + ((c + lane) & 63) * 4) = make_float4(r, c, lane, warp);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) {
      const long long blk = ((long long)blockIdx.x * nwarps + warp) * stores_per_warp + r;
      const int patch = (int)(blk % patches);
      const int q0 = (int)((blk / patches) * 32 % nq);
      asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                       reinterpret_cast<unsigned long long>(&tm)),
                   "r"(smem_u32(buf)), "r"(patch * 256), "r"(q0)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

__global__ void __launch_bounds__(1024) lsu_store_kernel(float* out, int h, int w, int nq, int rows_per_warp) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const long long slab = (long long)h * w;
  for (int r = 0; r < rows_per_warp; ++r) {
    const long long blk = ((long long)blockIdx.x * nwarps + warp) * rows_per_warp + r;
    const int x0 = (int)(blk % (w / 32)) * 32;
    const int y = (int)((blk / (w / 32)) % h);
    const int q0 = (int)((blk / ((long long)(w / 32) * h)) * 32 % nq);
    float* row = out + (q0 + (lane >> 3)) * slab + (long long)y * w + x0 + 4 * (lane & 7);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(row), "f"((float)r), "f"((float)i),
                   "f"((float)lane), "f"((float)warp)
                   : "memory");
      row += 4 * slab;
    }
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int h = 55, w = 128, nq = 7040;
  float* out;
  CK(cudaMalloc(&out, (size_t)nq * h * w * 4));
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  CUtensorMap tm;
  const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)nq};
  const cuuint64_t strides[2] = {(cuuint64_t)w * 4, (cuuint64_t)h * w * 4};
  const cuuint32_t box[3] = {32, 1, 32}, es[3] = {1, 1, 1};
  if (((EncodeFn)fp)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, out, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
    printf("encode failed\n");
    return 1;
  }
  // total row blocks (32 queries x 128 B) in the level-0 volume: 7040/32 * 55 * 4 = 48400 -> 198 MB
  const long long total_blocks = (long long)(nq / 32) * h * (w / 32);
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  auto run = [&](const char* name, int warps, int nbuf, int mode) {
    const int rows_per_warp = (int)((total_blocks + 148LL * warps - 1) / (148LL * warps));
    const double bytes = 148.0 * warps * rows_per_warp * 4096.0;
    float best = 1e9f;
    for (int rep = 0; rep < 5; ++rep) {
      cudaEventRecord(a);
      if (mode == 0) {
        const size_t smem = (size_t)warps * nbuf * 4096;
        if (nbuf == 1) {
          cudaFuncSetAttribute(tma_store_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
          tma_store_kernel<1><<<148, warps * 32, smem>>>(tm, h, w, nq, rows_per_warp);
        } else if (nbuf == 2) {
          cudaFuncSetAttribute(tma_store_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
          tma_store_kernel<2><<<148, warps * 32, smem>>>(tm, h, w, nq, rows_per_warp);
        } else {
          cudaFuncSetAttribute(tma_store_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
          tma_store_kernel<4><<<148, warps * 32, smem>>>(tm, h, w, nq, rows_per_warp);
        }
      } else {
        lsu_store_kernel<<<148, warps * 32>>>(out, h, w, nq, rows_per_warp);
      }
      cudaEventRecord(b);
      CK(cudaEventSynchronize(b));
      float ms;
      cudaEventElapsedTime(&ms, a, b);
      if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    printf("%-4s warps/SM %2d  tiles in flight/warp %d : %7.1f us  %6.2f TB/s  (%.1f B/clk/SM at 1.9 GHz)\n", name, warps, nbuf,
           best * 1e3, bytes / best / 1e9, bytes / (best * 1e-3) / 148.0 / 1.9e9);
  };
  {
    // patch-major slabs: 56 x 128 padded -> 7 x 4 patches of 256 floats
    const int patches = 28;
    CUtensorMap tp;
    const cuuint64_t d2[2] = {(cuuint64_t)patches * 256, (cuuint64_t)nq};
    const cuuint64_t s2[1] = {(cuuint64_t)patches * 256 * 4};
    const cuuint32_t b2[2] = {256, 32}, e2[2] = {1, 1};
    float* out2;
    CK(cudaMalloc(&out2, (size_t)nq * patches * 256 * 4));
    if (((EncodeFn)fp)(&tp, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, out2, d2, s2, b2, e2, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
      printf("encode 2 failed\n");
      return 1;
    }
    const long long stores = (long long)(nq / 32) * patches;  // 32 KB each -> 202 MB
    for (int warps : {2, 4, 6}) {
      const int spw = (int)((stores + 148LL * warps - 1) / (148LL * warps));
      const double bytes = 148.0 * warps * spw * 32768.0;
      float best = 1e9f;
      cudaFuncSetAttribute(tma_store_patch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, warps * 32768);
      for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(a);
        tma_store_patch_kernel<<<148, warps * 32, warps * 32768>>>(tp, patches, nq, spw);
        cudaEventRecord(b);
        CK(cudaEventSynchronize(b));
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
      }
      CK(cudaGetLastError());
      printf("TMA patch-major (1 KB runs) warps/SM %d: %7.1f us  %6.2f TB/s\n", warps, best * 1e3, bytes / best / 1e9);
    }
  }
  for (int warps : {4, 8, 16}) for (int nbuf : {1, 2, 4}) if (warps * nbuf * 4096 <= 200 * 1024) run("TMA", warps, nbuf, 0);
  for (int warps : {4, 8, 16, 32}) run("LSU", warps, 0, 1);
  return 0;
}
