// Probe of the tcgen05.mma TS form (A operand in TMEM, B in shared memory) for kind::f16, M = 128:
//   nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O2 -I jax_raft_b200/csrc -I include \
//        profiles/scratch/ts_mma_probe.cu -o profiles/_build/ts_mma_probe -lcuda
// One CTA of 128 threads: thread m writes row m of A (K = 64 bf16 = 32 words) into TMEM lane m, columns 128..159, with
// tcgen05.st.32x32b.x32 (two consecutive-K bf16 per 32-bit column, even k in the low half); B (N = 16 rows, K = 64,
// K-major, 128-byte swizzle) is written into shared memory by hand; thread 0 issues four K = 16 MMAs; D (columns 0..15)
// is read back with tcgen05.ld and compared on the host with A B^T.
#include <cuda_bf16.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "tc_ptx.cuh"

using namespace raftk::ptx;

constexpr int M = 128, N = 16, K = 64, A_COL = 128;

__global__ void __launch_bounds__(128) probe(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D) {
  __shared__ __align__(1024) uint8_t sb[N * 128];
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int t = threadIdx.x, warp = t >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "r"(256));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (t == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // B: row n at n * 128 bytes, 16-byte chunk c stored at chunk c ^ (n & 7)
  for (int i = t; i < N * 8; i += 128) {
    const int n = i >> 3, c = i & 7;
    *reinterpret_cast<uint4*>(sb + n * 128 + ((c ^ (n & 7)) << 4)) = *reinterpret_cast<const uint4*>(B + n * K + c * 8);
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tslot;
  {
    uint32_t r[32];
    const uint32_t* src = reinterpret_cast<const uint32_t*>(A + t * K);
#pragma unroll
    for (int i = 0; i < 32; ++i) r[i] = src[i];
    const uint32_t taddr = tmem + ((uint32_t)(32 * warp) << 16) + A_COL;
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]),
        "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]),
        "r"(r[31])
        : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (t == 0) {
    const uint32_t idesc = make_idesc_bf16(M, N);
    const uint64_t bdesc = umma_desc(smem_u32(sb));
    for (int ks = 0; ks < K / 16; ++ks) {
      const uint32_t a_addr = tmem + A_COL + ks * 8;
      const uint64_t bd = bdesc + (uint64_t)(ks * 2);
      const uint32_t acc = ks != 0;
      asm volatile(
          "{\n\t"
          ".reg .pred p;\n\t"
          "setp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
          "}" ::"r"(tmem),
          "r"(a_addr), "l"(bd), "r"(idesc), "r"(acc)
          : "memory");
    }
    tc_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  float v[16];
  tmem_ld_32x8(tmem + ((uint32_t)(32 * warp) << 16), v);
  tmem_ld_32x8(tmem + ((uint32_t)(32 * warp) << 16) + 8, v + 8);
  tmem_ld_wait();
  for (int n = 0; n < N; ++n) D[t * N + n] = v[n];
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256));
  }
}

int main() {
  std::vector<__nv_bfloat16> hA(M * K), hB(N * K);
  std::vector<float> fA(M * K), fB(N * K);
  srand(1);
  for (int i = 0; i < M * K; ++i) {
    hA[i] = __float2bfloat16((rand() % 17 - 8) / 8.0f);
    fA[i] = __bfloat162float(hA[i]);
  }
  for (int i = 0; i < N * K; ++i) {
    hB[i] = __float2bfloat16((rand() % 13 - 6) / 4.0f);
    fB[i] = __bfloat162float(hB[i]);
  }
  __nv_bfloat16 *dA, *dB;
  float* dD;
  cudaMalloc(&dA, M * K * 2);
  cudaMalloc(&dB, N * K * 2);
  cudaMalloc(&dD, M * N * 4);
  cudaMemcpy(dA, hA.data(), M * K * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), N * K * 2, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0xff, M * N * 4);
  probe<<<1, 128>>>(dA, dB, dD);
  cudaError_t e = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(e));
  std::vector<float> hD(M * N);
  cudaMemcpy(hD.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0;
  int bad = 0;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      double ref = 0;
      for (int k = 0; k < K; ++k) ref += (double)fA[m * K + k] * fB[n * K + k];
      const double err = fabs(ref - hD[m * N + n]);
      if (err > maxerr) maxerr = err;
      if (err > 1e-3 && bad++ < 6) printf("  D[%d][%d] = %f, expected %f\n", m, n, hD[m * N + n], ref);
    }
  printf("max |D - A B^T| = %g over %d x %d (%d mismatches)\n", maxerr, M, N, bad);
  return bad ? 1 : 0;
}
