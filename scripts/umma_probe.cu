// Probe: how many cycles does ONE tcgen05.mma (kind::f16, cta_group::1, M=128, K=16) occupy the tensor pipe as a
// function of N, operand source (A in smem vs TMEM) and B major-ness?  One CTA, one issuing thread, `reps`
// back-to-back dependent MMAs into the same accumulator, clock64 around issue .. commit .. mbarrier wait.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../depth_estimation_b200/csrc -I../include
//        scripts/umma_probe.cu -o /tmp/umma_probe   (smem holds garbage; only the timing matters)
#include <cstdio>
#include <cuda_runtime.h>

#include "common.cuh"

using namespace deb;

__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a_tmem), "l"(bdesc),
               "r"(idesc), "r"(acc) : "memory");
}

// mode 0: SS, B K-major.  1: SS, B MN-major.  2: TS, B K-major.  3: TS, B MN-major.
template <int NACC>
__global__ void __launch_bounds__(128, 1) probe(int N, int mode, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  if (warp == 0) {
    tmem_alloc(&slot, 512);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 1) {
    const uint32_t idesc = make_idesc_bf16(128, N) | ((mode & 1) ? (1u << 16) : 0u);
    const uint64_t adesc = make_sw128_kmajor_desc(smem_u32(smem));
    const uint64_t bdesc = (mode & 1) ? make_sw128_mnmajor_desc(smem_u32(smem + 32768))
                                      : make_sw128_kmajor_desc(smem_u32(smem + 32768));
    uint32_t phase = 0;
    for (int round = 0; round < 3; ++round) {
      const long long t0 = clock64();
      if (elect_one()) {
      if (mode >= 2) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
          asm volatile("tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, 1;" ::"r"(tm + (i % NACC) * 64),
                       "r"(tm + 256 + (i & 3) * 8), "l"(bdesc + 2 * (i & 3)), "r"(idesc) : "memory");
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i)
          asm volatile("tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, 1;" ::"r"(tm + (i % NACC) * 64),
                       "l"(adesc + 2 * (i & 3)), "l"(bdesc + 2 * (i & 3)), "r"(idesc) : "memory");
      }
      }
      __syncwarp();
      const long long t1 = clock64();
      if (elect_one()) umma_commit(&bar);
      __syncwarp();
      mbar_wait(&bar, phase);
      phase ^= 1;
      const long long t2 = clock64();
      if (threadIdx.x == 32) {
        out[round * 2] = t1 - t0;
        out[round * 2 + 1] = t2 - t0;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tm, 512);
  }
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
  cudaFuncSetAttribute(probe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
  cudaFuncSetAttribute(probe<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
  const char* names[4] = {"SS  B K-major ", "SS  B MN-major", "TS  B K-major ", "TS  B MN-major"};
  const int Ns[] = {64, 128, 256};
  const int naccs[] = {1, 2, 4};
  for (int mode = 0; mode < 4; mode += 1)
    for (int N : Ns) {
      for (int nacc : naccs) {
        if (N == 256 && nacc > 1) continue;
        if ((mode & 1) && N > 64) continue;
        if (N == 128 && nacc > 2) continue;
        const int reps = 16;
        if (nacc == 1) probe<1><<<1, 128, 96 * 1024>>>(N, mode, d);
        else if (nacc == 2) probe<2><<<1, 128, 96 * 1024>>>(N, mode, d);
        else probe<4><<<1, 128, 96 * 1024>>>(N, mode, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[6];
        cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        printf("%s N=%3d nacc=%d reps=%2d: issue %5lld cyc, complete %5lld cyc -> %6.1f cyc/MMA (floor %d)%s\n", names[mode], N,
               nacc, reps, h[4], h[5], (double)h[5] / reps, N / 2, e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
    }
  return 0;
}
