// HBM bandwidth as a function of the read:write mix.  K8 (bicubic up-resize) and the x2 bilinear upsample write 4-8x
// more than they read, K1 reads 4x more than it writes; the copy figure in MEASURED_PEAKS.json is a 1:1 mix.
// Each thread streams 16-byte vectors: per item `R` loads (from R disjoint regions) and `Wn` stores.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 scripts/hbm_rw_probe.cu -o scripts/_bin/hbm_rw_probe
#include <cstdio>
#include <cuda_runtime.h>

template <int R, int WN>
__global__ void __launch_bounds__(256) rw_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t items,
                                                 unsigned* sink) {
  unsigned acc = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < items; i += (size_t)gridDim.x * blockDim.x) {
    uint4 v = make_uint4((unsigned)i, 1, 2, 3);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint4 t = __ldcs(src + r * items + i);
      v.x ^= t.x; v.y += t.y; v.z ^= t.z; v.w += t.w;
    }
    if (WN == 0) acc += v.x + v.y + v.z + v.w;
#pragma unroll
    for (int w = 0; w < WN; ++w) __stcs(dst + w * items + i, v);
  }
  if (WN == 0 && acc == 0x12345678u) *sink = acc;
}

template <int R, int WN>
void run(const char* name, const uint4* src, uint4* dst, size_t total_vecs, unsigned* sink) {
  const size_t items = total_vecs / (R > WN ? R : WN);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e9f;
  for (int it = 0; it < 8; ++it) {
    cudaEventRecord(e0);
    rw_kernel<R, WN><<<148 * 16, 256>>>(src, dst, items, sink);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (it >= 2 && ms < best) best = ms;
  }
  const double bytes = (double)items * 16 * (R + WN);
  printf("{\"mix\": \"%s\", \"reads\": %d, \"writes\": %d, \"GB\": %.2f, \"ms\": %.3f, \"GBps\": %.0f}\n", name, R, WN, bytes / 1e9,
         best, bytes / best / 1e6);
}

int main() {
  const size_t total_vecs = (size_t)1 << 27;  // 2 GiB per side
  uint4 *src, *dst;
  unsigned* sink;
  cudaMalloc(&src, total_vecs * 16);
  cudaMalloc(&dst, total_vecs * 16);
  cudaMalloc(&sink, 4);
  cudaMemset(src, 1, total_vecs * 16);
  cudaMemset(dst, 0, total_vecs * 16);
  run<1, 0>("read only", src, dst, total_vecs, sink);
  run<0, 1>("write only", src, dst, total_vecs, sink);
  run<1, 1>("copy 1:1", src, dst, total_vecs, sink);
  run<2, 1>("2:1", src, dst, total_vecs, sink);
  run<4, 1>("4:1 (K1-like)", src, dst, total_vecs, sink);
  run<1, 2>("1:2 (K1 identity)", src, dst, total_vecs, sink);
  run<1, 4>("1:4 (upsample-like)", src, dst, total_vecs, sink);
  run<1, 8>("1:8 (K8-like)", src, dst, total_vecs, sink);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
