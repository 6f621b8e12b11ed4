// Probe: MUFU.EX2 / F2FP / FFMA issue rates per SM (one CTA, W warps, 8 independent chains per thread).
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
template <int mode>
__global__ void probe(int iters, float* sink, long long* out) {
  float v[8];
  for (int i = 0; i < 8; ++i) v[i] = threadIdx.x * 1e-3f + i;
  unsigned acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (mode == 0) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i])); }
      else if (mode == 1) { v[i] = fmaf(v[i], 1.0001f, 0.5f); }
      else if (mode == 2) { __nv_bfloat162 h = __floats2bfloat162_rn(v[i], v[(i + 1) & 7]); acc += *reinterpret_cast<unsigned*>(&h); v[i] += 1.0f; }
      else { float e; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(v[i])); v[i] = fmaf(e, 0.5f, v[i]); }
    }
  }
  const long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 8; ++i) s += v[i];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s + acc;
  if (threadIdx.x == 0) out[0] = t1 - t0;
}
int main() {
  float* sink; long long* d; cudaMalloc(&sink, 1 << 20); cudaMalloc(&d, 64);
  const char* names[] = {"MUFU.EX2 only", "FFMA only", "F2FP pack + FADD", "EX2 + FFMA"};
  for (int mode = 0; mode < 4; ++mode)
    for (int warps : {4, 8, 16, 32}) {
      const int iters = 256;
      if (mode == 0) probe<0><<<1, warps * 32, 0>>>(iters, sink, d);
      else if (mode == 1) probe<1><<<1, warps * 32, 0>>>(iters, sink, d);
      else if (mode == 2) probe<2><<<1, warps * 32, 0>>>(iters, sink, d);
      else probe<3><<<1, warps * 32, 0>>>(iters, sink, d);
      cudaDeviceSynchronize();
      long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
      const double ops = (double)warps * 32 * iters * 8;
      printf("%-18s warps=%2d: %7lld cycles -> %.1f lane-ops/clk/SM\n", names[mode], warps, h, ops / h);
    }
  return 0;
}
