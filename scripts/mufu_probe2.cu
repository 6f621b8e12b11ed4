// Probe: does the fp32x2 -> bf16x2 pack (F2FP) share the MUFU pipe?  One CTA, W warps, every warp times itself and
// the slowest warp is reported.  16 independent element slots per thread, consumers always use results that are a
// full iteration (16 MUFU issues) old, accumulators are per slot - no dependency chain shorter than an iteration.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 scripts/mufu_probe2.cu -o scripts/_bin/mufu_probe2
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
template <int mode>
__global__ void probe(int iters, float* sink, unsigned long long* out) {
  float v[16], e[16];
  unsigned acc[8];
  float s[16];
  for (int i = 0; i < 16; ++i) { v[i] = -(threadIdx.x * 1e-3f + i * 0.01f); e[i] = 0.5f; s[i] = 0.f; }
  for (int i = 0; i < 8; ++i) acc[i] = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      // consumers of the PREVIOUS iteration's exponentials
      if (mode == 1 || mode == 3) {
        __nv_bfloat162 h = __floats2bfloat162_rn(e[i], e[i + 1]);
        acc[i >> 1] += *reinterpret_cast<unsigned*>(&h);
      }
      if (mode == 2 || mode == 4) acc[i >> 1] += __byte_perm(__float_as_uint(e[i]) + 0x8000u, __float_as_uint(e[i + 1]) + 0x8000u, 0x7632);
      if (mode >= 3) { s[i] += e[i]; s[i + 1] += e[i + 1]; }
      float a0 = v[i], a1 = v[i + 1];
      if (mode >= 3) { a0 = fmaf(v[i], 1.0001f, -0.5f); a1 = fmaf(v[i + 1], 1.0001f, -0.5f); }
      asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e[i]) : "f"(a0));
      asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e[i + 1]) : "f"(a1));
    }
  }
  const long long t1 = clock64();
  float r = 0;
  for (int i = 0; i < 16; ++i) r += v[i] + e[i] + s[i];
  for (int i = 0; i < 8; ++i) r += acc[i];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = r;
  atomicMax(out, (unsigned long long)(t1 - t0));
}
template <int mode>
void run(const char* name, float* sink, unsigned long long* d) {
  for (int warps : {4, 8}) {
    const int iters = 256;
    cudaMemset(d, 0, 8);
    probe<mode><<<1, warps * 32, 0>>>(iters, sink, d);
    cudaDeviceSynchronize();
    unsigned long long h;
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    const double el = (double)iters * 16;
    printf("%-40s warps=%d: %7llu cycles -> %.2f cycles per element per scheduler\n", name, warps, h, (double)h / el / (warps / 4));
  }
}
int main() {
  float* sink; unsigned long long* d;
  cudaMalloc(&sink, 1 << 20); cudaMalloc(&d, 64);
  run<0>("MUFU.EX2 only", sink, d);
  run<1>("MUFU + F2FP pack", sink, d);
  run<2>("MUFU + integer pack", sink, d);
  run<3>("FFMA + MUFU + FADD + F2FP pack", sink, d);
  run<4>("FFMA + MUFU + FADD + integer pack", sink, d);
  return 0;
}
