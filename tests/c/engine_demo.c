/* A host with no Python: drives the handle-level C ABI of libdepthb200.so (include/depth_b200.h) end to end.
 * What a maintainer of a C/C++/Go/Rust port of juhoward/depth_estimation would write in place of
 * DepthEstimator.__init__ + predict (/root/reference/depth_midas.py:7-35).
 *
 *   engine_demo <model_type> <weights.bin> <frames.bin> <out.bin> [reps]
 *
 * weights.bin: int32 count, then per tensor: int32 key_len, key bytes, int32 ndim, int64 shape[ndim], float32 data.
 * frames.bin:  int32 B, H, W, then B*H*W*3 uint8 BGR.
 * out.bin:     float32 [B,H,W] relative inverse depth (the array predict returns), written after `reps` predictions
 *              (the first plans + captures the CUDA graph, the rest replay it).
 * tests/test_engine_gpu.py builds the inputs, runs this binary and compares out.bin with DepthEstimator.predict. */
#include <cuda_runtime_api.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "depth_b200.h"

#define CHECK(call)                                                              \
  do {                                                                           \
    int rc_ = (call);                                                            \
    if (rc_ != DE_OK) {                                                          \
      fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, de_last_error());      \
      return 1;                                                                  \
    }                                                                            \
  } while (0)
#define CUDA(call)                                                               \
  do {                                                                           \
    cudaError_t e_ = (call);                                                     \
    if (e_ != cudaSuccess) {                                                     \
      fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_));                \
      return 1;                                                                  \
    }                                                                            \
  } while (0)

static int read_exact(FILE* f, void* p, size_t n) { return fread(p, 1, n, f) == n ? 0 : -1; }

int main(int argc, char** argv) {
  if (argc < 5) {
    fprintf(stderr, "usage: %s model_type weights.bin frames.bin out.bin [reps]\n", argv[0]);
    return 2;
  }
  const int reps = argc > 5 ? atoi(argv[5]) : 2;
  de_handle h = NULL;
  CHECK(de_create(argv[1], 0, DE_PRECISION_DEFAULT, &h));

  FILE* f = fopen(argv[2], "rb");
  if (!f) { perror(argv[2]); return 1; }
  int32_t count = 0;
  if (read_exact(f, &count, 4)) return 1;
  for (int i = 0; i < count; ++i) {
    int32_t klen = 0, ndim = 0;
    char key[512];
    int64_t shape[8], n = 1;
    if (read_exact(f, &klen, 4) || klen <= 0 || klen >= (int)sizeof(key) || read_exact(f, key, (size_t)klen)) return 1;
    key[klen] = 0;
    if (read_exact(f, &ndim, 4) || ndim < 0 || ndim > 8 || read_exact(f, shape, sizeof(int64_t) * (size_t)ndim)) return 1;
    for (int d = 0; d < ndim; ++d) n *= shape[d];
    float* data = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
    if (!data || read_exact(f, data, sizeof(float) * (size_t)n)) return 1;
    CHECK(de_set_weight(h, key, data, shape, ndim, DE_DTYPE_F32));
    free(data);
  }
  fclose(f);
  CHECK(de_finalize_weights(h));

  f = fopen(argv[3], "rb");
  if (!f) { perror(argv[3]); return 1; }
  int32_t dims[3];
  if (read_exact(f, dims, 12)) return 1;
  const int B = dims[0], H = dims[1], W = dims[2];
  const size_t in_bytes = (size_t)B * H * W * 3, out_elems = (size_t)B * H * W;
  uint8_t* frames = (uint8_t*)malloc(in_bytes);
  if (!frames || read_exact(f, frames, in_bytes)) return 1;
  fclose(f);

  uint8_t* d_frames = NULL;
  float* d_map = NULL;
  cudaStream_t st;
  CUDA(cudaSetDevice(0));
  CUDA(cudaStreamCreate(&st));
  CUDA(cudaMalloc((void**)&d_frames, in_bytes));
  CUDA(cudaMalloc((void**)&d_map, out_elems * sizeof(float)));
  int net_h = 0, net_w = 0;
  CHECK(de_plan(h, B, H, W, &net_h, &net_w));
  DePredictIO io;
  memset(&io, 0, sizeof(io));
  io.frames = d_frames;
  io.out_map = d_map;
  float* map = (float*)malloc(out_elems * sizeof(float));
  for (int r = 0; r < reps; ++r) {                     /* frame in -> map out, like depth_midas.py:21-35 */
    CUDA(cudaMemcpyAsync(d_frames, frames, in_bytes, cudaMemcpyHostToDevice, st));
    CHECK(de_predict(h, B, H, W, &io, /*use_graph=*/1, st));
    CUDA(cudaMemcpyAsync(map, d_map, out_elems * sizeof(float), cudaMemcpyDeviceToHost, st));
    CUDA(cudaStreamSynchronize(st));
  }
  f = fopen(argv[4], "wb");
  if (!f || fwrite(map, sizeof(float), out_elems, f) != out_elems) { perror(argv[4]); return 1; }
  fclose(f);
  printf("%s: %d frame(s) %dx%d -> net %dx%d, %lld kernel launches, map[0]=%g\n", argv[1], B, W, H, net_w, net_h,
         (long long)de_launch_count(), map[0]);
  de_destroy(h);
  cudaFree(d_frames);
  cudaFree(d_map);
  free(frames);
  free(map);
  return 0;
}
