// Host-side helpers shared by the C-ABI translation units: error reporting, launch accounting,
// device queries and the driver entry point for tensor-map encoding (no link-time libcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace deb {

void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
int cuda_check_launch(const char* what);
void count_launch();
int num_sms();
// programmatic dependent launch per kernel class (1 GEMM, 2 attention / head, 4 LayerNorm, 8 depthwise); DEB_PDL=<mask>,
// default 11
bool pdl_enabled(int cls);

// Optional per-operator timing (de_profile_start / de_profile_stop): when enabled, every operator entry point brackets
// its launches with CUDA events on the stream it was given.  Off by default; not usable during stream capture.
struct ProfScope {
  int idx = -1;
  void* stream = nullptr;
  ProfScope(const char* name, double flops, void* stream, int64_t d0 = 0, int64_t d1 = 0, int64_t d2 = 0, int64_t d3 = 0);
  ~ProfScope();
};

typedef CUresult (*TensorMapEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                           const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                           CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                           CUtensorMapFloatOOBfill);
TensorMapEncodeTiledFn get_tensor_map_encoder();

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a PER-DEVICE setting: remember, per kernel and device, the largest
// opt-in made so far (one static SmemOptIn per kernel instantiation).
struct SmemOptIn {
  int bytes[32] = {0};
};
template <typename K>
inline cudaError_t ensure_smem(SmemOptIn& s, K kernel, int bytes) {
  int dev = 0;
  cudaGetDevice(&dev);
  const bool cached = dev >= 0 && dev < 32;
  if (cached && s.bytes[dev] >= bytes) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess && cached) s.bytes[dev] = bytes;
  return e;
}

// Launch with programmatic stream serialization (see common.cuh pdl_wait / pdl_launch).  Only for kernels that call
// pdl_wait() before their first global-memory access.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int cls, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled(cls) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace deb
