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

typedef CUresult (*TensorMapEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                           const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                           CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                           CUtensorMapFloatOOBfill);
TensorMapEncodeTiledFn get_tensor_map_encoder();

}  // namespace deb
