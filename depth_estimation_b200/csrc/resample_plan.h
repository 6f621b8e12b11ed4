// Cached per-shape resampling plans (tap tables on host + device, tile shape, smem extents) shared by
// the fused pre-processing (K1) and post-processing (K8) kernels.
#pragma once
#include <stdint.h>

#include <vector>

namespace deb {

struct AxisTaps {
  int src = 0, dst = 0, mode = 0;
  std::vector<int32_t> idx;  // [dst*4] border-clamped
  std::vector<float> w;      // [dst*4]
  int32_t* d_idx = nullptr;
  float* d_w = nullptr;
  // largest number of source samples touched by `span` consecutive destination samples
  int max_extent(int span) const;
};

// mode 0 = OpenCV INTER_CUBIC rule, 1 = ATen bicubic align_corners=False rule.
// Returns nullptr (and sets the error string) on CUDA failure.  Thread-safe; entries live forever.
const AxisTaps* get_axis_taps(int src, int dst, int mode);

}  // namespace deb
