// Tap-table cache for K1 / K8 (see resample_plan.h).
#include "resample_plan.h"

#include <cuda_runtime.h>

#include <map>
#include <memory>
#include <mutex>
#include <tuple>

#include "depth_b200.h"
#include "host_util.h"

namespace deb {

int AxisTaps::max_extent(int span) const {
  int best = 1;
  for (int d0 = 0; d0 < dst; d0 += span) {
    const int d1 = (d0 + span < dst ? d0 + span : dst) - 1;
    const int e = idx[d1 * 4 + 3] - idx[d0 * 4] + 1;
    if (e > best) best = e;
  }
  return best;
}

const AxisTaps* get_axis_taps(int src, int dst, int mode) {
  static std::mutex mu;
  static std::map<std::tuple<int, int, int, int>, std::unique_ptr<AxisTaps>> cache;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_tuple(dev, src, dst, mode);
  auto it = cache.find(key);
  if (it != cache.end()) return it->second.get();
  std::unique_ptr<AxisTaps> t(new AxisTaps);
  t->src = src;
  t->dst = dst;
  t->mode = mode;
  t->idx.resize((size_t)dst * 4);
  t->w.resize((size_t)dst * 4);
  if (de_cubic_taps(src, dst, mode, t->idx.data(), t->w.data()) != DE_OK) return nullptr;
  cudaError_t e = cudaMalloc(&t->d_idx, sizeof(int32_t) * dst * 4);
  if (e == cudaSuccess) e = cudaMalloc(&t->d_w, sizeof(float) * dst * 4);
  if (e == cudaSuccess) e = cudaMemcpy(t->d_idx, t->idx.data(), sizeof(int32_t) * dst * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(t->d_w, t->w.data(), sizeof(float) * dst * 4, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    cuda_fail(e, "get_axis_taps");
    return nullptr;
  }
  const AxisTaps* out = t.get();
  cache[key] = std::move(t);
  return out;
}

}  // namespace deb
