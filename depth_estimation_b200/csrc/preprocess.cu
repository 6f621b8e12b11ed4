// K1 - fused MiDaS input transform (replaces /root/reference/depth_midas.py:22-23):
//   uint8 BGR HWC frame -> (BGR->RGB) -> /255 -> cv2.INTER_CUBIC resize to (net_h, net_w) ->
//   (x-mean)/std -> bf16 NCHW (or the ViT patch matrix, or fp32 NCHW for parity checks).
// One CTA per output tile: (1) the source window is staged in shared memory with 16-byte coalesced
// loads of the raw byte rows, (2) horizontal 4-tap pass into an fp32 smem buffer, (3) vertical 4-tap
// pass + normalisation + packed stores.  The frame is read once and the tensor written once.
#include <algorithm>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"
#include "resample_plan.h"

namespace deb {

struct PreParams {
  const uint8_t* src;
  long long row_stride, frame_stride, total_bytes;
  int H, W, net_h, net_w;
  const int32_t* xidx;
  const float* xw;
  const int32_t* yidx;
  const float* yw;
  float scale[3], shift[3];  // RGB order: out = v*scale - shift
  int layout;
  void* out;
  int TW, TH, pitch, max_rows;
};

constexpr int kPreThreads = 256;

__global__ void __launch_bounds__(kPreThreads) preprocess_kernel(const PreParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* s_src = smem;                                                     // [max_rows][pitch]
  float* s_h = reinterpret_cast<float*>(smem + (size_t)p.max_rows * p.pitch);  // [max_rows][TW][3]
  const int tid = threadIdx.x;
  const int b = blockIdx.z;
  const int ox0 = blockIdx.x * p.TW, oy0 = blockIdx.y * p.TH;
  const int ncols = min(p.TW, p.net_w - ox0), nrows = min(p.TH, p.net_h - oy0);
  const int xs = __ldg(p.xidx + ox0 * 4), xe = __ldg(p.xidx + (ox0 + ncols - 1) * 4 + 3);
  const int ys = __ldg(p.yidx + oy0 * 4), ye = __ldg(p.yidx + (oy0 + nrows - 1) * 4 + 3);
  const int nr = ye - ys + 1;
  const int nb = (xe - xs + 1) * 3;
  const uint8_t* frame = p.src + (long long)b * p.frame_stride;
  const uint8_t* buf_end = p.src + p.total_bytes;

  // ---- phase 1: stage source bytes; smem row rr holds bytes starting at offset (gaddr & 15)
  const int nvec_max = p.pitch / 16;
  for (int i = tid; i < nr * nvec_max; i += kPreThreads) {
    const int rr = i / nvec_max, v = i % nvec_max;
    const uint8_t* g = frame + (long long)(ys + rr) * p.row_stride + (long long)xs * 3;
    const int mis = static_cast<int>(reinterpret_cast<uintptr_t>(g) & 15);
    if (v * 16 >= mis + nb) continue;
    const uint8_t* ga = g - mis + v * 16;
    uint8_t* sa = s_src + rr * p.pitch + v * 16;
    if (ga >= p.src && ga + 16 <= buf_end) {
      *reinterpret_cast<uint4*>(sa) = __ldg(reinterpret_cast<const uint4*>(ga));
    } else {
      for (int j = 0; j < 16; ++j)
        if (ga + j >= p.src && ga + j < buf_end) sa[j] = ga[j];
    }
  }
  __syncthreads();

  // ---- phase 2: horizontal taps (OpenCV HResizeCubic order), BGR -> RGB
  for (int i = tid; i < nr * ncols; i += kPreThreads) {
    const int rr = i / ncols, c = i % ncols;
    const uint8_t* g = frame + (long long)(ys + rr) * p.row_stride + (long long)xs * 3;
    const int mis = static_cast<int>(reinterpret_cast<uintptr_t>(g) & 15);
    const uint8_t* row = s_src + rr * p.pitch + mis;
    const int4 xi = __ldg(reinterpret_cast<const int4*>(p.xidx) + ox0 + c);
    const float4 w = __ldg(reinterpret_cast<const float4*>(p.xw) + ox0 + c);
    const uint8_t* p0 = row + (xi.x - xs) * 3;
    const uint8_t* p1 = row + (xi.y - xs) * 3;
    const uint8_t* p2 = row + (xi.z - xs) * 3;
    const uint8_t* p3 = row + (xi.w - xs) * 3;
    float* h = s_h + (rr * p.TW + c) * 3;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      float acc = (float)p0[ch] * w.x;
      acc = fmaf((float)p1[ch], w.y, acc);
      acc = fmaf((float)p2[ch], w.z, acc);
      acc = fmaf((float)p3[ch], w.w, acc);
      h[2 - ch] = acc;
    }
  }
  __syncthreads();

  // ---- phase 3: vertical taps + normalise + store, two adjacent columns per thread
  const int cpairs = ncols >> 1;
  const int gw = p.net_w >> 4, gh = p.net_h >> 4;
  for (int i = tid; i < 3 * nrows * cpairs; i += kPreThreads) {
    const int cp = i % cpairs, ry = (i / cpairs) % nrows, ch = i / (cpairs * nrows);
    const int oy = oy0 + ry;
    const int4 yi = __ldg(reinterpret_cast<const int4*>(p.yidx) + oy);
    const float4 w = __ldg(reinterpret_cast<const float4*>(p.yw) + oy);
    float v[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int c = cp * 2 + e;
      float acc = s_h[((yi.x - ys) * p.TW + c) * 3 + ch] * w.x;
      acc = fmaf(s_h[((yi.y - ys) * p.TW + c) * 3 + ch], w.y, acc);
      acc = fmaf(s_h[((yi.z - ys) * p.TW + c) * 3 + ch], w.z, acc);
      acc = fmaf(s_h[((yi.w - ys) * p.TW + c) * 3 + ch], w.w, acc);
      v[e] = fmaf(acc, p.scale[ch], -p.shift[ch]);
    }
    const int ox = ox0 + cp * 2;
    if (p.layout == 0) {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (((long long)b * 3 + ch) * p.net_h + oy) * p.net_w + ox;
      *reinterpret_cast<uint32_t*>(o) = pack_bf16(v[0], v[1]);
    } else if (p.layout == 1) {
      const long long prow = ((long long)b * gh + (oy >> 4)) * gw + (ox >> 4);
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + prow * 768 + ch * 256 + (oy & 15) * 16 + (ox & 15);
      *reinterpret_cast<uint32_t*>(o) = pack_bf16(v[0], v[1]);
    } else {
      float* o = reinterpret_cast<float*>(p.out) + (((long long)b * 3 + ch) * p.net_h + oy) * p.net_w + ox;
      *reinterpret_cast<float2*>(o) = make_float2(v[0], v[1]);
    }
  }
}

}  // namespace deb

extern "C" int de_preprocess(const uint8_t* bgr, int B, int H, int W, int64_t row_stride, int64_t frame_stride,
                             int net_h, int net_w, const float* mean_host, const float* std_host, int layout,
                             void* out, void* stream) {
  using namespace deb;
  if (!bgr || !out || B <= 0 || H <= 0 || W <= 0) return fail(DE_ERR_INVALID, "de_preprocess: bad args");
  if (net_h % 32 != 0 || net_w % 32 != 0 || net_h <= 0 || net_w <= 0)
    return fail(DE_ERR_INVALID, "de_preprocess: net size %dx%d is not a positive multiple of 32", net_h, net_w);
  if (layout < 0 || layout > 2) return fail(DE_ERR_INVALID, "de_preprocess: layout %d", layout);
  if (row_stride < (int64_t)W * 3 || frame_stride < row_stride * H) return fail(DE_ERR_INVALID, "de_preprocess: strides");
  const AxisTaps* tx = get_axis_taps(W, net_w, 0);
  const AxisTaps* ty = get_axis_taps(H, net_h, 0);
  if (!tx || !ty) return DE_ERR_CUDA;
  PreParams p{};
  p.src = bgr;
  p.row_stride = row_stride;
  p.frame_stride = frame_stride;
  p.total_bytes = (long long)frame_stride * (B - 1) + (long long)row_stride * (H - 1) + (long long)W * 3;
  p.H = H; p.W = W; p.net_h = net_h; p.net_w = net_w;
  p.xidx = tx->d_idx; p.xw = tx->d_w; p.yidx = ty->d_idx; p.yw = ty->d_w;
  for (int c = 0; c < 3; ++c) {
    p.scale[c] = (float)(1.0 / (255.0 * (double)std_host[c]));
    p.shift[c] = (float)((double)mean_host[c] / (double)std_host[c]);
  }
  p.layout = layout;
  p.out = out;
  // tile shape: widest tile whose staged window fits in shared memory
  static const int cand[][2] = {{64, 16}, {64, 8}, {32, 8}, {32, 4}, {16, 4}, {8, 2}, {2, 1}};
  size_t smem = 0;
  bool ok = false;
  for (auto& c : cand) {
    p.TW = c[0];
    p.TH = c[1];
    if (layout == 1 && p.TH < 16 && c[0] == 64 && c[1] == 8) { /* fine, just smaller runs */ }
    const int ext_x = tx->max_extent(p.TW), ext_y = ty->max_extent(p.TH);
    p.pitch = ((ext_x * 3 + 15) / 16) * 16 + 32;
    p.max_rows = ext_y;
    smem = (size_t)p.max_rows * p.pitch + (size_t)p.max_rows * p.TW * 3 * sizeof(float);
    if (smem <= 200 * 1024) {
      ok = true;
      break;
    }
  }
  if (!ok) return fail(DE_ERR_UNSUPPORTED, "de_preprocess: downscale factor too large for the staged window");
  if (smem > 48 * 1024) {
    static size_t configured = 0;
    if (smem > configured) {
      cudaError_t e = cudaFuncSetAttribute(preprocess_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(preprocess)");
      configured = 200 * 1024;
    }
  }
  dim3 grid((net_w + p.TW - 1) / p.TW, (net_h + p.TH - 1) / p.TH, B);
  preprocess_kernel<<<grid, kPreThreads, smem, reinterpret_cast<cudaStream_t>(stream)>>>(p);
  count_launch();
  return cuda_check_launch("preprocess_kernel");
}
