// K8 - fused output stage (replaces /root/reference/depth_midas.py:28-33 and the metric tail of
// /root/reference/face.py:106-108,137):
//   prediction [B,h,w] -> bicubic (Keys A=-0.75, align_corners=False, ATen fp32 coordinate rule) to the
//   source resolution [B,H,W] fp32, optionally also the bf16 "metric" map D*metric_scale (=D/2.54) and
//   the interpolated values at a few (row,col) sample points (the iris centres).
// One CTA per output tile: horizontal pass from global into smem, vertical pass, 16-byte streaming stores.
#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"
#include "resample_plan.h"

namespace deb {

struct PostParams {
  const void* src;
  int in_dtype;
  int h, w, H, W;
  const int32_t* xidx;
  const float* xw;
  const int32_t* yidx;
  const float* yw;
  float* out_map;
  __nv_bfloat16* out_metric;
  uint8_t* out_vis;
  float metric_scale;
  int TW, TH, max_rows, max_cols;
};

constexpr int kPostThreads = 256;

DEV float load_pred(const void* src, int in_dtype, long long i) {
  if (in_dtype == DE_DTYPE_F32) return __ldg(reinterpret_cast<const float*>(src) + i);
  return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(src)[i]);
}
// The reference's to_video_frame (midas_mediapipe.py:141-150 / depth_midas.py:83): img.astype(np.uint8), then the
// in-place uint8 `*= 5`, `+= 10` - both wrap modulo 256.  numpy's float32 -> uint8 cast on x86-64 truncates toward
// zero to int32 and keeps the low byte (NaN / out-of-int32-range give 0x80000000 -> 0).
DEV uint32_t vis_byte(float v) {
  int t;
  asm("cvt.rzi.s32.f32 %0, %1;" : "=r"(t) : "f"(v));
  if (!(fabsf(v) < 2147483648.f)) t = static_cast<int>(0x80000000u);   // cvttss2si's "integer indefinite"
  return (static_cast<uint32_t>(t) * 5u + 10u) & 0xffu;
}
// fixed evaluation order so the tile kernel and the sample kernel agree bit for bit
DEV float dot4(float a, float b, float c, float d, float4 w) {
  float acc = a * w.x;
  acc = fmaf(b, w.y, acc);
  acc = fmaf(c, w.z, acc);
  acc = fmaf(d, w.w, acc);
  return acc;
}

// Tile = 128 output columns x TH output rows, 256 threads.
//  phase A (horizontal): thread column c = tid & 127 is FIXED, so its 4 tap indices / weights live in registers for
//    the whole tile; the 4 taps of every needed source row come straight from global memory (the prediction is small
//    and L2/L1 resident - no staging copy), result into s_h[row][c].
//  phase B (vertical): a warp owns one output row at a time (row taps are warp-uniform -> broadcast loads), each
//    lane 4 adjacent columns: 4 conflict-free LDS.128, 16 FMA, one 16-byte streaming store (the warp writes 512
//    contiguous bytes).  The output is written once and never re-read: st.global.cs keeps it from evicting the taps.
// No integer division anywhere; the arithmetic order (dot4 horizontally, then dot4 vertically) is unchanged, so the
// map is bit-identical to the sample kernel below.
constexpr int kPostTW = 128;

__global__ void __launch_bounds__(kPostThreads) postprocess_kernel(const PostParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  float* s_h = reinterpret_cast<float*>(smem);          // [max_rows][128]
  const int tid = threadIdx.x;
  const int b = blockIdx.z;
  const int ox0 = blockIdx.x * kPostTW, oy0 = blockIdx.y * p.TH;
  const int ncols = min(kPostTW, p.W - ox0), nrows = min(p.TH, p.H - oy0);
  const int ys = __ldg(p.yidx + oy0 * 4), ye = __ldg(p.yidx + (oy0 + nrows - 1) * 4 + 3);
  const int nr = ye - ys + 1;
  const long long fbase = (long long)b * p.h * p.w;

  {
    const int c = tid & (kPostTW - 1);
    const int cx = min(ox0 + c, p.W - 1);               // columns past the edge recompute the last one (never stored)
    const int4 xi = __ldg(reinterpret_cast<const int4*>(p.xidx) + cx);
    const float4 w = __ldg(reinterpret_cast<const float4*>(p.xw) + cx);
    if (p.in_dtype == DE_DTYPE_F32) {
      const float* src = reinterpret_cast<const float*>(p.src) + fbase + (long long)ys * p.w;
      for (int rr = tid >> 7; rr < nr; rr += kPostThreads / kPostTW) {
        const float* row = src + (long long)rr * p.w;
        s_h[rr * kPostTW + c] = dot4(__ldg(row + xi.x), __ldg(row + xi.y), __ldg(row + xi.z), __ldg(row + xi.w), w);
      }
    } else {
      const __nv_bfloat16* src = reinterpret_cast<const __nv_bfloat16*>(p.src) + fbase + (long long)ys * p.w;
      for (int rr = tid >> 7; rr < nr; rr += kPostThreads / kPostTW) {
        const __nv_bfloat16* row = src + (long long)rr * p.w;
        s_h[rr * kPostTW + c] = dot4(__bfloat162float(row[xi.x]), __bfloat162float(row[xi.y]),
                                     __bfloat162float(row[xi.z]), __bfloat162float(row[xi.w]), w);
      }
    }
  }
  __syncthreads();

  const int cq = tid & 31, c0 = cq * 4;
  if (c0 >= ncols) return;
  const bool quad_in = c0 + 4 <= ncols;
  for (int ry = tid >> 5; ry < nrows; ry += kPostThreads / 32) {
    const int oy = oy0 + ry;
    const int4 yi = __ldg(reinterpret_cast<const int4*>(p.yidx) + oy);
    const float4 w = __ldg(reinterpret_cast<const float4*>(p.yw) + oy);
    const float4 a0 = *reinterpret_cast<const float4*>(s_h + (yi.x - ys) * kPostTW + c0);
    const float4 a1 = *reinterpret_cast<const float4*>(s_h + (yi.y - ys) * kPostTW + c0);
    const float4 a2 = *reinterpret_cast<const float4*>(s_h + (yi.z - ys) * kPostTW + c0);
    const float4 a3 = *reinterpret_cast<const float4*>(s_h + (yi.w - ys) * kPostTW + c0);
    float v[4];
    v[0] = dot4(a0.x, a1.x, a2.x, a3.x, w);
    v[1] = dot4(a0.y, a1.y, a2.y, a3.y, w);
    v[2] = dot4(a0.z, a1.z, a2.z, a3.z, w);
    v[3] = dot4(a0.w, a1.w, a2.w, a3.w, w);
    const long long o = ((long long)b * p.H + oy) * p.W + ox0 + c0;
    const bool full = quad_in && ((o & 3) == 0);
    if (p.out_map) {
      if (full) {
        __stcs(reinterpret_cast<float4*>(p.out_map + o), make_float4(v[0], v[1], v[2], v[3]));
      } else {
        for (int e = 0; e < 4 && c0 + e < ncols; ++e) p.out_map[o + e] = v[e];
      }
    }
    if (p.out_vis) {
      if (full) {
        *reinterpret_cast<uint32_t*>(p.out_vis + o) =
            vis_byte(v[0]) | (vis_byte(v[1]) << 8) | (vis_byte(v[2]) << 16) | (vis_byte(v[3]) << 24);
      } else {
        for (int e = 0; e < 4 && c0 + e < ncols; ++e) p.out_vis[o + e] = static_cast<uint8_t>(vis_byte(v[e]));
      }
    }
    if (p.out_metric) {
      if (full) {
        uint2 m;
        m.x = pack_bf16(v[0] * p.metric_scale, v[1] * p.metric_scale);
        m.y = pack_bf16(v[2] * p.metric_scale, v[3] * p.metric_scale);
        __stcs(reinterpret_cast<uint2*>(p.out_metric + o), m);
      } else {
        for (int e = 0; e < 4 && c0 + e < ncols; ++e) p.out_metric[o + e] = __float2bfloat16(v[e] * p.metric_scale);
      }
    }
  }
}

// out_samples[b, s] = interpolated value at output pixel (row, col) = sample_rc[b, s, :]
__global__ void sample_kernel(const PostParams p, int B, const int32_t* sample_rc, int n_samples, float* out_samples) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * n_samples) return;
  const int b = i / n_samples;
  int row = sample_rc[i * 2], col = sample_rc[i * 2 + 1];
  row = min(max(row, 0), p.H - 1);
  col = min(max(col, 0), p.W - 1);
  const int4 xi = __ldg(reinterpret_cast<const int4*>(p.xidx) + col);
  const float4 wx = __ldg(reinterpret_cast<const float4*>(p.xw) + col);
  const int4 yi = __ldg(reinterpret_cast<const int4*>(p.yidx) + row);
  const float4 wy = __ldg(reinterpret_cast<const float4*>(p.yw) + row);
  const long long fbase = (long long)b * p.h * p.w;
  const int ys[4] = {yi.x, yi.y, yi.z, yi.w};
  float hrow[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const long long rb = fbase + (long long)ys[k] * p.w;
    hrow[k] = dot4(load_pred(p.src, p.in_dtype, rb + xi.x), load_pred(p.src, p.in_dtype, rb + xi.y),
                   load_pred(p.src, p.in_dtype, rb + xi.z), load_pred(p.src, p.in_dtype, rb + xi.w), wx);
  }
  out_samples[i] = dot4(hrow[0], hrow[1], hrow[2], hrow[3], wy);
}

}  // namespace deb

extern "C" int de_postprocess(const void* pred, int in_dtype, int B, int h, int w, int H, int W, float* out_map,
                              void* out_metric, float metric_scale, uint8_t* out_vis, const int32_t* sample_rc,
                              int n_samples, float* out_samples, void* stream) {
  using namespace deb;
  if (!pred || B <= 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0) return fail(DE_ERR_INVALID, "de_postprocess: bad args");
  if (in_dtype != DE_DTYPE_F32 && in_dtype != DE_DTYPE_BF16) return fail(DE_ERR_INVALID, "de_postprocess: in_dtype");
  const AxisTaps* tx = get_axis_taps(w, W, 1);
  const AxisTaps* ty = get_axis_taps(h, H, 1);
  if (!tx || !ty) return DE_ERR_CUDA;
  PostParams p{};
  p.src = pred; p.in_dtype = in_dtype;
  p.h = h; p.w = w; p.H = H; p.W = W;
  p.xidx = tx->d_idx; p.xw = tx->d_w; p.yidx = ty->d_idx; p.yw = ty->d_w;
  p.out_map = out_map;
  p.out_metric = reinterpret_cast<__nv_bfloat16*>(out_metric);
  p.metric_scale = metric_scale;
  p.out_vis = out_vis;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (out_map || out_metric || out_vis) {
    static const int cand[] = {32, 16, 8, 4, 2, 1};           // output rows per tile; 128 columns always
    size_t smem = 0;
    bool ok = false;
    for (int th : cand) {
      p.TW = kPostTW;
      p.TH = th;
      p.max_cols = 0;
      p.max_rows = ty->max_extent(p.TH);
      smem = sizeof(float) * (size_t)p.max_rows * kPostTW;
      if (smem <= 40 * 1024) {                                  // several CTAs per SM: the kernel lives on occupancy
        ok = true;
        break;
      }
    }
    if (!ok) return fail(DE_ERR_UNSUPPORTED, "de_postprocess: resampling window too large");
    dim3 grid((W + p.TW - 1) / p.TW, (H + p.TH - 1) / p.TH, B);
    postprocess_kernel<<<grid, kPostThreads, smem, st>>>(p);
    count_launch();
    int rc = cuda_check_launch("postprocess_kernel");
    if (rc != DE_OK) return rc;
  }
  if (sample_rc && n_samples > 0 && out_samples) {
    const int n = B * n_samples;
    sample_kernel<<<(n + 127) / 128, 128, 0, st>>>(p, B, sample_rc, n_samples, out_samples);
    count_launch();
    return cuda_check_launch("sample_kernel");
  }
  return DE_OK;
}
