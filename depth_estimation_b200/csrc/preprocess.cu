// K1 - fused MiDaS input transform (replaces /root/reference/depth_midas.py:22-23):
//   uint8 BGR HWC frame -> (BGR->RGB) -> /255 -> cv2.INTER_CUBIC resize to (net_h, net_w) ->
//   (x-mean)/std -> bf16 NCHW (or the ViT patch matrix, or fp32 NCHW for parity checks).
// The frame is read once and the tensor written once.  Two kernels:
//   * preprocess_copy_kernel - source size == net size (BASELINE configs #2/#3: 384x384 frames).  cv2.resize
//     degenerates to a copy there (and the cubic taps to [0,1,0,0]), so this is a pure streaming kernel:
//     12 input bytes -> 4 pixels x 3 channels per thread, 8-byte stores.
//   * preprocess_kernel - general resample.  A CTA owns 64 output columns x TH output rows; each of its 8 warps
//     takes whole output rows: (V) the 4 vertical taps are applied to the interleaved source BYTES of the row
//     strip, one aligned 32-bit word (4 bytes) per lane and iteration -> fp32 row in the warp's private smem
//     line; (H) the lane's two output columns (tap tables in registers) gather 4 taps x 3 channels from that
//     line, normalise and store.  Rows never leave their warp, so there is no block-wide barrier, and no integer
//     division anywhere.  The vertical pass does 16 byte->float conversions + 16 FMAs per input word; at 2.8x
//     down-scaling (1080p -> 384x672) that FMA-pipe work, not HBM, bounds the kernel.
#include <algorithm>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"
#include "resample_plan.h"

namespace deb {

struct PreParams {
  const uint8_t* src;
  long long row_stride, frame_stride, total_bytes;
  int B, H, W, net_h, net_w;
  const int32_t* xidx;
  const float* xw;
  const int32_t* yidx;
  const float* yw;
  float scale[3], shift[3];  // RGB order: out = v*scale - shift
  int layout;
  void* out;
  int TH, max_words, aligned;
  // format 1: NV12 source - src is the Y plane (row_stride / frame_stride describe it), uv the
  // interleaved half-resolution U,V plane; the BGR bytes the kernels consume are produced on the fly
  int format;
  const uint8_t* uv;
  long long uv_row_stride, uv_frame_stride;
};

constexpr int kPreThreads = 256;
#ifndef PRE_MINB
#define PRE_MINB 5          // 48 registers -> 5 CTAs/SM; measured best of {4,5,6} x unroll {1,2} (+7-10 % over 4)
#endif
#ifndef PRE_UNROLL
#define PRE_UNROLL 1
#endif
constexpr int kPreUnroll = PRE_UNROLL;
constexpr int kPreTW = 64;                                  // output columns per CTA (2 per lane)

// byte k of a little-endian word -> float, exactly: 0x4B0000bb is 2^23 + b
DEV float byte_to_float(uint32_t word, int k) {
  return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7540u | k)) - 8388608.0f;
}

DEV void store_px2(const PreParams& p, int b, int ch, int oy, int ox, float v0, float v1) {
  if (p.layout == 0) {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (((long long)b * 3 + ch) * p.net_h + oy) * p.net_w + ox;
    *reinterpret_cast<uint32_t*>(o) = pack_bf16(v0, v1);
  } else if (p.layout == 1) {
    const int gw = p.net_w >> 4, gh = p.net_h >> 4;
    const long long prow = ((long long)b * gh + (oy >> 4)) * gw + (ox >> 4);
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + prow * 768 + ch * 256 + (oy & 15) * 16 + (ox & 15);
    *reinterpret_cast<uint32_t*>(o) = pack_bf16(v0, v1);
  } else {
    float* o = reinterpret_cast<float*>(p.out) + (((long long)b * 3 + ch) * p.net_h + oy) * p.net_w + ox;
    *reinterpret_cast<float2*>(o) = make_float2(v0, v1);
  }
}

// aligned 32-bit load of source bytes [off, off+4) of the frame buffer; falls back to byte loads when the word is
// not 4-byte aligned or would cross the end of the buffer (bytes past the end read as 0 and are never used)
DEV uint32_t load_word(const PreParams& p, const uint8_t* ptr) {
  if (p.aligned && ptr + 4 <= p.src + p.total_bytes) return __ldg(reinterpret_cast<const uint32_t*>(ptr));
  uint32_t w = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (ptr + k < p.src + p.total_bytes) w |= static_cast<uint32_t>(__ldg(ptr + k)) << (8 * k);
  return w;
}

// NV12 -> the 32-bit word of interleaved BGR bytes [byte_off, byte_off+4) of frame row `row`, bit-exact with
// cv2.cvtColor(COLOR_YUV2BGR_NV12) (OpenCV color_yuv: ITU-R BT.601 limited range, 20-bit fixed point, saturating).
DEV uint32_t nv12_pixel_bgr(const PreParams& p, int b, int row, int px) {
  if (px >= p.W) return 0u;
  const int yv = __ldg(p.src + (long long)b * p.frame_stride + (long long)row * p.row_stride + px);
  const uint8_t* uvp = p.uv + (long long)b * p.uv_frame_stride + (long long)(row >> 1) * p.uv_row_stride + (px & ~1);
  const int u = static_cast<int>(__ldg(uvp)) - 128, v = static_cast<int>(__ldg(uvp + 1)) - 128;
  const int y = max(0, yv - 16) * 1220542 + (1 << 19);
  const int r = min(max((y + 1673527 * v) >> 20, 0), 255);
  const int g = min(max((y - 852492 * v - 409993 * u) >> 20, 0), 255);
  const int bl = min(max((y + 2116026 * u) >> 20, 0), 255);
  return static_cast<uint32_t>(bl) | (static_cast<uint32_t>(g) << 8) | (static_cast<uint32_t>(r) << 16);
}
DEV uint32_t nv12_word(const PreParams& p, int b, int row, int byte_off) {
  const int px = byte_off / 3, c0 = byte_off - px * 3;
  // 4 bytes starting at channel c0 of pixel px never reach past pixel px + 1
  const unsigned long long two = static_cast<unsigned long long>(nv12_pixel_bgr(p, b, row, px)) |
                                 (static_cast<unsigned long long>(nv12_pixel_bgr(p, b, row, px + 1)) << 24);
  return static_cast<uint32_t>(two >> (8 * c0));
}

DEV float vdot(float a, float b, float c, float d, float4 w) {
  float acc = a * w.x;
  acc = fmaf(b, w.y, acc);
  acc = fmaf(c, w.z, acc);
  acc = fmaf(d, w.w, acc);
  return acc;
}

__global__ void __launch_bounds__(kPreThreads, PRE_MINB) preprocess_kernel(const PreParams p) {
  extern __shared__ __align__(16) float s_line[];             // [8 warps][max_words * 4]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z;
  const int ox0 = blockIdx.x * kPreTW, oy0 = blockIdx.y * p.TH;
  const int ncols = min(kPreTW, p.net_w - ox0), nrows = min(p.TH, p.net_h - oy0);
  const int xs = __ldg(p.xidx + ox0 * 4), xe = __ldg(p.xidx + (ox0 + ncols - 1) * 4 + 3);
  const int sb0 = (xs * 3) & ~3;                               // first staged byte of the row strip (word aligned)
  const int nwords = (xe * 3 + 3 - sb0 + 3) >> 2;
  float* line = s_line + warp * (p.max_words * 4);
  const uint8_t* frame = p.src + (long long)b * p.frame_stride + sb0;
  const int y_last = __ldg(p.yidx + (oy0 + nrows - 1) * 4 + 3);
  const bool fast = p.format == 0 && p.aligned && (long long)b * p.frame_stride + sb0 + (long long)y_last * p.row_stride + 4LL * nwords <= p.total_bytes;

  // this lane's two output columns: tap offsets into the fp32 line (R channel = byte 2 of the BGR pixel)
  const bool active = 2 * lane < ncols;
  const int c0 = ox0 + 2 * lane;
  int4 o0 = make_int4(0, 0, 0, 0), o1 = o0;
  float4 w0 = make_float4(0.f, 0.f, 0.f, 0.f), w1 = w0;
  if (active) {
    o0 = __ldg(reinterpret_cast<const int4*>(p.xidx) + c0);
    o1 = __ldg(reinterpret_cast<const int4*>(p.xidx) + c0 + 1);
    w0 = __ldg(reinterpret_cast<const float4*>(p.xw) + c0);
    w1 = __ldg(reinterpret_cast<const float4*>(p.xw) + c0 + 1);
    o0.x = o0.x * 3 - sb0; o0.y = o0.y * 3 - sb0; o0.z = o0.z * 3 - sb0; o0.w = o0.w * 3 - sb0;
    o1.x = o1.x * 3 - sb0; o1.y = o1.y * 3 - sb0; o1.z = o1.z * 3 - sb0; o1.w = o1.w * 3 - sb0;
  }

  for (int ry = warp; ry < nrows; ry += kPreThreads / 32) {
    const int oy = oy0 + ry;
    const int4 yi = __ldg(reinterpret_cast<const int4*>(p.yidx) + oy);
    const float4 wy = __ldg(reinterpret_cast<const float4*>(p.yw) + oy);
    const uint8_t* r0 = frame + (long long)yi.x * p.row_stride;
    const uint8_t* r1 = frame + (long long)yi.y * p.row_stride;
    const uint8_t* r2 = frame + (long long)yi.z * p.row_stride;
    const uint8_t* r3 = frame + (long long)yi.w * p.row_stride;
    // ---- V: four vertical taps on the raw bytes, one 32-bit word per lane and iteration
#pragma unroll kPreUnroll
    for (int wi = lane; wi < nwords; wi += 32) {
      uint32_t u0, u1, u2, u3;
      if (fast) {                                               // whole tile aligned and inside the buffer (CTA-uniform)
        u0 = __ldg(reinterpret_cast<const uint32_t*>(r0) + wi);
        u1 = __ldg(reinterpret_cast<const uint32_t*>(r1) + wi);
        u2 = __ldg(reinterpret_cast<const uint32_t*>(r2) + wi);
        u3 = __ldg(reinterpret_cast<const uint32_t*>(r3) + wi);
      } else if (p.format == 0) {
        u0 = load_word(p, r0 + 4 * wi); u1 = load_word(p, r1 + 4 * wi);
        u2 = load_word(p, r2 + 4 * wi); u3 = load_word(p, r3 + 4 * wi);
      } else {
        u0 = nv12_word(p, b, yi.x, sb0 + 4 * wi); u1 = nv12_word(p, b, yi.y, sb0 + 4 * wi);
        u2 = nv12_word(p, b, yi.z, sb0 + 4 * wi); u3 = nv12_word(p, b, yi.w, sb0 + 4 * wi);
      }
      float4 v;
      v.x = vdot(byte_to_float(u0, 0), byte_to_float(u1, 0), byte_to_float(u2, 0), byte_to_float(u3, 0), wy);
      v.y = vdot(byte_to_float(u0, 1), byte_to_float(u1, 1), byte_to_float(u2, 1), byte_to_float(u3, 1), wy);
      v.z = vdot(byte_to_float(u0, 2), byte_to_float(u1, 2), byte_to_float(u2, 2), byte_to_float(u3, 2), wy);
      v.w = vdot(byte_to_float(u0, 3), byte_to_float(u1, 3), byte_to_float(u2, 3), byte_to_float(u3, 3), wy);
      reinterpret_cast<float4*>(line)[wi] = v;
    }
    __syncwarp();
    // ---- H: four horizontal taps per channel for the lane's two columns, normalise, store (BGR -> RGB)
    if (active) {
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        const float* l = line + (2 - ch);
        const float a = vdot(l[o0.x], l[o0.y], l[o0.z], l[o0.w], w0);
        const float c = vdot(l[o1.x], l[o1.y], l[o1.z], l[o1.w], w1);
        store_px2(p, b, ch, oy, c0, fmaf(a, p.scale[ch], -p.shift[ch]), fmaf(c, p.scale[ch], -p.shift[ch]));
      }
    }
    __syncwarp();
  }
}

// source size == net size, generic alignment: 4 pixels (12 bytes, 3 words) per thread
__global__ void __launch_bounds__(128) preprocess_copy_kernel(const PreParams p) {
  const int x4 = blockIdx.x * 128 + threadIdx.x;
  if (x4 * 4 >= p.W) return;
  const int y = blockIdx.y, b = blockIdx.z;
  const uint8_t* ptr = p.src + (long long)b * p.frame_stride + (long long)y * p.row_stride + x4 * 12;
  uint32_t u0, u1, u2;
  if (p.format == 0) {
    u0 = load_word(p, ptr); u1 = load_word(p, ptr + 4); u2 = load_word(p, ptr + 8);
  } else {
    u0 = nv12_word(p, b, y, x4 * 12); u1 = nv12_word(p, b, y, x4 * 12 + 4); u2 = nv12_word(p, b, y, x4 * 12 + 8);
  }
  // bytes: u0 = B0 G0 R0 B1 | u1 = G1 R1 B2 G2 | u2 = R2 B3 G3 R3
  const float r[4] = {byte_to_float(u0, 2), byte_to_float(u1, 1), byte_to_float(u2, 0), byte_to_float(u2, 3)};
  const float g[4] = {byte_to_float(u0, 1), byte_to_float(u1, 0), byte_to_float(u1, 3), byte_to_float(u2, 2)};
  const float bl[4] = {byte_to_float(u0, 0), byte_to_float(u0, 3), byte_to_float(u1, 2), byte_to_float(u2, 1)};
  const int ox = x4 * 4;
  store_px2(p, b, 0, y, ox, fmaf(r[0], p.scale[0], -p.shift[0]), fmaf(r[1], p.scale[0], -p.shift[0]));
  store_px2(p, b, 0, y, ox + 2, fmaf(r[2], p.scale[0], -p.shift[0]), fmaf(r[3], p.scale[0], -p.shift[0]));
  store_px2(p, b, 1, y, ox, fmaf(g[0], p.scale[1], -p.shift[1]), fmaf(g[1], p.scale[1], -p.shift[1]));
  store_px2(p, b, 1, y, ox + 2, fmaf(g[2], p.scale[1], -p.shift[1]), fmaf(g[3], p.scale[1], -p.shift[1]));
  store_px2(p, b, 2, y, ox, fmaf(bl[0], p.scale[2], -p.shift[2]), fmaf(bl[1], p.scale[2], -p.shift[2]));
  store_px2(p, b, 2, y, ox + 2, fmaf(bl[2], p.scale[2], -p.shift[2]), fmaf(bl[3], p.scale[2], -p.shift[2]));
}

// source size == net size, 16-byte aligned rows, bf16 output: 16 pixels per thread - three 16-byte loads, and per
// channel one 32-byte run (a full patch row in the patch-matrix layout) written with two 16-byte stores.
// pixel i of the 48 bytes: B at 3i, G at 3i+1, R at 3i+2.
template <int CH>
DEV void copy16_channel(const uint32_t (&u)[12], float scale, float shift, uint4& lo, uint4& hi) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int byte = 3 * i + (2 - CH);                          // RGB channel CH lives at BGR byte 2-CH
    v[i] = fmaf(byte_to_float(u[byte >> 2], byte & 3), scale, -shift);
  }
  lo = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
  hi = make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15]));
}

__global__ void __launch_bounds__(256) preprocess_copy16_kernel(const PreParams p) {
  const int nx = p.W >> 4;
  const long long item = (long long)blockIdx.x * 256 + threadIdx.x;     // (frame row, 16-pixel group), flattened
  const long long rowi = item / nx;
  if (rowi >= (long long)p.B * p.H) return;
  const int x16 = static_cast<int>(item - rowi * nx);
  const int b = static_cast<int>(rowi / p.H), y = static_cast<int>(rowi - (long long)b * p.H);
  const uint4* ptr = reinterpret_cast<const uint4*>(p.src + (long long)b * p.frame_stride + (long long)y * p.row_stride) + x16 * 3;
  const uint4 q0 = __ldg(ptr), q1 = __ldg(ptr + 1), q2 = __ldg(ptr + 2);
  const uint32_t u[12] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w};
  __nv_bfloat16* base = reinterpret_cast<__nv_bfloat16*>(p.out);
  long long o, cstride;                                         // element offset of (b, ch 0, y, x16*16); channel stride
  if (p.layout == 0) {
    cstride = (long long)p.net_h * p.net_w;
    o = ((long long)b * 3 * p.net_h + y) * p.net_w + x16 * 16;
  } else {
    cstride = 256;
    o = (((long long)b * (p.net_h >> 4) + (y >> 4)) * nx + x16) * 768 + (y & 15) * 16;
  }
  uint4 lo, hi;
  copy16_channel<0>(u, p.scale[0], p.shift[0], lo, hi);
  __stcs(reinterpret_cast<uint4*>(base + o), lo);
  __stcs(reinterpret_cast<uint4*>(base + o) + 1, hi);
  copy16_channel<1>(u, p.scale[1], p.shift[1], lo, hi);
  __stcs(reinterpret_cast<uint4*>(base + o + cstride), lo);
  __stcs(reinterpret_cast<uint4*>(base + o + cstride) + 1, hi);
  copy16_channel<2>(u, p.scale[2], p.shift[2], lo, hi);
  __stcs(reinterpret_cast<uint4*>(base + o + 2 * cstride), lo);
  __stcs(reinterpret_cast<uint4*>(base + o + 2 * cstride) + 1, hi);
}

}  // namespace deb

static int run_preprocess(const uint8_t* bgr, const uint8_t* uv, int64_t uv_row_stride, int64_t uv_frame_stride, int B,
                          int H, int W, int64_t row_stride, int64_t frame_stride, int net_h, int net_w,
                          const float* mean_host, const float* std_host, int layout, void* out, void* stream) {
  using namespace deb;
  const int format = uv ? 1 : 0;
  if (!bgr || !out || B <= 0 || H <= 0 || W <= 0) return fail(DE_ERR_INVALID, "de_preprocess: bad args");
  if (net_h % 32 != 0 || net_w % 32 != 0 || net_h <= 0 || net_w <= 0)
    return fail(DE_ERR_INVALID, "de_preprocess: net size %dx%d is not a positive multiple of 32", net_h, net_w);
  if (layout < 0 || layout > 2) return fail(DE_ERR_INVALID, "de_preprocess: layout %d", layout);
  if (format == 0 && (row_stride < (int64_t)W * 3 || frame_stride < row_stride * H)) return fail(DE_ERR_INVALID, "de_preprocess: strides");
  if (format == 1 && (row_stride < W || frame_stride < row_stride * H || (W & 1) || (H & 1) || uv_row_stride < W ||
                      uv_frame_stride < uv_row_stride * (H / 2)))
    return fail(DE_ERR_INVALID, "de_preprocess_nv12: strides / odd frame size");
  const AxisTaps* tx = get_axis_taps(W, net_w, 0);
  const AxisTaps* ty = get_axis_taps(H, net_h, 0);
  if (!tx || !ty) return DE_ERR_CUDA;
  PreParams p{};
  p.src = bgr;
  p.row_stride = row_stride;
  p.frame_stride = frame_stride;
  p.total_bytes = (long long)frame_stride * (B - 1) + (long long)row_stride * (H - 1) + (long long)W * 3;
  p.B = B; p.H = H; p.W = W; p.net_h = net_h; p.net_w = net_w;
  p.xidx = tx->d_idx; p.xw = tx->d_w; p.yidx = ty->d_idx; p.yw = ty->d_w;
  for (int c = 0; c < 3; ++c) {
    p.scale[c] = (float)(1.0 / (255.0 * (double)std_host[c]));
    p.shift[c] = (float)((double)mean_host[c] / (double)std_host[c]);
  }
  p.layout = layout;
  p.out = out;
  p.format = format;
  p.uv = uv;
  p.uv_row_stride = uv_row_stride;
  p.uv_frame_stride = uv_frame_stride;
  p.aligned = ((reinterpret_cast<uintptr_t>(bgr) | (uintptr_t)row_stride | (uintptr_t)frame_stride) & 3) == 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (H == net_h && W == net_w) {                             // cv2.resize is a copy for equal sizes
    const bool wide = format == 0 && layout != 2 && W % 16 == 0 &&
                      ((reinterpret_cast<uintptr_t>(bgr) | (uintptr_t)row_stride | (uintptr_t)frame_stride |
                        reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    if (wide) {
      const long long items = (long long)B * H * (W / 16);
      preprocess_copy16_kernel<<<(unsigned)((items + 255) / 256), 256, 0, st>>>(p);
    } else {
      dim3 grid((W / 4 + 127) / 128, H, B);
      preprocess_copy_kernel<<<grid, 128, 0, st>>>(p);
    }
    count_launch();
    return cuda_check_launch("preprocess_copy_kernel");
  }
  p.TH = 32;
  p.max_words = (tx->max_extent(kPreTW) * 3 + 6) / 4 + 1;
  const size_t smem = (size_t)(kPreThreads / 32) * p.max_words * 16;
  if (smem > 200 * 1024) return fail(DE_ERR_UNSUPPORTED, "de_preprocess: downscale factor too large for the staged row strip");
  if (smem > 48 * 1024) {
    static bool configured = false;
    if (!configured) {
      cudaError_t e = cudaFuncSetAttribute(preprocess_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(preprocess)");
      configured = true;
    }
  }
  dim3 grid((net_w + kPreTW - 1) / kPreTW, (net_h + p.TH - 1) / p.TH, B);
  preprocess_kernel<<<grid, kPreThreads, smem, st>>>(p);
  count_launch();
  return cuda_check_launch("preprocess_kernel");
}

extern "C" int de_preprocess(const uint8_t* bgr, int B, int H, int W, int64_t row_stride, int64_t frame_stride,
                             int net_h, int net_w, const float* mean_host, const float* std_host, int layout,
                             void* out, void* stream) {
  return run_preprocess(bgr, nullptr, 0, 0, B, H, W, row_stride, frame_stride, net_h, net_w, mean_host, std_host, layout,
                        out, stream);
}

extern "C" int de_preprocess_nv12(const uint8_t* y_plane, const uint8_t* uv_plane, int B, int H, int W, int64_t y_row_stride,
                                  int64_t y_frame_stride, int64_t uv_row_stride, int64_t uv_frame_stride, int net_h,
                                  int net_w, const float* mean_host, const float* std_host, int layout, void* out,
                                  void* stream) {
  if (!uv_plane) return deb::fail(DE_ERR_INVALID, "de_preprocess_nv12: null UV plane");
  return run_preprocess(y_plane, uv_plane, uv_row_stride, uv_frame_stride, B, H, W, y_row_stride, y_frame_stride, net_h,
                        net_w, mean_host, std_host, layout, out, stream);
}
