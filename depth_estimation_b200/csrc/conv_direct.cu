// Direct (CUDA-core) convolution kernels for the shapes that are not dense contractions:
//   * the 3-channel stems (EfficientNet-Lite3 3x3 s2, ResNetV2 7x7 s2) - K = 27 / 147, reads the NCHW
//     bf16 tensor K1 writes, emits NHWC bf16;
//   * (the depthwise convolutions of the MBConv blocks live in depthwise.cu);
//   * MaxPool 3x3 s2 TF-same (ResNetV2 stem).
// Restates [UPSTREAM] geffnet Conv2dSame / timm StdConv2dSame+MaxPool2dSame padding rule:
//   pad_total = max((ceil(i/s)-1)*s + k - i, 0), pad_before = pad_total / 2 (extra pixel bottom/right).
#include <stdlib.h>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

__host__ __device__ inline int same_pad_before(int in, int k, int s) {
  const int out = (in + s - 1) / s;
  int total = (out - 1) * s + k - in;
  if (total < 0) total = 0;
  return total / 2;
}

// ---------------------------------------------------------------- stem: NCHW bf16 (3 ch) -> NHWC bf16
constexpr int kStemCo = 16;  // output channels per thread

// Two adjacent output pixels x 16 output channels per thread: every weight float4 read from shared memory feeds
// four packed FFMA2 (two channels each, both pixels) instead of four scalar FMAs of one pixel - half the shared-memory
// reads and half the FMA-pipe instructions per MAC of the one-pixel version.
// Accumulation order per output element is unchanged (bias, then taps in (ky, kx, c) order).
__global__ void __launch_bounds__(128) stem_conv_kernel(const void* __restrict__ x, int x_f32,
                                                        const float* __restrict__ w,  // [Cout][k*k*3] (kh,kw,c)
                                                        const float* __restrict__ bias, void* __restrict__ out, int out_f32,
                                                        int B, int H, int W, int Cout, int k, int stride, int Ho, int Wo,
                                                        int pad_y, int pad_x, int act) {
  extern __shared__ float s_w[];  // [k*k*3][Cout] transposed for broadcast float4 reads
  const int taps = k * k * 3;
  // coalesced global reads (adjacent lanes -> adjacent taps of one filter), transposed on the way into smem.  Reading
  // with adjacent lanes on adjacent FILTERS (588-byte stride: 32 lines per warp load, 73 dependent rounds) made this
  // prologue ~20 us per CTA - most of the kernel's time at batch 1.
  for (int i = threadIdx.x; i < taps * Cout; i += blockDim.x) {
    const int co = i / taps, t = i - co * taps;
    s_w[t * Cout + co] = __ldg(w + i);
  }
  __syncthreads();
  const int cgs = Cout / kStemCo;
  const int Wo2 = (Wo + 1) >> 1;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)B * Ho * Wo2 * cgs) return;
  // pixel-major thread order (adjacent threads -> adjacent pixel pairs) keeps the input reads coalesced
  const int ox0 = static_cast<int>(idx % Wo2) * 2;
  long long t = idx / Wo2;
  const int oy = static_cast<int>(t % Ho);
  t /= Ho;
  const int cg = static_cast<int>(t % cgs);
  const int b = static_cast<int>(t / cgs);
  const bool two = ox0 + 1 < Wo;
  float2 acc[2][kStemCo / 2];
#pragma unroll
  for (int j = 0; j < kStemCo / 2; ++j) {
    const float2 bv = bias ? make_float2(bias[cg * kStemCo + 2 * j], bias[cg * kStemCo + 2 * j + 1]) : make_float2(0.f, 0.f);
    acc[0][j] = bv;
    acc[1][j] = bv;
  }
  const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x) + (long long)b * 3 * H * W;
  const float* xf = reinterpret_cast<const float*>(x) + (long long)b * 3 * H * W;
  for (int ky = 0; ky < k; ++ky) {
    const int iy = oy * stride + ky - pad_y;
    if (iy < 0 || iy >= H) continue;
    for (int kx = 0; kx < k; ++kx) {
      const int ix0 = ox0 * stride + kx - pad_x, ix1 = ix0 + stride;
      const bool in0 = ix0 >= 0 && ix0 < W, in1 = two && ix1 >= 0 && ix1 < W;
      if (!in0 && !in1) continue;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const long long e = ((long long)c * H + iy) * W;
        float v0 = 0.f, v1 = 0.f;
        if (in0) v0 = x_f32 ? __ldg(xf + e + ix0) : __bfloat162float(xb[e + ix0]);
        if (in1) v1 = x_f32 ? __ldg(xf + e + ix1) : __bfloat162float(xb[e + ix1]);
        const float4* wp = reinterpret_cast<const float4*>(s_w + ((ky * k + kx) * 3 + c) * Cout + cg * kStemCo);
        const float2 p0 = make_float2(v0, v0), p1 = make_float2(v1, v1);
        // a pixel whose tap falls into the padding must skip the tap (x + 0*w would still be exact, but keeps -0 / NaN
        // semantics identical to the one-pixel kernel only if skipped)
#pragma unroll
        for (int j4 = 0; j4 < kStemCo / 4; ++j4) {
          const float4 ww = wp[j4];
          if (in0) {
            acc[0][2 * j4] = ffma2(p0, make_float2(ww.x, ww.y), acc[0][2 * j4]);
            acc[0][2 * j4 + 1] = ffma2(p0, make_float2(ww.z, ww.w), acc[0][2 * j4 + 1]);
          }
          if (in1) {
            acc[1][2 * j4] = ffma2(p1, make_float2(ww.x, ww.y), acc[1][2 * j4]);
            acc[1][2 * j4 + 1] = ffma2(p1, make_float2(ww.z, ww.w), acc[1][2 * j4 + 1]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int px = 0; px < 2; ++px) {
    if (px == 1 && !two) break;
    float r[kStemCo];
#pragma unroll
    for (int j = 0; j < kStemCo / 2; ++j) {
      float a = acc[px][j].x, c = acc[px][j].y;
      if (act == DE_ACT_RELU6) { a = fminf(fmaxf(a, 0.f), 6.f); c = fminf(fmaxf(c, 0.f), 6.f); }
      else if (act == DE_ACT_RELU) { a = fmaxf(a, 0.f); c = fmaxf(c, 0.f); }
      r[2 * j] = a;
      r[2 * j + 1] = c;
    }
    const long long e = (((long long)b * Ho + oy) * Wo + ox0 + px) * Cout + cg * kStemCo;
    if (out_f32) {
      float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + e);
#pragma unroll
      for (int j4 = 0; j4 < kStemCo / 4; ++j4) op[j4] = make_float4(r[j4 * 4], r[j4 * 4 + 1], r[j4 * 4 + 2], r[j4 * 4 + 3]);
    } else {
      uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(out) + e);
      op[0] = make_uint4(pack_bf16(r[0], r[1]), pack_bf16(r[2], r[3]), pack_bf16(r[4], r[5]), pack_bf16(r[6], r[7]));
      op[1] = make_uint4(pack_bf16(r[8], r[9]), pack_bf16(r[10], r[11]), pack_bf16(r[12], r[13]), pack_bf16(r[14], r[15]));
    }
  }
}

// EfficientNet-Lite3 stem (3x3, stride 2, 3 -> 32 channels, bf16 in / out) staged through shared memory: a CTA owns an
// 8 x 32-pixel output tile; its 17 x 65 x 3 input window is read once with coalesced loads (zero outside the image =
// TF-same padding) and kept as fp32, the 27 x 32 filter bank sits next to it.  A thread computes two adjacent pixels x
// all 32 channels: per (channel, ky) it reads its five window columns with one LDS.128 + one LDS.32, every weight
// float4 (a warp-wide broadcast) feeds four FFMA2.  The kernel above issues one dependent 2-byte global load per
// (tap, channel, pixel) behind a bounds branch - 34 us per 32 frames of 128x256 (an 8th of the HBM rate) - this one has
// no global load inside the tap loops.  Same accumulation order per output (bias, then (ky, kx, c)).
constexpr int kStTH = 8, kStTW = 32;                      // output tile
constexpr int kStWH = 2 * (kStTH - 1) + 3, kStWW = 2 * (kStTW - 1) + 3;   // 17 x 65 window
constexpr int kStWP = 68;                                 // padded window row (16-byte aligned rows)
__global__ void __launch_bounds__(128) stem3x3s2_tiled_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                                                              const float* __restrict__ bias, __nv_bfloat16* __restrict__ out,
                                                              int B, int H, int W, int Ho, int Wo, int pad_y, int pad_x, int act) {
  __shared__ __align__(16) float s_in[3 * kStWH * kStWP];
  __shared__ __align__(16) float s_wt[27 * 32];             // [tap = (ky*3 + kx)*3 + c][32]
  __shared__ float s_bias[32];
  const int b = blockIdx.z, oy0 = blockIdx.y * kStTH, ox0 = blockIdx.x * kStTW;
  const int iy0 = oy0 * 2 - pad_y, ix0 = ox0 * 2 - pad_x;
  for (int i = threadIdx.x; i < 27 * 32; i += 128) {
    const int co = i / 27, t = i - co * 27;                 // coalesced read of [32][27], transposed into smem
    s_wt[t * 32 + co] = __ldg(w + i);
  }
  if (threadIdx.x < 32) s_bias[threadIdx.x] = bias ? __ldg(bias + threadIdx.x) : 0.f;
  const __nv_bfloat16* xb = x + (long long)b * 3 * H * W;
  for (int i = threadIdx.x; i < 3 * kStWH * kStWP; i += 128) {
    const int col = i % kStWP, r = i / kStWP;               // r = c * 17 + row
    const int c = r / kStWH, row = r - c * kStWH;
    const int iy = iy0 + row, ix = ix0 + col;
    float v = 0.f;
    if (col < kStWW && iy >= 0 && iy < H && ix >= 0 && ix < W) v = __bfloat162float(xb[((long long)c * H + iy) * W + ix]);
    s_in[i] = v;
  }
  __syncthreads();
  const int px = threadIdx.x & 15, ty = threadIdx.x >> 4;  // pixel pair (columns 2px, 2px+1) of tile row ty
  float2 acc[2][16];
#pragma unroll
  for (int j = 0; j < 16; ++j) acc[0][j] = acc[1][j] = make_float2(s_bias[2 * j], s_bias[2 * j + 1]);
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    float in[3][5];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float* rp = s_in + (c * kStWH + 2 * ty + ky) * kStWP + 4 * px;
      const float4 v4 = *reinterpret_cast<const float4*>(rp);
      in[c][0] = v4.x; in[c][1] = v4.y; in[c][2] = v4.z; in[c][3] = v4.w; in[c][4] = rp[4];
    }
#pragma unroll
    for (int kx = 0; kx < 3; ++kx)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float4* wp = reinterpret_cast<const float4*>(s_wt + ((ky * 3 + kx) * 3 + c) * 32);
        const float2 p0 = make_float2(in[c][kx], in[c][kx]), p1 = make_float2(in[c][kx + 2], in[c][kx + 2]);
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 ww = wp[j4];
          acc[0][2 * j4] = ffma2(p0, make_float2(ww.x, ww.y), acc[0][2 * j4]);
          acc[0][2 * j4 + 1] = ffma2(p0, make_float2(ww.z, ww.w), acc[0][2 * j4 + 1]);
          acc[1][2 * j4] = ffma2(p1, make_float2(ww.x, ww.y), acc[1][2 * j4]);
          acc[1][2 * j4 + 1] = ffma2(p1, make_float2(ww.z, ww.w), acc[1][2 * j4 + 1]);
        }
      }
  }
  const int oy = oy0 + ty;
  if (oy >= Ho) return;
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int ox = ox0 + 2 * px + q;
    if (ox >= Wo) break;
    uint32_t o[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float a = acc[q][j].x, c = acc[q][j].y;
      if (act == DE_ACT_RELU6) { a = fminf(fmaxf(a, 0.f), 6.f); c = fminf(fmaxf(c, 0.f), 6.f); }
      else if (act == DE_ACT_RELU) { a = fmaxf(a, 0.f); c = fmaxf(c, 0.f); }
      o[j] = pack_bf16(a, c);
    }
    uint4* op = reinterpret_cast<uint4*>(out + (((long long)b * Ho + oy) * Wo + ox) * 32);
#pragma unroll
    for (int j = 0; j < 4; ++j) op[j] = make_uint4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
  }
}

// ---------------------------------------------------------------- stem as a GEMM operand: fp32 NCHW -> split3 im2col
// One warp per output pixel: row [hi(kp) | hi(kp) | lo(kp)] of bf16, column t = (ky*k + kx)*3 + c for t < k*k*3
// (the order of the stem filter bank), zero beyond; hi = bf16(v), lo = bf16(v - hi).  Against weights packed
// [Wh | Wl | Wh] the bf16 tensor-core GEMM then computes the stem at bf16x3 (~16 bit) precision - the same contract as
// the ResNetV2 stages - instead of 147 CUDA-core taps per output (DPT_Hybrid: 1.55 ms per 32-frame step, 94 us at
// batch 1 for the direct kernel).  Lanes own column PAIRS, so every store is a 4-byte word.
// A warp owns kStemRun consecutive output pixels; the (ky, kx, c) decode of a lane's columns is pixel-independent and
// done once (the per-pixel version spent its time in integer divisions: ~25 us per 384^2 frame).
constexpr int kStemRun = 16;
constexpr int kStemMaxJ = 3;                       // column pairs per lane: kp <= 192
__global__ void __launch_bounds__(256) stem_im2col_split3_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                                 int B, int H, int W, int k, int stride, int Ho, int Wo,
                                                                 int pad_y, int pad_x, int kp) {
  const long long npix = (long long)B * Ho * Wo;
  const long long pix0 = ((long long)blockIdx.x * 8 + (threadIdx.x >> 5)) * kStemRun;
  if (pix0 >= npix) return;
  const int lane = threadIdx.x & 31;
  const int taps = k * k * 3;
  int dy[kStemMaxJ][2], dx[kStemMaxJ][2], co[kStemMaxJ][2];
#pragma unroll
  for (int j = 0; j < kStemMaxJ; ++j)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int t = 2 * (lane + 32 * j) + e;
      const int tap = t / 3, c = t - tap * 3;
      const int ky = tap / k;
      dy[j][e] = t < taps ? ky : -100000;          // out-of-range row: the bounds test below yields 0
      dx[j][e] = tap - ky * k;
      co[j][e] = c * H * W;
    }
  const long long pix1 = pix0 + kStemRun < npix ? pix0 + kStemRun : npix;
  for (long long pix = pix0; pix < pix1; ++pix) {
    const int ox = static_cast<int>(pix % Wo);
    const long long r = pix / Wo;
    const int oy = static_cast<int>(r % Ho), b = static_cast<int>(r / Ho);
    const float* xb = x + (long long)b * 3 * H * W;
    uint32_t* row = reinterpret_cast<uint32_t*>(out + pix * 3 * kp);
    const int iy0 = oy * stride - pad_y, ix0 = ox * stride - pad_x;
#pragma unroll
    for (int j = 0; j < kStemMaxJ; ++j) {
      const int pj = lane + 32 * j;
      if (2 * pj < kp) {
        float v[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int iy = iy0 + dy[j][e], ix = ix0 + dx[j][e];
          v[e] = (iy >= 0 && iy < H && ix >= 0 && ix < W) ? __ldg(xb + co[j][e] + iy * W + ix) : 0.f;
        }
        const uint32_t hi = pack_bf16(v[0], v[1]);
        const uint32_t lo = pack_bf16(v[0] - bf16_lo(hi), v[1] - bf16_hi(hi));
        row[pj] = hi;
        row[(kp >> 1) + pj] = hi;
        row[kp + pj] = lo;
      }
    }
  }
}

// ---------------------------------------------------------------- maxpool 3x3 s2 TF-same, NHWC bf16
// in: bf16 or fp32 NHWC.  out_mode 0: bf16 NHWC; 1: the [hi | hi | lo] split3 operand layout (misc.cu store_split3).
__global__ void maxpool3x3s2_kernel(const void* __restrict__ x, int x_f32, __nv_bfloat16* __restrict__ out, int out_mode,
                                    int B, int H, int W, int C, int Ho, int Wo, int pad_y, int pad_x) {
  const int cgs = C >> 3;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)B * Ho * Wo * cgs) return;
  const int c = static_cast<int>(idx % cgs) * 8;
  long long t = idx / cgs;
  const int ox = static_cast<int>(t % Wo);
  t /= Wo;
  const int oy = static_cast<int>(t % Ho);
  const int b = static_cast<int>(t / Ho);
  float m[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) m[j] = -INFINITY;
  for (int ky = 0; ky < 3; ++ky) {
    const int iy = oy * 2 + ky - pad_y;
    if (iy < 0 || iy >= H) continue;
    for (int kx = 0; kx < 3; ++kx) {
      const int ix = ox * 2 + kx - pad_x;
      if (ix < 0 || ix >= W) continue;
      const long long e = (((long long)b * H + iy) * W + ix) * C + c;
      if (x_f32) {
        const float4 u0 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(x) + e);
        const float4 u1 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(x) + e + 4);
        m[0] = fmaxf(m[0], u0.x); m[1] = fmaxf(m[1], u0.y); m[2] = fmaxf(m[2], u0.z); m[3] = fmaxf(m[3], u0.w);
        m[4] = fmaxf(m[4], u1.x); m[5] = fmaxf(m[5], u1.y); m[6] = fmaxf(m[6], u1.z); m[7] = fmaxf(m[7], u1.w);
      } else {
        const uint4 v = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(x) + e);
        m[0] = fmaxf(m[0], bf16_lo(v.x)); m[1] = fmaxf(m[1], bf16_hi(v.x));
        m[2] = fmaxf(m[2], bf16_lo(v.y)); m[3] = fmaxf(m[3], bf16_hi(v.y));
        m[4] = fmaxf(m[4], bf16_lo(v.z)); m[5] = fmaxf(m[5], bf16_hi(v.z));
        m[6] = fmaxf(m[6], bf16_lo(v.w)); m[7] = fmaxf(m[7], bf16_hi(v.w));
      }
    }
  }
  const long long row = ((long long)b * Ho + oy) * Wo + ox;
  uint32_t hi[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) hi[k] = pack_bf16(m[2 * k], m[2 * k + 1]);
  if (out_mode == 0) {
    *reinterpret_cast<uint4*>(out + row * C + c) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
  } else {
    uint32_t lo[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) lo[k] = pack_bf16(m[2 * k] - bf16_lo(hi[k]), m[2 * k + 1] - bf16_hi(hi[k]));
    __nv_bfloat16* o = out + row * 3 * C + c;
    *reinterpret_cast<uint4*>(o) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4*>(o + C) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4*>(o + 2 * C) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
}

}  // namespace deb

using namespace deb;

extern "C" int de_stem_conv(const void* x_nchw, int x_dtype, const float* w, const float* bias, void* out_nhwc, int out_dtype,
                            int B, int H, int W, int Cout, int k, int stride, int act, void* stream) {
  deb::ProfScope prof_("stem_conv", 0.0, stream, B, H, W, Cout);
  if (!x_nchw || !w || !out_nhwc || Cout % kStemCo != 0 || k < 1 || k > 7)
    return fail(DE_ERR_INVALID, "de_stem_conv: bad args (Cout %% 16, k <= 7)");
  const int Ho = (H + stride - 1) / stride, Wo = (W + stride - 1) / stride;
  static const bool direct = getenv("DEB_STEM_UNTILED") && atoi(getenv("DEB_STEM_UNTILED")) != 0;   // A/B switch
  if (k == 3 && stride == 2 && Cout == 32 && x_dtype != DE_DTYPE_F32 && out_dtype != DE_DTYPE_F32 && B <= 65535 && !direct) {
    const dim3 grid((Wo + kStTW - 1) / kStTW, (Ho + kStTH - 1) / kStTH, B);
    stem3x3s2_tiled_kernel<<<grid, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(x_nchw), w, bias, reinterpret_cast<__nv_bfloat16*>(out_nhwc), B, H, W, Ho, Wo,
        same_pad_before(H, k, stride), same_pad_before(W, k, stride), act);
    count_launch();
    return cuda_check_launch("stem3x3s2_tiled_kernel");
  }
  const long long n = (long long)B * Ho * ((Wo + 1) / 2) * (Cout / kStemCo);
  const size_t smem = sizeof(float) * k * k * 3 * Cout;
  stem_conv_kernel<<<(unsigned)((n + 127) / 128), 128, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
      x_nchw, x_dtype == DE_DTYPE_F32, w, bias, out_nhwc, out_dtype == DE_DTYPE_F32, B, H, W, Cout, k, stride, Ho, Wo, same_pad_before(H, k, stride), same_pad_before(W, k, stride), act);
  count_launch();
  return cuda_check_launch("stem_conv_kernel");
}

extern "C" int de_stem_im2col_split3(const float* x_nchw, void* out_bf16, int B, int H, int W, int k, int stride, int kp,
                                     void* stream) {
  deb::ProfScope prof_("stem_im2col_split3", 0.0, stream, B, H, W, kp);
  if (!x_nchw || !out_bf16 || B <= 0 || k < 1 || stride < 1 || kp % 8 != 0 || kp < k * k * 3 || kp > 64 * kStemMaxJ)
    return fail(DE_ERR_INVALID, "de_stem_im2col_split3: bad args (kp %% 8 == 0, 3*k*k <= kp <= 192)");
  const int Ho = (H + stride - 1) / stride, Wo = (W + stride - 1) / stride;
  const long long n = (long long)B * Ho * Wo;
  const long long nwarps = (n + kStemRun - 1) / kStemRun;
  stem_im2col_split3_kernel<<<(unsigned)((nwarps + 7) / 8), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x_nchw, reinterpret_cast<__nv_bfloat16*>(out_bf16), B, H, W, k, stride, Ho, Wo, same_pad_before(H, k, stride),
      same_pad_before(W, k, stride), kp);
  count_launch();
  return cuda_check_launch("stem_im2col_split3_kernel");
}

extern "C" int de_maxpool3x3s2(const void* x, int x_dtype, void* out, int out_mode, int B, int H, int W, int C,
                               void* stream) {
  deb::ProfScope prof_("maxpool3x3s2", 0.0, stream, B, H, W, C);
  if (!x || !out || C % 8 != 0) return fail(DE_ERR_INVALID, "de_maxpool3x3s2: bad args");
  const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
  const long long n = (long long)B * Ho * Wo * (C / 8);
  maxpool3x3s2_kernel<<<(unsigned)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, x_dtype == DE_DTYPE_F32, reinterpret_cast<__nv_bfloat16*>(out), out_mode, B, H, W, C, Ho, Wo,
      same_pad_before(H, 3, 2), same_pad_before(W, 3, 2));
  count_launch();
  return cuda_check_launch("maxpool3x3s2_kernel");
}
