// K1 - fused MiDaS input transform (replaces /root/reference/depth_midas.py:22-23):
//   uint8 BGR HWC frame -> (BGR->RGB) -> /255 -> cv2.INTER_CUBIC resize to (net_h, net_w) ->
//   (x-mean)/std -> bf16 NCHW (or the ViT patch matrix, or fp32 NCHW for parity checks).
// The frame is read once and the tensor written once.  Kernels:
//   * preprocess_copy16_kernel - source size == net size (BASELINE configs #2/#3: 384x384 frames).  cv2.resize
//     degenerates to a copy there (and the cubic taps to [0,1,0,0]), so this is a pure streaming kernel: two items of
//     16 pixels per thread (six 16-byte loads issued up front), per channel one 32-byte run written with a single
//     256-bit store (sm_100 STG.256) - a full sector, and a full patch row of the patch-matrix layout.  The
//     128-bit-store version reached 0.62 of the HBM copy figure, this one 0.96 (profiles/r01_prepost_v2.md).
//     preprocess_copy_kernel is the generic-alignment fallback (4 pixels per thread).
//   * preprocess_kernel - general resample, horizontal pass first (the order cv2's resize uses: it is also the
//     cheap order when down-scaling, because the byte -> float conversions and the first four taps are spent on
//     net_w instead of W columns per source row).  A CTA owns 64 output columns x TH output rows.
//     (H) thread = (output column, two source rows): the column's 4 taps x 3 channels are 12 contiguous source bytes;
//     four aligned 32-bit loads + three funnel shifts bring them into three registers, 12 exact byte -> float
//     conversions and the taps run as packed fp32 pairs (FADD2 / FFMA2: B,G of a row share a tap weight, so do the R
//     values of the two rows), and the three channel values go to shared memory as fp32 [source row][RGB][64].
//     Border columns (clamped, non-contiguous taps), unaligned buffers and NV12 sources fetch per tap instead.
//     (V) thread = (output row, channel, 4 columns): four float4 reads from the staged rows, vertical taps,
//     normalise, one 8/16-byte store.  TH is chosen on the host so that the staged rows (TH * scale + 3) fit the
//     shared-memory budget; the vertical halo re-reads 3 source rows per tile (L2 hits).
//     Tried and dropped (round 1, numbers in DESIGN.md): vertical-first on raw bytes (2.7x the conversions), and a
//     persistent variant that stages the raw strip with per-row bulk async copies two tiles ahead (producer warp,
//     mbarriers) - 0.26-0.37 of the copy figure against 0.38-0.40 for this kernel: per-tile overhead and the
//     H -> V barrier, not bytes in flight, are what is left.
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
  int TH, max_rows, aligned;
  // format 1: NV12 source - src is the Y plane (row_stride / frame_stride describe it), uv the
  // interleaved half-resolution U,V plane; the BGR bytes the kernels consume are produced on the fly
  int format;
  const uint8_t* uv;
  long long uv_row_stride, uv_frame_stride;
};

constexpr int kPreThreads = 256;
#ifndef PRE_MINB
#define PRE_MINB 5
#endif
#ifndef PRE_SMEM_KB
#define PRE_SMEM_KB 40      // staged fp32 rows per CTA (768 B each): 5 CTAs / SM (measured best of 3..6)
#endif
constexpr int kPreTW = 64;                                  // output columns per CTA

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

// one source pixel as 0x00RRGGBB-style word (B in byte 0), any alignment / either source format
DEV uint32_t pixel_bgr(const PreParams& p, int b, int row, int px) {
  if (p.format == 1) return nv12_pixel_bgr(p, b, row, px);
  const uint8_t* q = p.src + (long long)b * p.frame_stride + (long long)row * p.row_stride + (long long)px * 3;
  return static_cast<uint32_t>(__ldg(q)) | (static_cast<uint32_t>(__ldg(q + 1)) << 8) | (static_cast<uint32_t>(__ldg(q + 2)) << 16);
}

// horizontal taps of one output column on one source row: a0..a2 hold the 12 bytes B0 G0 R0 B1 ... R3
// (tap t, BGR channel c at byte 3t+c); result in RGB order
DEV void htaps(uint32_t a0, uint32_t a1, uint32_t a2, const float4 w, float& r, float& g, float& bl) {
  bl = byte_to_float(a0, 0) * w.x;
  g = byte_to_float(a0, 1) * w.x;
  r = byte_to_float(a0, 2) * w.x;
  bl = fmaf(byte_to_float(a0, 3), w.y, bl);
  g = fmaf(byte_to_float(a1, 0), w.y, g);
  r = fmaf(byte_to_float(a1, 1), w.y, r);
  bl = fmaf(byte_to_float(a1, 2), w.z, bl);
  g = fmaf(byte_to_float(a1, 3), w.z, g);
  r = fmaf(byte_to_float(a2, 0), w.z, r);
  bl = fmaf(byte_to_float(a2, 1), w.w, bl);
  g = fmaf(byte_to_float(a2, 2), w.w, g);
  r = fmaf(byte_to_float(a2, 3), w.w, r);
}

// {byte kx of x, byte ky of y} -> float2, exactly
DEV float2 bytes_to_float2(uint32_t x, int kx, uint32_t y, int ky) {
  const float2 m = make_float2(__uint_as_float(__byte_perm(x, 0x4B000000u, 0x7540u | kx)),
                               __uint_as_float(__byte_perm(y, 0x4B000000u, 0x7540u | ky)));
  return fadd2(m, make_float2(-8388608.0f, -8388608.0f));
}
// htaps() for two source rows at once (a*: row A, b*: row B), same per-channel operation order, half the FMA-pipe
// instructions: (B,G) of a row share a tap weight, the two rows' R values share it too
DEV void htaps2(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t b0, uint32_t b1, uint32_t b2, const float4 w,
                float2& bgA, float2& bgB, float2& rr) {
  const float2 w0 = make_float2(w.x, w.x), w1 = make_float2(w.y, w.y), w2 = make_float2(w.z, w.z), w3 = make_float2(w.w, w.w);
  bgA = fmul2(bytes_to_float2(a0, 0, a0, 1), w0);
  bgB = fmul2(bytes_to_float2(b0, 0, b0, 1), w0);
  rr = fmul2(bytes_to_float2(a0, 2, b0, 2), w0);
  bgA = ffma2(bytes_to_float2(a0, 3, a1, 0), w1, bgA);
  bgB = ffma2(bytes_to_float2(b0, 3, b1, 0), w1, bgB);
  rr = ffma2(bytes_to_float2(a1, 1, b1, 1), w1, rr);
  bgA = ffma2(bytes_to_float2(a1, 2, a1, 3), w2, bgA);
  bgB = ffma2(bytes_to_float2(b1, 2, b1, 3), w2, bgB);
  rr = ffma2(bytes_to_float2(a2, 0, b2, 0), w2, rr);
  bgA = ffma2(bytes_to_float2(a2, 1, a2, 2), w3, bgA);
  bgB = ffma2(bytes_to_float2(b2, 1, b2, 2), w3, bgB);
  rr = ffma2(bytes_to_float2(a2, 3, b2, 3), w3, rr);
}

DEV void store_px4(const PreParams& p, int b, int ch, int oy, int ox, float4 v) {
  if (p.layout == 0) {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (((long long)b * 3 + ch) * p.net_h + oy) * p.net_w + ox;
    *reinterpret_cast<uint2*>(o) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
  } else if (p.layout == 1) {
    const int gw = p.net_w >> 4, gh = p.net_h >> 4;
    const long long prow = ((long long)b * gh + (oy >> 4)) * gw + (ox >> 4);
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + prow * 768 + ch * 256 + (oy & 15) * 16 + (ox & 15);
    *reinterpret_cast<uint2*>(o) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
  } else {
    float* o = reinterpret_cast<float*>(p.out) + (((long long)b * 3 + ch) * p.net_h + oy) * p.net_w + ox;
    *reinterpret_cast<float4*>(o) = v;
  }
}

// (V) item = (output row, RGB channel, group of 4 columns): vertical taps on the staged fp32 rows, normalise, store
template <int kThreads>
DEV void vertical_phase(const PreParams& p, const float* s_h, int b, int ox0, int oy0, int ncols, int nrows, int ys) {
  const int nitems = nrows * 48;
  for (int item = threadIdx.x; item < nitems; item += kThreads) {
    const int ry = item / 48, rem = item - ry * 48;
    const int ch = rem >> 4, cg = rem & 15;
    if (cg * 4 >= ncols) continue;
    const int oy = oy0 + ry;
    const int4 yi = __ldg(reinterpret_cast<const int4*>(p.yidx) + oy);
    const float4 wy = __ldg(reinterpret_cast<const float4*>(p.yw) + oy);
    const float* src = s_h + ch * kPreTW + cg * 4;
    const float4 t0 = *reinterpret_cast<const float4*>(src + (yi.x - ys) * (3 * kPreTW));
    const float4 t1 = *reinterpret_cast<const float4*>(src + (yi.y - ys) * (3 * kPreTW));
    const float4 t2 = *reinterpret_cast<const float4*>(src + (yi.z - ys) * (3 * kPreTW));
    const float4 t3 = *reinterpret_cast<const float4*>(src + (yi.w - ys) * (3 * kPreTW));
    const float2 sc = make_float2(p.scale[ch], p.scale[ch]), sf = make_float2(-p.shift[ch], -p.shift[ch]);
    const float2 w0 = make_float2(wy.x, wy.x), w1 = make_float2(wy.y, wy.y), w2 = make_float2(wy.z, wy.z), w3 = make_float2(wy.w, wy.w);
    float2 lo = fmul2(make_float2(t0.x, t0.y), w0), hi = fmul2(make_float2(t0.z, t0.w), w0);
    lo = ffma2(make_float2(t1.x, t1.y), w1, lo); hi = ffma2(make_float2(t1.z, t1.w), w1, hi);
    lo = ffma2(make_float2(t2.x, t2.y), w2, lo); hi = ffma2(make_float2(t2.z, t2.w), w2, hi);
    lo = ffma2(make_float2(t3.x, t3.y), w3, lo); hi = ffma2(make_float2(t3.z, t3.w), w3, hi);
    lo = ffma2(lo, sc, sf); hi = ffma2(hi, sc, sf);
    store_px4(p, b, ch, oy, ox0 + cg * 4, make_float4(lo.x, lo.y, hi.x, hi.y));
  }
}

__global__ void __launch_bounds__(kPreThreads, PRE_MINB) preprocess_kernel(const PreParams p) {
  extern __shared__ __align__(16) float s_h[];                // [max_rows][3 (RGB)][64]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z;
  const int ox0 = blockIdx.x * kPreTW, oy0 = blockIdx.y * p.TH;
  const int ncols = min(kPreTW, p.net_w - ox0), nrows = min(p.TH, p.net_h - oy0);
  const int ys = __ldg(p.yidx + oy0 * 4), ye = __ldg(p.yidx + (oy0 + nrows - 1) * 4 + 3);
  const int nsrc = ye - ys + 1;

  // ---- H: this thread's output column; warps (2k, 2k+1) share source rows k, k+4, ...
  const int col = (warp & 1) * 32 + lane;
  if ((warp & 1) * 32 < ncols) {                              // warp-uniform (net_w is a multiple of 32)
    const int4 xi = __ldg(reinterpret_cast<const int4*>(p.xidx) + ox0 + col);
    const float4 xw = __ldg(reinterpret_cast<const float4*>(p.xw) + ox0 + col);
    const int o = xi.x * 3;
    // aligned fast path: taps contiguous, word loads legal, and the 16 bytes read stay inside the buffer
    const bool lane_fast = p.format == 0 && p.aligned && xi.y == xi.x + 1 && xi.z == xi.x + 2 && xi.w == xi.x + 3 &&
                           (long long)b * p.frame_stride + (long long)ye * p.row_stride + (o & ~3) + 16 <= p.total_bytes;
    float* dst = s_h + col;
    if (__all_sync(0xffffffffu, lane_fast)) {
      const uint32_t* base = reinterpret_cast<const uint32_t*>(p.src + (long long)b * p.frame_stride + (long long)ys * p.row_stride) + (o >> 2);
      const long long rs = p.row_stride >> 2;
      const int sh = (o & 3) * 8;
      // two source rows per iteration (rows sr and sr + 4): eight independent loads in flight, packed fp32 taps
      constexpr int kStep = kPreThreads / 64;
      for (int sr = warp >> 1; sr < nsrc; sr += 2 * kStep) {
        const uint32_t* rp = base + sr * rs;
        const bool two = sr + kStep < nsrc;                        // warp-uniform
        const uint32_t* rq = two ? rp + kStep * rs : rp;
        const uint32_t u0 = __ldg(rp), u1 = __ldg(rp + 1), u2 = __ldg(rp + 2), u3 = __ldg(rp + 3);
        const uint32_t v0 = __ldg(rq), v1 = __ldg(rq + 1), v2 = __ldg(rq + 2), v3 = __ldg(rq + 3);
        float2 bgA, bgB, rr;
        htaps2(__funnelshift_r(u0, u1, sh), __funnelshift_r(u1, u2, sh), __funnelshift_r(u2, u3, sh),
               __funnelshift_r(v0, v1, sh), __funnelshift_r(v1, v2, sh), __funnelshift_r(v2, v3, sh), xw, bgA, bgB, rr);
        float* d = dst + sr * (3 * kPreTW);
        d[0] = rr.x; d[kPreTW] = bgA.y; d[2 * kPreTW] = bgA.x;
        if (two) {
          d += kStep * 3 * kPreTW;
          d[0] = rr.y; d[kPreTW] = bgB.y; d[2 * kPreTW] = bgB.x;
        }
      }
    } else {
      for (int sr = warp >> 1; sr < nsrc; sr += kPreThreads / 64) {
        const int row = ys + sr;
        const uint32_t p0 = pixel_bgr(p, b, row, xi.x), p1 = pixel_bgr(p, b, row, xi.y);
        const uint32_t p2 = pixel_bgr(p, b, row, xi.z), p3 = pixel_bgr(p, b, row, xi.w);
        float r, g, bl;
        htaps(p0 | (p1 << 24), (p1 >> 8) | (p2 << 16), (p2 >> 16) | (p3 << 8), xw, r, g, bl);
        dst[(sr * 3 + 0) * kPreTW] = r;
        dst[(sr * 3 + 1) * kPreTW] = g;
        dst[(sr * 3 + 2) * kPreTW] = bl;
      }
    }
  }
  __syncthreads();

  vertical_phase<kPreThreads>(p, s_h, b, ox0, oy0, ncols, nrows, ys);
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

#ifndef PRE_COPY_ILP
#define PRE_COPY_ILP 2
#endif
constexpr int kCopyIlp = PRE_COPY_ILP;

__global__ void __launch_bounds__(256) preprocess_copy16_kernel(const PreParams p) {
  const int nx = p.W >> 4;
  const long long nitems = (long long)p.B * p.H * nx;
  uint4 q[kCopyIlp][3];
  long long o[kCopyIlp];
  bool ok[kCopyIlp];
  const long long cstride = p.layout == 0 ? (long long)p.net_h * p.net_w : 256;   // channel stride in elements
  // all loads of the thread's items first (48 bytes each), then convert and store
#pragma unroll
  for (int k = 0; k < kCopyIlp; ++k) {
    const long long item = ((long long)blockIdx.x * kCopyIlp + k) * 256 + threadIdx.x;   // (frame row, 16-pixel group)
    ok[k] = item < nitems;
    if (!ok[k]) continue;
    const long long rowi = item / nx;
    const int x16 = static_cast<int>(item - rowi * nx);
    const int b = static_cast<int>(rowi / p.H), y = static_cast<int>(rowi - (long long)b * p.H);
    const uint4* ptr = reinterpret_cast<const uint4*>(p.src + (long long)b * p.frame_stride + (long long)y * p.row_stride) + x16 * 3;
    q[k][0] = __ldg(ptr); q[k][1] = __ldg(ptr + 1); q[k][2] = __ldg(ptr + 2);
    o[k] = p.layout == 0 ? ((long long)b * 3 * p.net_h + y) * p.net_w + x16 * 16
                         : (((long long)b * (p.net_h >> 4) + (y >> 4)) * nx + x16) * 768 + (y & 15) * 16;
  }
  __nv_bfloat16* base = reinterpret_cast<__nv_bfloat16*>(p.out);
#pragma unroll
  for (int k = 0; k < kCopyIlp; ++k) {
    if (!ok[k]) continue;
    const uint32_t u[12] = {q[k][0].x, q[k][0].y, q[k][0].z, q[k][0].w, q[k][1].x, q[k][1].y,
                            q[k][1].z, q[k][1].w, q[k][2].x, q[k][2].y, q[k][2].z, q[k][2].w};
    uint4 lo, hi;
    copy16_channel<0>(u, p.scale[0], p.shift[0], lo, hi);
    st_stream_256(base + o[k], lo, hi);
    copy16_channel<1>(u, p.scale[1], p.shift[1], lo, hi);
    st_stream_256(base + o[k] + cstride, lo, hi);
    copy16_channel<2>(u, p.scale[2], p.shift[2], lo, hi);
    st_stream_256(base + o[k] + 2 * cstride, lo, hi);
  }
}

}  // namespace deb

static int run_preprocess(const uint8_t* bgr, const uint8_t* uv, int64_t uv_row_stride, int64_t uv_frame_stride, int B,
                          int H, int W, int64_t row_stride, int64_t frame_stride, int net_h, int net_w,
                          const float* mean_host, const float* std_host, int layout, void* out, void* stream) {
  deb::ProfScope prof_("preprocess", 0.0, stream, B, H, W, (int64_t)net_h * net_w);
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
                        reinterpret_cast<uintptr_t>(out)) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 31) == 0;
    if (wide) {
      const long long items = (long long)B * H * (W / 16);
      preprocess_copy16_kernel<<<(unsigned)((items + 256 * kCopyIlp - 1) / (256 * kCopyIlp)), 256, 0, st>>>(p);
    } else {
      dim3 grid((W / 4 + 127) / 128, H, B);
      preprocess_copy_kernel<<<grid, 128, 0, st>>>(p);
    }
    count_launch();
    return cuda_check_launch("preprocess_copy_kernel");
  }
  // largest power-of-two tile height whose staged source rows fit the shared-memory budget (TH = 1 needs 4 rows)
  p.TH = 32;
  while (p.TH > 1 && (size_t)ty->max_extent(p.TH) * 3 * kPreTW * sizeof(float) > (size_t)PRE_SMEM_KB * 1024) p.TH >>= 1;
  p.max_rows = ty->max_extent(p.TH);
  const size_t smem = (size_t)p.max_rows * 3 * kPreTW * sizeof(float);
  {
    static SmemOptIn optin;
    cudaError_t e = ensure_smem(optin, preprocess_kernel, PRE_SMEM_KB * 1024);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(preprocess)");
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
