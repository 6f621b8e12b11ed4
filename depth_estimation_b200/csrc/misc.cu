// Bandwidth-bound helper kernels between the tensor-core GEMMs: LayerNorm, token casts, cls fill,
// bilinear x2 upsample (NHWC), strided 3x3 im2col.  All are 16-byte vectorised and grid-stride free
// (one element group per thread, grids far larger than 148 SMs at the BASELINE shapes).
#include <mutex>
#include <stdlib.h>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

// ---------------------------------------------------------------- (residual add +) LayerNorm, one warp per row
// x (fp32 token stream) is optionally updated in place with x += y (the branch output a GEMM just wrote),
// then normalised to bf16.  Folding the residual add here keeps the GEMM epilogues load-free (TMA-store
// path) and hides the extra read behind the massive parallelism of a bandwidth kernel.
template <int VEC>  // VEC float4 per lane: dim = 128*VEC
__global__ void __launch_bounds__(256) layernorm_kernel(float* __restrict__ x, long long ldx, const void* __restrict__ y,
                                                        int y_bf16, long long ldy, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, int rows, float eps,
                                                        __nv_bfloat16* __restrict__ out, long long ldo, int split) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  pdl_launch();
  pdl_wait();
  if (row >= rows) return;
  float4* xr = reinterpret_cast<float4*>(x + (long long)row * ldx);
  float4 v[VEC];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < VEC; ++i) v[i] = xr[lane + i * 32];
  if (y && y_bf16) {                               // branch output stored as bf16 by the GEMM epilogue
    const uint2* yr = reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(y) + (long long)row * ldy);
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
      const uint2 t = yr[lane + i * 32];
      v[i].x += bf16_lo(t.x); v[i].y += bf16_hi(t.x); v[i].z += bf16_lo(t.y); v[i].w += bf16_hi(t.y);
      xr[lane + i * 32] = v[i];
    }
  } else if (y) {
    const float4* yr = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(y) + (long long)row * ldy);
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
      const float4 t = yr[lane + i * 32];
      v[i].x += t.x; v[i].y += t.y; v[i].z += t.z; v[i].w += t.w;
      xr[lane + i * 32] = v[i];
    }
  }
  if (!out) return;
#pragma unroll
  for (int i = 0; i < VEC; ++i) sum += v[i].x + v[i].y + v[i].z + v[i].w;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float mean = sum / (128.f * VEC);
  float var = 0.f;
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
    var += a * a + b * b + c * c + d * d;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) var += __shfl_xor_sync(0xffffffffu, var, o);
  const float rstd = rsqrtf(var / (128.f * VEC) + eps);
  uint2* orow = reinterpret_cast<uint2*>(out + (long long)row * ldo);
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + lane + i * 32);
    const float4 bb = __ldg(reinterpret_cast<const float4*>(beta) + lane + i * 32);
    const float y0 = (v[i].x - mean) * rstd * g.x + bb.x, y1 = (v[i].y - mean) * rstd * g.y + bb.y;
    const float y2 = (v[i].z - mean) * rstd * g.z + bb.z, y3 = (v[i].w - mean) * rstd * g.w + bb.w;
    uint2 o;
    o.x = pack_bf16(y0, y1);
    o.y = pack_bf16(y2, y3);
    orow[lane + i * 32] = o;
    if (split) {                                   // [hi | hi | lo], each 128*VEC wide (bf16x3 operand)
      orow[32 * VEC + lane + i * 32] = o;
      uint2 lo;
      lo.x = pack_bf16(y0 - bf16_lo(o.x), y1 - bf16_hi(o.x));
      lo.y = pack_bf16(y2 - bf16_lo(o.y), y3 - bf16_hi(o.y));
      orow[64 * VEC + lane + i * 32] = lo;
    }
  }
}

// ---------------------------------------------------------------- fp32 -> bf16 token cast (drop cls)
__global__ void cast_tokens_kernel(const float* __restrict__ x, long long ldx, long long out_rows, int dim, int grp,
                                   int skip, __nv_bfloat16* __restrict__ out, int split) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one 8-element group per thread
  const int per_row = dim >> 3;
  if (i >= out_rows * per_row) return;
  const long long orow = i / per_row;
  const int c = static_cast<int>(i % per_row) * 8;
  long long irow = orow;
  if (grp > 0) irow = (orow / (grp - skip)) * grp + skip + (orow % (grp - skip));
  const float4 a = *reinterpret_cast<const float4*>(x + irow * ldx + c);
  const float4 b = *reinterpret_cast<const float4*>(x + irow * ldx + c + 4);
  uint4 o;
  o.x = pack_bf16(a.x, a.y); o.y = pack_bf16(a.z, a.w); o.z = pack_bf16(b.x, b.y); o.w = pack_bf16(b.z, b.w);
  if (!split) {
    *reinterpret_cast<uint4*>(out + orow * dim + c) = o;
    return;
  }
  uint4 lo;
  lo.x = pack_bf16(a.x - bf16_lo(o.x), a.y - bf16_hi(o.x)); lo.y = pack_bf16(a.z - bf16_lo(o.y), a.w - bf16_hi(o.y));
  lo.z = pack_bf16(b.x - bf16_lo(o.z), b.y - bf16_hi(o.z)); lo.w = pack_bf16(b.z - bf16_lo(o.w), b.w - bf16_hi(o.w));
  __nv_bfloat16* op = out + orow * 3 * dim + c;
  *reinterpret_cast<uint4*>(op) = o;
  *reinterpret_cast<uint4*>(op + dim) = o;
  *reinterpret_cast<uint4*>(op + 2 * dim) = lo;
}

// fp32 NCHW network input [B,3,h,w] (K1 layout 2) -> split3 ViT patch matrix [B*(h/16)*(w/16), 3*768] = [hi | hi | lo],
// column = c*256 + py*16 + px - the bf16x3 form of K1's layout 1.  One thread per (patch, channel, patch row): 16 pixels.
__global__ void patchify_split3_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int B, int h, int w) {
  const int gw = w >> 4, gh = h >> 4;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)B * gh * gw * 48) return;
  const int py = static_cast<int>(i % 16), c = static_cast<int>((i / 16) % 3);
  const long long patch = i / 48;
  const int gx = static_cast<int>(patch % gw), gy = static_cast<int>((patch / gw) % gh), b = static_cast<int>(patch / ((long long)gw * gh));
  const float* src = x + (((long long)b * 3 + c) * h + gy * 16 + py) * w + gx * 16;
  __nv_bfloat16* dst = out + patch * (3 * 768) + c * 256 + py * 16;
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    const float4 a = *reinterpret_cast<const float4*>(src + 8 * v), bq = *reinterpret_cast<const float4*>(src + 8 * v + 4);
    uint4 o, lo;
    o.x = pack_bf16(a.x, a.y); o.y = pack_bf16(a.z, a.w); o.z = pack_bf16(bq.x, bq.y); o.w = pack_bf16(bq.z, bq.w);
    lo.x = pack_bf16(a.x - bf16_lo(o.x), a.y - bf16_hi(o.x)); lo.y = pack_bf16(a.z - bf16_lo(o.y), a.w - bf16_hi(o.y));
    lo.z = pack_bf16(bq.x - bf16_lo(o.z), bq.y - bf16_hi(o.z)); lo.w = pack_bf16(bq.z - bf16_lo(o.w), bq.w - bf16_hi(o.w));
    *reinterpret_cast<uint4*>(dst + 8 * v) = o;
    *reinterpret_cast<uint4*>(dst + 768 + 8 * v) = o;
    *reinterpret_cast<uint4*>(dst + 1536 + 8 * v) = lo;
  }
}

__global__ void fill_cls_kernel(float* __restrict__ x, long long ldx, int B, int T, int dim,
                                const float* __restrict__ cls, const float* __restrict__ pos) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * dim) return;
  const int b = i / dim, c = i % dim;
  x[(long long)b * T * ldx + c] = cls[c] + pos[c];
}

// ---------------------------------------------------------------- bilinear x2, NHWC bf16
// grid = (ceil(OW*cg/256), OH, B): the row and the image come from the block index and the column from one 32-bit
// division (a shift when C/8 is a power of two), instead of four 64-bit div/mod chains per thread - this kernel
// moves 2.5 GB per DPT_Large batch-32 step and used to spend most of its time on index arithmetic.
// Each thread owns one (ox, 8-channel group) column position and kUpRows consecutive output rows: the horizontal
// lerp of a source row is computed once (fp32, 8 channels) and reused by every output row that touches it - 2x
// upsampling needs at most kUpRows/2 + 2 source rows for kUpRows output rows - so a thread issues ~6 16-byte loads
// and 4 16-byte streaming stores instead of 4 loads per store, with a quarter of the index arithmetic.
constexpr int kUpRows = 4;
// kV = 16-byte vectors (8 channels each) per thread.  kV = 2 (C % 16 == 0): 32 bytes per output pixel and thread,
// written with one 256-bit streaming store (a full sector) - the write side is 4/5 of this kernel's traffic.
template <int kV>
DEV void up_hrow(const __nv_bfloat16* rowp, int x0off, int x1off, float hx, float lx, float* h) {
#pragma unroll
  for (int v = 0; v < kV; ++v) {
    const uint4 a = *reinterpret_cast<const uint4*>(rowp + x0off + 8 * v);
    const uint4 b = *reinterpret_cast<const uint4*>(rowp + x1off + 8 * v);
    const uint32_t* pa = &a.x;
    const uint32_t* pb = &b.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      h[8 * v + 2 * k] = hx * bf16_lo(pa[k]) + lx * bf16_lo(pb[k]);
      h[8 * v + 2 * k + 1] = hx * bf16_hi(pa[k]) + lx * bf16_hi(pb[k]);
    }
  }
}
template <int kV>
__global__ void __launch_bounds__(256) upsample2x_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                         int B, int H, int W, int C, int align_corners, int cg_shift) {
  constexpr int kCh = 8 * kV;
  const int cg = C / kCh;
  const int OH = 2 * H, OW = 2 * W;
  const unsigned idx = blockIdx.x * 256u + threadIdx.x;          // (ox, channel group) inside the row
  const int ox = cg_shift >= 0 ? static_cast<int>(idx >> cg_shift) : static_cast<int>(idx / (unsigned)cg);
  if (ox >= OW) return;
  const int c = static_cast<int>(idx - (unsigned)ox * (unsigned)cg) * kCh;
  const int oy0 = blockIdx.y * kUpRows, b = blockIdx.z;
  // ATen area_pixel_compute_scale: (in-1)/(out-1) in float, src = scale*dst (align_corners); else 0.5*(dst+0.5)-0.5 >= 0
  const float rh = OH > 1 ? (float)(H - 1) / (float)(OH - 1) : 0.f;
  const float rw = OW > 1 ? (float)(W - 1) / (float)(OW - 1) : 0.f;
  const float sx = align_corners ? rw * ox : fmaxf(0.5f * (ox + 0.5f) - 0.5f, 0.f);
  const int x0 = min(static_cast<int>(sx), W - 1), x1 = min(x0 + 1, W - 1);
  const float lx = sx - x0, hx = 1.f - lx;
  const __nv_bfloat16* base = x + (long long)b * H * W * C + c;
  const int x0off = x0 * C, x1off = x1 * C;
  float h0[kCh], h1[kCh];
  int cy0 = -1, cy1 = -1;                                        // source rows currently held in h0 / h1
  __nv_bfloat16* op = out + (((long long)b * OH + oy0) * OW + ox) * C + c;
#pragma unroll
  for (int r = 0; r < kUpRows; ++r) {
    const int oy = oy0 + r;
    if (oy >= OH) break;
    const float sy = align_corners ? rh * oy : fmaxf(0.5f * (oy + 0.5f) - 0.5f, 0.f);
    const int y0 = min(static_cast<int>(sy), H - 1), y1 = min(y0 + 1, H - 1);
    const float ly = sy - y0, hy = 1.f - ly;
    if (y0 != cy0) {                                             // block-uniform branches (oy is per block)
      if (y0 == cy1) {
#pragma unroll
        for (int j = 0; j < kCh; ++j) h0[j] = h1[j];
      } else {
        up_hrow<kV>(base + (long long)y0 * W * C, x0off, x1off, hx, lx, h0);
      }
      cy0 = y0;
    }
    if (y1 != cy1) {
      if (y1 == cy0) {
#pragma unroll
        for (int j = 0; j < kCh; ++j) h1[j] = h0[j];
      } else {
        up_hrow<kV>(base + (long long)y1 * W * C, x0off, x1off, hx, lx, h1);
      }
      cy1 = y1;
    }
    uint32_t o[4 * kV];
#pragma unroll
    for (int k = 0; k < 4 * kV; ++k)
      o[k] = pack_bf16(hy * h0[2 * k] + ly * h1[2 * k], hy * h0[2 * k + 1] + ly * h1[2 * k + 1]);
    __nv_bfloat16* dst = op + (long long)r * OW * C;
    if constexpr (kV == 2) st_stream_256(dst, make_uint4(o[0], o[1], o[2], o[3]), make_uint4(o[4], o[5], o[6], o[7]));
    else __stcs(reinterpret_cast<uint4*>(dst), make_uint4(o[0], o[1], o[2], o[3]));
  }
}

// ---------------------------------------------------------------- im2col 3x3, stride s, explicit pad
__global__ void im2col3x3_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int B, int H,
                                 int W, int C, int stride, int pad_y, int pad_x, int Ho, int Wo) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int cg = C >> 3;
  if (i >= (long long)B * Ho * Wo * 9 * cg) return;
  const int c = static_cast<int>(i % cg) * 8;
  long long t = i / cg;
  const int tap = static_cast<int>(t % 9);
  t /= 9;
  const int ox = static_cast<int>(t % Wo);
  t /= Wo;
  const int oy = static_cast<int>(t % Ho);
  const int b = static_cast<int>(t / Ho);
  const int iy = oy * stride + tap / 3 - pad_y, ix = ox * stride + tap % 3 - pad_x;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (iy >= 0 && iy < H && ix >= 0 && ix < W)
    v = *reinterpret_cast<const uint4*>(x + (((long long)b * H + iy) * W + ix) * C + c);
  *reinterpret_cast<uint4*>(out + (((long long)b * Ho + oy) * Wo + ox) * 9 * C + (long long)tap * C + c) = v;
}

// ---------------------------------------------------------------- spatial subsample (1x1 stride-2 conv input)
__global__ void subsample_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int B, int H,
                                 int W, int C, int stride, int Ho, int Wo) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int cg = C >> 3;
  if (i >= (long long)B * Ho * Wo * cg) return;
  const int c = static_cast<int>(i % cg) * 8;
  long long t = i / cg;
  const int ox = static_cast<int>(t % Wo);
  t /= Wo;
  const int oy = static_cast<int>(t % Ho);
  const int b = static_cast<int>(t / Ho);
  *reinterpret_cast<uint4*>(out + (((long long)b * Ho + oy) * Wo + ox) * C + c) =
      *reinterpret_cast<const uint4*>(x + (((long long)b * H + oy * stride) * W + ox * stride) * C + c);
}

// ---------------------------------------------------------------- GroupNorm (NHWC bf16), deterministic
// pass 1: per-block partial (sum, sumsq) per group, fixed-order reductions (no atomics: bf16 storage makes
//         last-bit noise visible as rounding flips, so the statistics must be bit-reproducible)
// pass 2: fixed-order sum over blocks -> (mean, rstd)
// pass 3: y = act((x-mean)*rstd*gamma + beta [+ add]) -> bf16 (and optionally an fp32 copy)
// 8 consecutive channels of a bf16 or fp32 NHWC tensor as floats
DEV void load8(const void* x, int x_f32, long long elem, float* f) {
  if (x_f32) {
    const float* ptr = reinterpret_cast<const float*>(x) + elem;
    if ((reinterpret_cast<uintptr_t>(ptr) & 31) == 0) {        // one 256-bit load: a full sector per lane (sm_100 LDG.256)
      uint32_t r[8];
      asm volatile("ld.global.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                   : "l"(ptr));
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(r[j]);
    } else {
      const float4 u0 = *reinterpret_cast<const float4*>(ptr);
      const float4 u1 = *reinterpret_cast<const float4*>(ptr + 4);
      f[0] = u0.x; f[1] = u0.y; f[2] = u0.z; f[3] = u0.w; f[4] = u1.x; f[5] = u1.y; f[6] = u1.z; f[7] = u1.w;
    }
  } else {
    const uint4 t = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(x) + elem);
    f[0] = bf16_lo(t.x); f[1] = bf16_hi(t.x); f[2] = bf16_lo(t.y); f[3] = bf16_hi(t.y);
    f[4] = bf16_lo(t.z); f[5] = bf16_hi(t.z); f[6] = bf16_lo(t.w); f[7] = bf16_hi(t.w);
  }
}
// "split3" operand layout of the precise (bf16x3) convolutions: a row of C fp32 values becomes 3C bf16 values
// [hi | hi | lo], hi = bf16(v), lo = bf16(v - hi).  Against weights packed [Wh | Wl | Wh] along K, one ordinary bf16
// GEMM then accumulates Ah*Wh + Ah*Wl + Al*Wh in fp32 - ~16 mantissa bits instead of 8 (weights.split3_*).
DEV void store_split3(__nv_bfloat16* out, long long row, int C, int c0, const float* f) {
  uint32_t hi[4], lo[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    hi[k] = pack_bf16(f[2 * k], f[2 * k + 1]);
    lo[k] = pack_bf16(f[2 * k] - bf16_lo(hi[k]), f[2 * k + 1] - bf16_hi(hi[k]));
  }
  __nv_bfloat16* o = out + row * 3 * C + c0;
  *reinterpret_cast<uint4*>(o) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
  *reinterpret_cast<uint4*>(o + C) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
  *reinterpret_cast<uint4*>(o + 2 * C) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
}

// Bilinear x2 for the bf16x3 ("split3") decoder: fp32 or bf16 NHWC in; fp32 NHWC, plain bf16 or split3 [hi | hi | lo] out.
// One thread per (output pixel, 8 channels); same coordinate rule as upsample2x_kernel.
__global__ void __launch_bounds__(256) upsample2x_ex_kernel(const void* __restrict__ x, int x_f32, void* __restrict__ out,
                                                            int out_mode, int B, int H, int W, int C, int align_corners) {
  const int cg = C >> 3;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int OH = 2 * H, OW = 2 * W;
  if (i >= (long long)B * OH * OW * cg) return;
  const int c = static_cast<int>(i % cg) * 8;
  const long long pix = i / cg;
  const int ox = static_cast<int>(pix % OW);
  const int oy = static_cast<int>((pix / OW) % OH);
  const int b = static_cast<int>(pix / ((long long)OW * OH));
  const float rh = OH > 1 ? (float)(H - 1) / (float)(OH - 1) : 0.f;
  const float rw = OW > 1 ? (float)(W - 1) / (float)(OW - 1) : 0.f;
  const float sx = align_corners ? rw * ox : fmaxf(0.5f * (ox + 0.5f) - 0.5f, 0.f);
  const float sy = align_corners ? rh * oy : fmaxf(0.5f * (oy + 0.5f) - 0.5f, 0.f);
  const int x0 = min(static_cast<int>(sx), W - 1), x1 = min(x0 + 1, W - 1);
  const int y0 = min(static_cast<int>(sy), H - 1), y1 = min(y0 + 1, H - 1);
  const float lx = sx - x0, hx = 1.f - lx, ly = sy - y0, hy = 1.f - ly;
  const long long base = (long long)b * H * W;
  float a[8], bb[8], cc[8], d[8], f[8];
  load8(x, x_f32, ((base + (long long)y0 * W + x0) * C) + c, a);
  load8(x, x_f32, ((base + (long long)y0 * W + x1) * C) + c, bb);
  load8(x, x_f32, ((base + (long long)y1 * W + x0) * C) + c, cc);
  load8(x, x_f32, ((base + (long long)y1 * W + x1) * C) + c, d);
#pragma unroll
  for (int j = 0; j < 8; ++j) f[j] = hy * (hx * a[j] + lx * bb[j]) + ly * (hx * cc[j] + lx * d[j]);
  if (out_mode == 2) {
    float* o = reinterpret_cast<float*>(out) + pix * C + c;
    *reinterpret_cast<float4*>(o) = make_float4(f[0], f[1], f[2], f[3]);
    *reinterpret_cast<float4*>(o + 4) = make_float4(f[4], f[5], f[6], f[7]);
  } else if (out_mode == 1) {
    store_split3(reinterpret_cast<__nv_bfloat16*>(out), pix, C, c, f);
  } else {
    *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(out) + pix * C + c) =
        make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
  }
}

// Pass 1 of 2: per-CTA partial sums; the LAST CTA of a sample to finish (arrival counter in the workspace, left at
// zero again) reduces that sample's partials in a fixed order - one warp per group, lane-strided double sums and a
// shuffle tree - and writes (mean, rstd).  Deterministic, and no separate finalize launch: at batch 1 the old
// single-thread-per-group finalize kernel cost 18 us per GroupNorm, a quarter of DPT_Hybrid's frame latency.
__global__ void __launch_bounds__(256) groupnorm_partial_kernel(const void* __restrict__ x, int x_f32, int HW, int C,
                                                                int groups, int px_per_block,
                                                                double* __restrict__ partial, unsigned* __restrict__ counters,
                                                                float2* __restrict__ stats, double n, float eps) {
  __shared__ __align__(16) float s_su[256 * 8], s_sq[256 * 8];
  __shared__ bool s_last;
  const int b = blockIdx.y;
  const int vec = C >> 3;                    // 8-channel vectors per pixel
  const int lanes = 256 / vec;               // pixels processed concurrently
  const int v = threadIdx.x % vec, pl = threadIdx.x / vec;
  const int p0 = blockIdx.x * px_per_block;
  const int p1 = min(p0 + px_per_block, HW);
  float su[8], sq[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) su[j] = sq[j] = 0.f;
  // four pixels per thread in flight (the one-load-at-a-time loop ran at 0.55 of the HBM copy figure at batch 32)
  for (int p = p0 + pl; p < p1; p += 4 * lanes) {
    float f[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (p + u * lanes < p1) load8(x, x_f32, ((long long)b * HW + p + u * lanes) * C + v * 8, f[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (p + u * lanes < p1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { su[j] += f[u][j]; sq[j] = fmaf(f[u][j], f[u][j], sq[j]); }
      }
  }
  // s_[pl][channel]
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    s_su[pl * C + v * 8 + j] = su[j];
    s_sq[pl * C + v * 8 + j] = sq[j];
  }
  __syncthreads();
  if (threadIdx.x < groups) {
    const int cpg = C / groups, g = threadIdx.x;
    double a = 0.0, q = 0.0;
    for (int c = g * cpg; c < (g + 1) * cpg; ++c)
      for (int l = 0; l < lanes; ++l) {
        a += (double)s_su[l * C + c];
        q += (double)s_sq[l * C + c];
      }
    double* o = partial + (((long long)b * gridDim.x + blockIdx.x) * groups + g) * 2;
    o[0] = a;
    o[1] = q;
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(counters + b, 1u);
    s_last = prev == gridDim.x - 1;
    if (s_last) counters[b] = 0;                 // every CTA of this sample has arrived: ready for the next call
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // thread = (group g, slice sl): sums the partials of CTAs sl, sl + slices, ... (independent, coalesced 16-byte loads -
  // a warp reads 32 consecutive groups of one CTA), then one thread per group adds the slices in a fixed order.
  // (The first version gave each warp whole groups: 5 dependent load rounds per group, ~5 us at batch 1.)
  double2* s_red = reinterpret_cast<double2*>(s_su);          // [slices][groups] - 4 KB of the 8 KB s_su array
  const int nblocks = gridDim.x;
  const int slices = 256 / groups;                             // groups <= 64 (checked on the host)
  const int g = threadIdx.x % groups, sl = threadIdx.x / groups;
  if (sl < slices) {
    double a = 0.0, q = 0.0;
    const double2* pp = reinterpret_cast<const double2*>(partial) + (long long)b * nblocks * groups + g;
#pragma unroll 4
    for (int k = sl; k < nblocks; k += slices) {
      const double2 v = __ldcg(pp + (long long)k * groups);
      a += v.x;
      q += v.y;
    }
    s_red[sl * groups + g] = make_double2(a, q);
  }
  __syncthreads();
  if (threadIdx.x < groups) {
    double a = 0.0, q = 0.0;
    for (int i = 0; i < slices; ++i) {
      const double2 v = s_red[i * groups + threadIdx.x];
      a += v.x;
      q += v.y;
    }
    const double m = a / n;
    double var = q / n - m * m;
    if (var < 0) var = 0;
    stats[b * groups + threadIdx.x] = make_float2((float)m, (float)(1.0 / sqrt(var + (double)eps)));
  }
}

__global__ void groupnorm_apply_kernel(const void* __restrict__ x, int x_f32, const float2* __restrict__ stats,
                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                       const void* __restrict__ add, int add_dtype, __nv_bfloat16* __restrict__ out,
                                       int out_mode, float* __restrict__ out_f32, __nv_bfloat16* __restrict__ out_plain,
                                       int B, int HW, int C, int groups, int act) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int vec = C >> 3;
  if (i >= (long long)B * HW * vec) return;
  const int c0 = static_cast<int>(i % vec) * 8;
  const int b = static_cast<int>(i / ((long long)HW * vec));
  const int cpg = C / groups;
  float f[8];
  load8(x, x_f32, i * 8, f);
  float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (add) {
    if (add_dtype == DE_DTYPE_F32) {
      const float4 u0 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(add) + i * 8);
      const float4 u1 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(add) + i * 8 + 4);
      a[0] = u0.x; a[1] = u0.y; a[2] = u0.z; a[3] = u0.w; a[4] = u1.x; a[5] = u1.y; a[6] = u1.z; a[7] = u1.w;
    } else {
      const uint4 u = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(add) + i * 8);
      a[0] = bf16_lo(u.x); a[1] = bf16_hi(u.x); a[2] = bf16_lo(u.y); a[3] = bf16_hi(u.y);
      a[4] = bf16_lo(u.z); a[5] = bf16_hi(u.z); a[6] = bf16_lo(u.w); a[7] = bf16_hi(u.w);
    }
  }
  int gprev = -1;
  float2 st = make_float2(0.f, 0.f);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = (c0 + j) / cpg;
    if (g != gprev) {
      st = __ldg(stats + b * groups + g);
      gprev = g;
    }
    float y = (f[j] - st.x) * st.y * __ldg(gamma + c0 + j) + __ldg(beta + c0 + j) + a[j];
    if (act == DE_ACT_RELU) y = fmaxf(y, 0.f);
    f[j] = y;
  }
  const uint4 plain = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
  if (out) {
    if (out_mode == 1) store_split3(out, i / vec, C, c0, f);
    else *reinterpret_cast<uint4*>(out + i * 8) = plain;
  }
  if (out_plain) *reinterpret_cast<uint4*>(out_plain + i * 8) = plain;
  if (out_f32) {
    *reinterpret_cast<float4*>(out_f32 + i * 8) = make_float4(f[0], f[1], f[2], f[3]);
    *reinterpret_cast<float4*>(out_f32 + i * 8 + 4) = make_float4(f[4], f[5], f[6], f[7]);
  }
}

// One-launch GroupNorm for grids that are co-resident on the GPU (the batch-1 latency path: 49 GroupNorms per DPT_Hybrid
// frame, two launches each before).  Same partial sums, same fixed-order reduction - hence the same (mean, rstd) bits -
// as the two-kernel path, but every CTA keeps its pixels in shared memory (fp32), waits on the sample's arrival counter
// until all partials are published, reduces them itself and normalises straight out of shared memory: x is read once,
// and the second launch, its tail and its launch gap are gone.  The spin is safe because the host only picks this
// kernel when gridDim.x * gridDim.y CTAs fit on the device at once (cudaOccupancy...), so every CTA the waiters depend
// on is running or will be scheduled as soon as an unrelated kernel frees an SM.  Counters: arrive[b] at counters[b],
// depart[b] at counters[2048 + b]; the last CTA to leave resets both.
__global__ void __launch_bounds__(256) groupnorm_fused_kernel(const void* __restrict__ x, int x_f32, int HW, int C, int groups,
                                                              int px_per_block, double* __restrict__ partial,
                                                              unsigned* __restrict__ counters, double n, float eps,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              const void* __restrict__ add, int add_dtype,
                                                              __nv_bfloat16* __restrict__ out, int out_mode,
                                                              float* __restrict__ out_f32, __nv_bfloat16* __restrict__ out_plain,
                                                              int act) {
  extern __shared__ __align__(16) float gn_tile[];             // [px_per_block][C] fp32
  __shared__ __align__(16) float s_su[256 * 8], s_sq[256 * 8];
  __shared__ float2 s_stats[64];
  __shared__ bool s_last;
  const int b = blockIdx.y;
  const int vec = C >> 3, lanes = 256 / vec;
  const int v = threadIdx.x % vec, pl = threadIdx.x / vec;
  const int p0 = blockIdx.x * px_per_block;
  const int p1 = min(p0 + px_per_block, HW);
  float su[8], sq[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) su[j] = sq[j] = 0.f;
  for (int p = p0 + pl; p < p1; p += 4 * lanes) {
    float f[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (p + u * lanes < p1) load8(x, x_f32, ((long long)b * HW + p + u * lanes) * C + v * 8, f[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (p + u * lanes < p1) {
        float* t = gn_tile + (size_t)(p + u * lanes - p0) * C + v * 8;
        *reinterpret_cast<float4*>(t) = make_float4(f[u][0], f[u][1], f[u][2], f[u][3]);
        *reinterpret_cast<float4*>(t + 4) = make_float4(f[u][4], f[u][5], f[u][6], f[u][7]);
#pragma unroll
        for (int j = 0; j < 8; ++j) { su[j] += f[u][j]; sq[j] = fmaf(f[u][j], f[u][j], sq[j]); }
      }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    s_su[pl * C + v * 8 + j] = su[j];
    s_sq[pl * C + v * 8 + j] = sq[j];
  }
  __syncthreads();
  if (threadIdx.x < groups) {
    const int cpg = C / groups, g = threadIdx.x;
    double a = 0.0, q = 0.0;
    for (int c = g * cpg; c < (g + 1) * cpg; ++c)
      for (int l = 0; l < lanes; ++l) {
        a += (double)s_su[l * C + c];
        q += (double)s_sq[l * C + c];
      }
    double* o = partial + (((long long)b * gridDim.x + blockIdx.x) * groups + g) * 2;
    o[0] = a;
    o[1] = q;
    __threadfence();
  }
  __syncthreads();
  const unsigned nblocks = gridDim.x;
  if (threadIdx.x == 0) {
    atomicAdd(counters + b, 1u);
    unsigned spins = 0;
    while (*reinterpret_cast<volatile unsigned*>(counters + b) < nblocks) {
      __nanosleep(64);
      if (++spins > (1u << 24)) __trap();                      // ~1 s: a protocol bug must not hang the GPU
    }
    __threadfence();
  }
  __syncthreads();
  // every CTA reduces the sample's partials in the same fixed order as the two-kernel path's last CTA
  {
    double2* s_red = reinterpret_cast<double2*>(s_su);
    const int slices = 256 / groups;
    const int g = threadIdx.x % groups, sl = threadIdx.x / groups;
    if (sl < slices) {
      double a = 0.0, q = 0.0;
      const double2* pp = reinterpret_cast<const double2*>(partial) + (long long)b * nblocks * groups + g;
      // all of the thread's partials are requested before the first add (one L2 round trip instead of one per
      // unrolled batch: ~1.2 us of every batch-1 GroupNorm); the adds run in the same order as before
      constexpr int kAhead = 20;                                // 148 CTAs / 8 slices = 19 partials per thread
      double2 t[kAhead];
#pragma unroll
      for (int i = 0; i < kAhead; ++i) {
        const int k = sl + i * slices;
        t[i] = k < (int)nblocks ? __ldcg(pp + (long long)k * groups) : make_double2(0.0, 0.0);
      }
#pragma unroll
      for (int i = 0; i < kAhead; ++i)
        if (sl + i * slices < (int)nblocks) {
          a += t[i].x;
          q += t[i].y;
        }
      for (int k = sl + kAhead * slices; k < (int)nblocks; k += slices) {
        const double2 u = __ldcg(pp + (long long)k * groups);
        a += u.x;
        q += u.y;
      }
      s_red[sl * groups + g] = make_double2(a, q);
    }
    __syncthreads();
    if (threadIdx.x < groups) {
      double a = 0.0, q = 0.0;
      for (int i = 0; i < slices; ++i) {
        const double2 t = s_red[i * groups + threadIdx.x];
        a += t.x;
        q += t.y;
      }
      const double m = a / n;
      double var = q / n - m * m;
      if (var < 0) var = 0;
      s_stats[threadIdx.x] = make_float2((float)m, (float)(1.0 / sqrt(var + (double)eps)));
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {                                      // done with the partials and the arrival counter
    const unsigned prev = atomicAdd(counters + 2048 + b, 1u);
    if (prev == nblocks - 1) {
      counters[b] = 0;
      counters[2048 + b] = 0;
    }
  }
  // normalise out of shared memory
  const int cpg = C / groups, c0 = v * 8;
  float ga[8], be[8];
  float2 st[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    ga[j] = __ldg(gamma + c0 + j);
    be[j] = __ldg(beta + c0 + j);
    st[j] = s_stats[(c0 + j) / cpg];
  }
  for (int p = p0 + pl; p < p1; p += lanes) {
    const long long i8 = ((long long)b * HW + p) * C + c0;
    const float* t = gn_tile + (size_t)(p - p0) * C + c0;
    const float4 t0 = *reinterpret_cast<const float4*>(t), t1 = *reinterpret_cast<const float4*>(t + 4);
    float f[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (add) {
      if (add_dtype == DE_DTYPE_F32) {
        const float4 u0 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(add) + i8);
        const float4 u1 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(add) + i8 + 4);
        a[0] = u0.x; a[1] = u0.y; a[2] = u0.z; a[3] = u0.w; a[4] = u1.x; a[5] = u1.y; a[6] = u1.z; a[7] = u1.w;
      } else {
        const uint4 u = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(add) + i8);
        a[0] = bf16_lo(u.x); a[1] = bf16_hi(u.x); a[2] = bf16_lo(u.y); a[3] = bf16_hi(u.y);
        a[4] = bf16_lo(u.z); a[5] = bf16_hi(u.z); a[6] = bf16_lo(u.w); a[7] = bf16_hi(u.w);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float y = (f[j] - st[j].x) * st[j].y * ga[j] + be[j] + a[j];
      if (act == DE_ACT_RELU) y = fmaxf(y, 0.f);
      f[j] = y;
    }
    const uint4 plain = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
    if (out) {
      if (out_mode == 1) store_split3(out, (long long)b * HW + p, C, c0, f);
      else *reinterpret_cast<uint4*>(out + i8) = plain;
    }
    if (out_plain) *reinterpret_cast<uint4*>(out_plain + i8) = plain;
    if (out_f32) {
      *reinterpret_cast<float4*>(out_f32 + i8) = make_float4(f[0], f[1], f[2], f[3]);
      *reinterpret_cast<float4*>(out_f32 + i8 + 4) = make_float4(f[4], f[5], f[6], f[7]);
    }
  }
}

}  // namespace deb

using namespace deb;

static int layernorm_impl(float* x, int64_t ldx, const void* y, int y_dtype, int64_t ldy, const float* gamma,
                          const float* beta, int rows, int dim, float eps, void* out_bf16, int64_t ldo, int split, void* stream) {
  deb::ProfScope prof_("layernorm", 0.0, stream, rows, dim);
  if (split && (!out_bf16 || ldo < 3LL * dim)) return fail(DE_ERR_INVALID, "de_layernorm_split3: needs an output with ldo >= 3*dim");
  if (!x || rows <= 0 || (out_bf16 && (!gamma || !beta)) || (!out_bf16 && !y))
    return fail(DE_ERR_INVALID, "de_layernorm: bad args");
  if (dim % 128 != 0 || dim > 1024 || ldx % 4 != 0 || ldo % 4 != 0 || (y && ldy % 4 != 0))
    return fail(DE_ERR_UNSUPPORTED, "de_layernorm: dim %d must be a multiple of 128, <= 1024", dim);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out_bf16);
  const int grid = (rows + 7) / 8;
  switch (dim / 128) {
#define LN_CASE(V) case V: launch_pdl(4, layernorm_kernel<V>, dim3(grid), dim3(256), 0, st, x, (long long)ldx, y, y_dtype == DE_DTYPE_BF16 ? 1 : 0, (long long)ldy, gamma, beta, rows, eps, o, (long long)ldo, split); break;
    LN_CASE(1) LN_CASE(2) LN_CASE(3) LN_CASE(4) LN_CASE(5) LN_CASE(6) LN_CASE(7) LN_CASE(8)
#undef LN_CASE
  }
  count_launch();
  return cuda_check_launch("layernorm_kernel");
}

extern "C" int de_layernorm(float* x, int64_t ldx, const void* y, int y_dtype, int64_t ldy, const float* gamma,
                            const float* beta, int rows, int dim, float eps, void* out_bf16, int64_t ldo, void* stream) {
  return layernorm_impl(x, ldx, y, y_dtype, ldy, gamma, beta, rows, dim, eps, out_bf16, ldo, 0, stream);
}
extern "C" int de_layernorm_split3(float* x, int64_t ldx, const float* gamma, const float* beta, int rows, int dim, float eps,
                                   void* out_split3, int64_t ldo, void* stream) {
  return layernorm_impl(x, ldx, nullptr, 0, 0, gamma, beta, rows, dim, eps, out_split3, ldo, 1, stream);
}

static int cast_tokens_impl(const float* x, int64_t ldx, int rows, int dim, int grp, int skip, void* out_bf16, int split,
                            void* stream) {
  deb::ProfScope prof_("cast_tokens", 0.0, stream, rows, dim);
  if (!x || !out_bf16 || rows <= 0 || dim % 8 != 0) return fail(DE_ERR_INVALID, "de_cast_tokens: bad args");
  long long out_rows = rows;
  if (grp > 0) {
    if (rows % grp != 0 || skip >= grp) return fail(DE_ERR_INVALID, "de_cast_tokens: rows %% grp != 0");
    out_rows = (long long)(rows / grp) * (grp - skip);
  }
  const long long n = out_rows * (dim / 8);
  cast_tokens_kernel<<<(unsigned)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, ldx, out_rows, dim, grp, skip, reinterpret_cast<__nv_bfloat16*>(out_bf16), split);
  count_launch();
  return cuda_check_launch("cast_tokens_kernel");
}
extern "C" int de_cast_tokens(const float* x, int64_t ldx, int rows, int dim, int grp, int skip, void* out_bf16,
                              void* stream) {
  return cast_tokens_impl(x, ldx, rows, dim, grp, skip, out_bf16, 0, stream);
}
extern "C" int de_cast_tokens_split3(const float* x, int64_t ldx, int rows, int dim, int grp, int skip, void* out_split3,
                                     void* stream) {
  return cast_tokens_impl(x, ldx, rows, dim, grp, skip, out_split3, 1, stream);
}
extern "C" int de_patchify_split3(const float* x_nchw, void* out_split3, int B, int h, int w, void* stream) {
  deb::ProfScope prof_("patchify_split3", 0.0, stream, B, h, w);
  if (!x_nchw || !out_split3 || B < 1 || h % 16 != 0 || w % 16 != 0 || h < 16 || w < 16)
    return fail(DE_ERR_INVALID, "de_patchify_split3: bad args (h, w multiples of 16)");
  const long long n = (long long)B * (h / 16) * (w / 16) * 48;
  patchify_split3_kernel<<<(unsigned)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x_nchw, reinterpret_cast<__nv_bfloat16*>(out_split3), B, h, w);
  count_launch();
  return cuda_check_launch("patchify_split3_kernel");
}

extern "C" int de_fill_cls(float* x, int64_t ldx, int B, int T, int dim, const float* cls, const float* pos,
                           void* stream) {
  deb::ProfScope prof_("fill_cls", 0.0, stream, B, T, dim);
  if (!x || !cls || !pos) return fail(DE_ERR_INVALID, "de_fill_cls: bad args");
  fill_cls_kernel<<<(B * dim + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(x, ldx, B, T, dim, cls, pos);
  count_launch();
  return cuda_check_launch("fill_cls_kernel");
}

extern "C" int de_upsample2x(const void* x, void* out, int B, int H, int W, int C, int align_corners, void* stream) {
  deb::ProfScope prof_("upsample2x", 0.0, stream, B, H, W, C);
  if (!x || !out || C % 8 != 0) return fail(DE_ERR_INVALID, "de_upsample2x: bad args (C %% 8)");
  if (B > 65535 || 2 * H > 65535) return fail(DE_ERR_UNSUPPORTED, "de_upsample2x: B / 2H exceed the grid limits");
  const bool wide = C % 16 == 0 && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 31) == 0;
  const int cg = wide ? C / 16 : C / 8;
  int cg_shift = -1;
  for (int sft = 0; sft < 12; ++sft)
    if ((1 << sft) == cg) cg_shift = sft;
  dim3 grid((unsigned)((2LL * W * cg + 255) / 256), (2 * H + kUpRows - 1) / kUpRows, B);
  if (wide)
    upsample2x_kernel<2><<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), B, H, W, C, align_corners, cg_shift);
  else
    upsample2x_kernel<1><<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), B, H, W, C, align_corners, cg_shift);
  count_launch();
  return cuda_check_launch("upsample2x_kernel");
}

extern "C" int de_upsample2x_ex(const void* x, int x_dtype, void* out, int out_mode, int B, int H, int W, int C,
                                int align_corners, void* stream) {
  deb::ProfScope prof_("upsample2x_ex", 0.0, stream, B, H, W, C);
  if (!x || !out || C % 8 != 0 || out_mode < 0 || out_mode > 2 || (x_dtype != DE_DTYPE_BF16 && x_dtype != DE_DTYPE_F32))
    return fail(DE_ERR_INVALID, "de_upsample2x_ex: bad args (C %% 8, out_mode 0..2)");
  const long long n = 4LL * B * H * W * (C / 8);
  upsample2x_ex_kernel<<<(unsigned)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, x_dtype == DE_DTYPE_F32, out, out_mode, B, H, W, C, align_corners);
  count_launch();
  return cuda_check_launch("upsample2x_ex_kernel");
}

extern "C" int de_im2col3x3(const void* x, void* out, int B, int H, int W, int C, int stride, int pad_y, int pad_x,
                            int Ho, int Wo, void* stream) {
  deb::ProfScope prof_("im2col3x3", 0.0, stream, B, H, W, C);
  if (!x || !out || C % 8 != 0 || stride < 1 || Ho < 1 || Wo < 1) return fail(DE_ERR_INVALID, "de_im2col3x3: bad args");
  const long long n = (long long)B * Ho * Wo * 9 * (C / 8);
  im2col3x3_kernel<<<(unsigned)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), B, H, W, C, stride, pad_y, pad_x,
      Ho, Wo);
  count_launch();
  return cuda_check_launch("im2col3x3_kernel");
}

extern "C" int de_subsample(const void* x, void* out, int B, int H, int W, int C, int stride, void* stream) {
  deb::ProfScope prof_("subsample", 0.0, stream, B, H, W, C);
  if (!x || !out || C % 8 != 0 || stride < 1) return fail(DE_ERR_INVALID, "de_subsample: bad args");
  const int Ho = (H + stride - 1) / stride, Wo = (W + stride - 1) / stride;
  const long long n = (long long)B * Ho * Wo * (C / 8);
  subsample_kernel<<<(unsigned)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), B, H, W, C, stride, Ho, Wo);
  count_launch();
  return cuda_check_launch("subsample_kernel");
}

// Workspace layout of de_groupnorm: [arrival counters: kGnCounterBytes | stats float2[B*groups] (16-aligned) | partials]
static const int64_t kGnCounterBytes = 16384;          // 4096 samples per call
struct GnLayout {
  int px_per_block, nblocks;
  int64_t stats_bytes, total;
};
static GnLayout gn_layout(int B, int HW, int C, int groups) {
  GnLayout L;
  L.px_per_block = (HW + 147) / 148;      // enough blocks to fill the GPU even at B=1
  const int lanes = 256 / (C / 8);
  if (L.px_per_block < lanes * 4) L.px_per_block = lanes * 4;
  L.nblocks = (HW + L.px_per_block - 1) / L.px_per_block;
  L.stats_bytes = (((int64_t)sizeof(float2) * B * groups + 15) / 16) * 16;
  L.total = kGnCounterBytes + L.stats_bytes + (int64_t)sizeof(double) * 2 * B * L.nblocks * groups;
  return L;
}
extern "C" int64_t de_groupnorm_ws_bytes(int B, int HW, int C, int groups) {
  if (B <= 0 || HW <= 0 || C < 8 || C % 8 != 0 || groups < 1 || 256 % (C / 8) != 0) return 0;
  return gn_layout(B, HW, C, groups).total;
}

extern "C" int de_groupnorm(const void* x, int x_dtype, const float* gamma, const float* beta, const void* add,
                            int add_dtype, void* out, int out_mode, float* out_f32, void* out_plain, int B, int HW, int C,
                            int groups, float eps, int act, void* ws, int64_t ws_bytes, void* stream) {
  deb::ProfScope prof_("groupnorm", 0.0, stream, B, HW, C);
  if (!x || !gamma || !beta || !ws || (!out && !out_f32 && !out_plain)) return fail(DE_ERR_INVALID, "de_groupnorm: null pointer");
  if (C % 8 != 0 || C > 2048 || groups > 64 || groups < 1 || C % groups != 0 || 256 % (C / 8) != 0)
    return fail(DE_ERR_UNSUPPORTED, "de_groupnorm: C=%d groups=%d unsupported", C, groups);
  if (B < 1 || (int64_t)B * 4 > kGnCounterBytes) return fail(DE_ERR_UNSUPPORTED, "de_groupnorm: B=%d (1..4096)", B);
  if ((x_dtype != DE_DTYPE_BF16 && x_dtype != DE_DTYPE_F32) || out_mode < 0 || out_mode > 1)
    return fail(DE_ERR_INVALID, "de_groupnorm: x_dtype %d / out_mode %d", x_dtype, out_mode);
  const int x_f32 = x_dtype == DE_DTYPE_F32;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  GnLayout L = gn_layout(B, HW, C, groups);
  if (ws_bytes < L.total) return fail(DE_ERR_INVALID, "de_groupnorm: workspace %lld < %lld bytes (de_groupnorm_ws_bytes)", (long long)ws_bytes, (long long)L.total);
  const int px_per_block = L.px_per_block, nblocks = L.nblocks;
  // arrival counters at a FIXED offset (the first kGnCounterBytes of the workspace; zero before the first call, zero
  // after every call), so calls with different B / HW can share a workspace without a counter ever aliasing a partial
  unsigned* counters = reinterpret_cast<unsigned*>(ws);
  float2* stats = reinterpret_cast<float2*>(reinterpret_cast<char*>(ws) + kGnCounterBytes);
  double* partial = reinterpret_cast<double*>(reinterpret_cast<char*>(ws) + kGnCounterBytes + L.stats_bytes);
  dim3 grid(nblocks, B);
  {
    // one-launch path when the whole grid is co-resident (batch-1 latency path); DEB_GN_FUSED=0 disables it
    static int fused_on = -1;
    if (fused_on < 0) {
      const char* e = getenv("DEB_GN_FUSED");
      fused_on = (e && e[0] == '0') ? 0 : 1;
    }
    const size_t tile_bytes = (size_t)px_per_block * C * sizeof(float);
    if (fused_on && B <= 2048 && tile_bytes <= 160 * 1024) {
      static SmemOptIn optin;
      int occ = 0;
      if (ensure_smem(optin, groupnorm_fused_kernel, (int)tile_bytes) == cudaSuccess &&
          cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, groupnorm_fused_kernel, 256, tile_bytes) == cudaSuccess &&
          (long long)occ * num_sms() >= (long long)nblocks * B) {
        groupnorm_fused_kernel<<<grid, 256, tile_bytes, st>>>(
            x, x_f32, HW, C, groups, px_per_block, partial, counters, (double)HW * (C / groups), eps, gamma, beta, add,
            add_dtype, reinterpret_cast<__nv_bfloat16*>(out), out_mode, out_f32, reinterpret_cast<__nv_bfloat16*>(out_plain), act);
        count_launch();
        return cuda_check_launch("groupnorm_fused_kernel");
      }
    }
  }
  groupnorm_partial_kernel<<<grid, 256, 0, st>>>(x, x_f32, HW, C, groups, px_per_block, partial, counters, stats,
                                                 (double)HW * (C / groups), eps);
  count_launch();
  const long long n = (long long)B * HW * (C / 8);
  groupnorm_apply_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(
      x, x_f32, stats, gamma, beta, add, add_dtype, reinterpret_cast<__nv_bfloat16*>(out), out_mode, out_f32,
      reinterpret_cast<__nv_bfloat16*>(out_plain), B, HW, C, groups, act);
  count_launch();
  return cuda_check_launch("groupnorm kernels");
}

// ---------------------------------------------------------------------------------------------- de_groupnorm_static
// GroupNorm for callers that may not allocate (CUDA-graph capture, tight serving loops, harnesses that forbid
// cudaMalloc): the workspace of de_groupnorm lives in a file-scope device array instead of a caller buffer.  Every GPU
// that loads the library holds these kGnStaticBytes until exit.  In host code `g_gn_static_ws` is only a stand-in for the
// symbol; the device address comes from cudaGetSymbolAddress, looked up once per device with that device current.
static const int64_t kGnStaticBytes = 4 << 20;          // DE_GROUPNORM_STATIC_MAX_BYTES in include/depth_b200.h
static __device__ __align__(256) unsigned char g_gn_static_ws[kGnStaticBytes];

static const int kGnStaticMaxDevices = 64;
static void* g_gn_static_ptr[kGnStaticMaxDevices];      // per device ordinal; written once under the mutex
static std::mutex g_gn_static_mu;

static int gn_static_ws(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(DE_ERR_CUDA, "de_groupnorm_static: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kGnStaticMaxDevices) return fail(DE_ERR_UNSUPPORTED, "de_groupnorm_static: device ordinal %d (max %d)", dev, kGnStaticMaxDevices - 1);
  std::lock_guard<std::mutex> lock(g_gn_static_mu);
  if (!g_gn_static_ptr[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_gn_static_ws);     // the CURRENT device's copy of the array
    if (e != cudaSuccess || !ptr)
      return fail(DE_ERR_CUDA, "de_groupnorm_static: cudaGetSymbolAddress(g_gn_static_ws) on device %d: %s", dev, cudaGetErrorString(e));
    g_gn_static_ptr[dev] = ptr;
  }
  *out = g_gn_static_ptr[dev];
  return DE_OK;
}

extern "C" int64_t de_groupnorm_static_max_bytes(void) { return kGnStaticBytes; }

extern "C" int de_groupnorm_static(const void* x, int x_dtype, const float* gamma, const float* beta, const void* add,
                                   int add_dtype, void* out, int out_mode, float* out_f32, void* out_plain, int B, int HW,
                                   int C, int groups, float eps, int act, void* stream) {
  const int64_t need = de_groupnorm_ws_bytes(B, HW, C, groups);
  if (need <= 0) return fail(DE_ERR_UNSUPPORTED, "de_groupnorm_static: unsupported shape B=%d HW=%d C=%d groups=%d", B, HW, C, groups);
  if (need > kGnStaticBytes)
    return fail(DE_ERR_UNSUPPORTED,
                "de_groupnorm_static: B=%d HW=%d C=%d groups=%d needs %lld bytes of scratch, the built-in array holds %lld "
                "(DE_GROUPNORM_STATIC_MAX_BYTES); use de_groupnorm with a caller workspace",
                B, HW, C, groups, (long long)need, (long long)kGnStaticBytes);
  void* ws = nullptr;
  int rc = gn_static_ws(&ws);
  if (rc != DE_OK) return rc;
  // The array keeps whatever the last call (of any shape) left.  Only the arrival counters must start at zero; stats and
  // partial sums are fully written before they are read.  A completed call leaves the counters zero already - the
  // memset (a memset node under capture, no allocation) also covers a predecessor that was aborted mid-kernel.
  cudaError_t e = cudaMemsetAsync(ws, 0, kGnCounterBytes, reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail(DE_ERR_CUDA, "de_groupnorm_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
  return de_groupnorm(x, x_dtype, gamma, beta, add, add_dtype, out, out_mode, out_f32, out_plain, B, HW, C, groups, eps,
                      act, ws, kGnStaticBytes, stream);
}
