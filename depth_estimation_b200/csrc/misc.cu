// Bandwidth-bound helper kernels between the tensor-core GEMMs: LayerNorm, token casts, cls fill,
// bilinear x2 upsample (NHWC), strided 3x3 im2col.  All are 16-byte vectorised and grid-stride free
// (one element group per thread, grids far larger than 148 SMs at the BASELINE shapes).
#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

// ---------------------------------------------------------------- LayerNorm (one warp per row)
template <int VEC>  // VEC float4 per lane: dim = 128*VEC
__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ x, long long ldx,
                                                        const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, int rows, float eps,
                                                        __nv_bfloat16* __restrict__ out, long long ldo) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float4* xr = reinterpret_cast<const float4*>(x + (long long)row * ldx);
  float4 v[VEC];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    v[i] = xr[lane + i * 32];
    sum += v[i].x + v[i].y + v[i].z + v[i].w;
  }
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
    uint2 o;
    o.x = pack_bf16((v[i].x - mean) * rstd * g.x + bb.x, (v[i].y - mean) * rstd * g.y + bb.y);
    o.y = pack_bf16((v[i].z - mean) * rstd * g.z + bb.z, (v[i].w - mean) * rstd * g.w + bb.w);
    orow[lane + i * 32] = o;
  }
}

// ---------------------------------------------------------------- fp32 -> bf16 token cast (drop cls)
__global__ void cast_tokens_kernel(const float* __restrict__ x, long long ldx, long long out_rows, int dim, int grp,
                                   int skip, __nv_bfloat16* __restrict__ out) {
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
  *reinterpret_cast<uint4*>(out + orow * dim + c) = o;
}

__global__ void fill_cls_kernel(float* __restrict__ x, long long ldx, int B, int T, int dim,
                                const float* __restrict__ cls, const float* __restrict__ pos) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * dim) return;
  const int b = i / dim, c = i % dim;
  x[(long long)b * T * ldx + c] = cls[c] + pos[c];
}

// ---------------------------------------------------------------- bilinear x2, NHWC bf16
__global__ void upsample2x_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int B, int H,
                                  int W, int C, int align_corners) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one 8-channel group per thread
  const int cg = C >> 3;
  const int OH = 2 * H, OW = 2 * W;
  if (i >= (long long)B * OH * OW * cg) return;
  const int c = static_cast<int>(i % cg) * 8;
  long long t = i / cg;
  const int ox = static_cast<int>(t % OW);
  t /= OW;
  const int oy = static_cast<int>(t % OH);
  const int b = static_cast<int>(t / OH);
  float sy, sx;
  if (align_corners) {
    // ATen area_pixel_compute_scale: (in-1)/(out-1) in float; src = scale*dst
    const float rh = OH > 1 ? (float)(H - 1) / (float)(OH - 1) : 0.f;
    const float rw = OW > 1 ? (float)(W - 1) / (float)(OW - 1) : 0.f;
    sy = rh * oy;
    sx = rw * ox;
  } else {
    sy = fmaxf(0.5f * (oy + 0.5f) - 0.5f, 0.f);
    sx = fmaxf(0.5f * (ox + 0.5f) - 0.5f, 0.f);
  }
  const int y0 = min(static_cast<int>(sy), H - 1), x0 = min(static_cast<int>(sx), W - 1);
  const int y1 = min(y0 + 1, H - 1), x1 = min(x0 + 1, W - 1);
  const float ly = sy - y0, lx = sx - x0;
  const float hy = 1.f - ly, hx = 1.f - lx;
  const __nv_bfloat16* base = x + (long long)b * H * W * C + c;
  const uint4 v00 = *reinterpret_cast<const uint4*>(base + ((long long)y0 * W + x0) * C);
  const uint4 v01 = *reinterpret_cast<const uint4*>(base + ((long long)y0 * W + x1) * C);
  const uint4 v10 = *reinterpret_cast<const uint4*>(base + ((long long)y1 * W + x0) * C);
  const uint4 v11 = *reinterpret_cast<const uint4*>(base + ((long long)y1 * W + x1) * C);
  const uint32_t* a = &v00.x;
  const uint32_t* bq = &v01.x;
  const uint32_t* cq = &v10.x;
  const uint32_t* d = &v11.x;
  uint32_t o[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float lo = hy * (hx * bf16_lo(a[k]) + lx * bf16_lo(bq[k])) + ly * (hx * bf16_lo(cq[k]) + lx * bf16_lo(d[k]));
    const float hi = hy * (hx * bf16_hi(a[k]) + lx * bf16_hi(bq[k])) + ly * (hx * bf16_hi(cq[k]) + lx * bf16_hi(d[k]));
    o[k] = pack_bf16(lo, hi);
  }
  *reinterpret_cast<uint4*>(out + (((long long)b * OH + oy) * OW + ox) * C + c) = make_uint4(o[0], o[1], o[2], o[3]);
}

// ---------------------------------------------------------------- im2col 3x3, stride s, pad 1
__global__ void im2col3x3_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int B, int H,
                                 int W, int C, int stride, int Ho, int Wo) {
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
  const int iy = oy * stride + tap / 3 - 1, ix = ox * stride + tap % 3 - 1;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (iy >= 0 && iy < H && ix >= 0 && ix < W)
    v = *reinterpret_cast<const uint4*>(x + (((long long)b * H + iy) * W + ix) * C + c);
  *reinterpret_cast<uint4*>(out + (((long long)b * Ho + oy) * Wo + ox) * 9 * C + (long long)tap * C + c) = v;
}

}  // namespace deb

using namespace deb;

extern "C" int de_layernorm(const float* x, int64_t ldx, const float* gamma, const float* beta, int rows, int dim,
                            float eps, void* out_bf16, int64_t ldo, void* stream) {
  if (!x || !gamma || !beta || !out_bf16 || rows <= 0) return fail(DE_ERR_INVALID, "de_layernorm: bad args");
  if (dim % 128 != 0 || dim > 1024 || ldx % 4 != 0 || ldo % 4 != 0)
    return fail(DE_ERR_UNSUPPORTED, "de_layernorm: dim %d must be a multiple of 128, <= 1024", dim);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out_bf16);
  const int grid = (rows + 7) / 8;
  switch (dim / 128) {
#define LN_CASE(V) case V: layernorm_kernel<V><<<grid, 256, 0, st>>>(x, ldx, gamma, beta, rows, eps, o, ldo); break;
    LN_CASE(1) LN_CASE(2) LN_CASE(3) LN_CASE(4) LN_CASE(5) LN_CASE(6) LN_CASE(7) LN_CASE(8)
#undef LN_CASE
  }
  count_launch();
  return cuda_check_launch("layernorm_kernel");
}

extern "C" int de_cast_tokens(const float* x, int64_t ldx, int rows, int dim, int grp, int skip, void* out_bf16,
                              void* stream) {
  if (!x || !out_bf16 || rows <= 0 || dim % 8 != 0) return fail(DE_ERR_INVALID, "de_cast_tokens: bad args");
  long long out_rows = rows;
  if (grp > 0) {
    if (rows % grp != 0 || skip >= grp) return fail(DE_ERR_INVALID, "de_cast_tokens: rows %% grp != 0");
    out_rows = (long long)(rows / grp) * (grp - skip);
  }
  const long long n = out_rows * (dim / 8);
  cast_tokens_kernel<<<(unsigned)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, ldx, out_rows, dim, grp, skip, reinterpret_cast<__nv_bfloat16*>(out_bf16));
  count_launch();
  return cuda_check_launch("cast_tokens_kernel");
}

extern "C" int de_fill_cls(float* x, int64_t ldx, int B, int T, int dim, const float* cls, const float* pos,
                           void* stream) {
  if (!x || !cls || !pos) return fail(DE_ERR_INVALID, "de_fill_cls: bad args");
  fill_cls_kernel<<<(B * dim + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(x, ldx, B, T, dim, cls, pos);
  count_launch();
  return cuda_check_launch("fill_cls_kernel");
}

extern "C" int de_upsample2x(const void* x, void* out, int B, int H, int W, int C, int align_corners, void* stream) {
  if (!x || !out || C % 8 != 0) return fail(DE_ERR_INVALID, "de_upsample2x: bad args (C %% 8)");
  const long long n = (long long)B * 4 * H * W * (C / 8);
  upsample2x_kernel<<<(unsigned)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), B, H, W, C, align_corners);
  count_launch();
  return cuda_check_launch("upsample2x_kernel");
}

extern "C" int de_im2col3x3(const void* x, void* out, int B, int H, int W, int C, int stride, void* stream) {
  if (!x || !out || C % 8 != 0 || stride < 1) return fail(DE_ERR_INVALID, "de_im2col3x3: bad args");
  const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
  const long long n = (long long)B * Ho * Wo * 9 * (C / 8);
  im2col3x3_kernel<<<(unsigned)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), B, H, W, C, stride, Ho, Wo);
  count_launch();
  return cuda_check_launch("im2col3x3_kernel");
}
