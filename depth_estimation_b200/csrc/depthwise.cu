// Depthwise k3 / k5 convolution, stride 1 / 2, TF-"same" padding, folded BatchNorm bias + ReLU6 - the middle of
// EfficientNet-Lite3's MBConv blocks ([UPSTREAM] geffnet InvertedResidual.conv_dw / bn2 / act, reached through the hub
// model of /root/reference/depth_midas.py:10,26 for model_type "MiDaS_small").  NHWC bf16 in, NHWC bf16 out.
//
// A bandwidth kernel staged through shared memory (north_star (2)): persistent CTAs walk (channel chunk, image, tile)
// work items; the input window of a tile - 8x16 outputs x 64 channels plus the k-1 halo - arrives as ONE 4-D TMA box
// (the tensor map's out-of-bounds zero fill IS the TF-same padding, negative start coordinates included), double
// buffered on two mbarriers so the next window streams in while the current one is convolved.  Low-resolution images
// (4x8 / 8x8 / 4x16 pixels per frame) use tile shapes that pack four / two frames into one work item, so the last
// blocks of the network do not convolve 75 % padding.  Each thread owns 8
// channels (one 16-byte shared-memory vector per pixel; a quarter-warp reads 128 contiguous bytes - conflict free) and a
// strip of 4 adjacent outputs, so a window row is read and converted once per ky instead of once per tap; the filter
// taps live in shared memory as fp32 pairs and every MAC is a packed FFMA2 (two channels per issue slot).  The first
// version of this operator (one thread per output, k*k global 16-byte loads of x and w each) spent ~80 instructions
// per output element; this one ~10 (k3) / ~25 (k5).
// Accumulation: fp32, bias first, taps in (ky, kx) order - the same order as the direct kernel it replaces.
#include <stdlib.h>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

constexpr int kDwCC = 64;                     // channels per work item (128 bytes per pixel)
constexpr int kDwPix = 128;                   // outputs per work item: TH x TW pixels of NB = 128 / (TH * TW) frames
constexpr int kDwP = 4;                       // outputs per thread along x
constexpr int kDwThreads = 256;               // 8 channel groups x 32 strips

// Tile shapes: 8 x 16 pixels of one frame, or - for the low-resolution blocks, where an 8 x 16 tile would be up to 75 %
// padding (4 x 8 pixels per frame at the end of the 128 x 256 network) - 8 x 8 / 4 x 16 pixels of two frames or 4 x 8
// pixels of four frames: the TMA box then spans NB frames, each with its own zero-filled halo.
template <int K, int S, int TH, int TW>
struct DwCfg {
  static constexpr int NB = kDwPix / (TH * TW);           // frames per work item
  static constexpr int SPR = TW / kDwP;                   // strips per tile row
  static constexpr int WH = (TH - 1) * S + K;             // staged window of one frame, rows
  static constexpr int WW = (TW - 1) * S + K;             // staged window, columns
  static constexpr int NC = (kDwP - 1) * S + K;           // window columns one strip touches
  static constexpr int kStageBytes = NB * WH * WW * kDwCC * 2;
  static constexpr int kWBytes = K * K * kDwCC * 4;       // fp32 taps of one channel chunk
  static constexpr int kSmem = 2 * kStageBytes + kWBytes + kDwCC * 4 + 128;
};

struct DwParams {
  const __nv_bfloat16* w;   // [K*K][C] tap-major
  const float* bias;        // [C]
  __nv_bfloat16* out;       // [B,Ho,Wo,C]
  int B, C, Ho, Wo, pad_y, pad_x, act;
  int tiles_x, tiles_y, bgroups, chunks, ntiles;
};

template <int K, int S, int TH, int TW>
__global__ void __launch_bounds__(kDwThreads, S == 1 ? 3 : 1) depthwise_tma_kernel(const __grid_constant__ CUtensorMap tmX, const DwParams p) {
  using Cfg = DwCfg<K, S, TH, TW>;
  extern __shared__ uint8_t dw_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dw_smem_raw) + 127) & ~static_cast<uintptr_t>(127));
  auto stage_ptr = [&](int s) { return smem + s * Cfg::kStageBytes; };
  float4* s_w = reinterpret_cast<float4*>(smem + 2 * Cfg::kStageBytes);       // [tap][half][cg] float4
  float4* s_b = s_w + K * K * 16;                                              // [half][cg]
  __shared__ __align__(8) uint64_t full[2];

  const int tid = threadIdx.x;
  const int cg = tid & 7, strip = tid >> 3;
  const int sx = (strip % Cfg::SPR) * kDwP, sy = (strip / Cfg::SPR) % TH, nb = strip / (Cfg::SPR * TH);
  if (tid == 0) {
    tma_prefetch_desc(&tmX);
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_fence_init();
  }
  __syncthreads();
  const int per_chunk = p.bgroups * p.tiles_y * p.tiles_x;

  auto issue = [&](int t, int s) {            // one thread: window of work item t -> stage s
    const int chunk = t / per_chunk;
    int r = t - chunk * per_chunk;
    const int bg = r / (p.tiles_y * p.tiles_x);
    r -= bg * (p.tiles_y * p.tiles_x);
    const int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
    mbar_arrive_expect_tx(&full[s], Cfg::kStageBytes);   // frames past B are zero-filled and still count
    tma_load_4d(stage_ptr(s), &tmX, &full[s], chunk * kDwCC, tx * TW * S - p.pad_x, ty * TH * S - p.pad_y, bg * Cfg::NB);
  };

  pdl_wait();                                 // x is the previous kernel's output
  // blocked partition: a CTA walks a contiguous range of work items - neighbouring tiles of one (chunk, image), so the
  // tap table is reloaded only when the range crosses a channel chunk and halo rows are re-read from L2 while hot
  int t = static_cast<int>((long long)blockIdx.x * p.ntiles / gridDim.x);
  const int t_end = static_cast<int>((long long)(blockIdx.x + 1) * p.ntiles / gridDim.x);
  if (tid == 0 && t < t_end) issue(t, 0);
  int cur_chunk = -1;
  for (int n = 0; t < t_end; ++t, ++n) {
    const int s = n & 1;
    if (tid == 0 && t + 1 < t_end) issue(t + 1, s ^ 1);   // stage s^1 was drained at the end of n-1
    const int chunk = t / per_chunk;
    int r = t - chunk * per_chunk;
    const int bg = r / (p.tiles_y * p.tiles_x);
    r -= bg * (p.tiles_y * p.tiles_x);
    const int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
    const int b = bg * Cfg::NB + nb;
    const int c0 = chunk * kDwCC;
    if (chunk != cur_chunk) {                 // taps + bias of this channel chunk -> fp32 in shared memory
      cur_chunk = chunk;
      for (int i = tid; i < K * K * 8; i += kDwThreads) {
        const int tap = i >> 3, g = i & 7;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (c0 + g * 8 < p.C) v = __ldg(reinterpret_cast<const uint4*>(p.w + (long long)tap * p.C + c0 + g * 8));
        s_w[tap * 16 + g] = make_float4(bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y));
        s_w[tap * 16 + 8 + g] = make_float4(bf16_lo(v.z), bf16_hi(v.z), bf16_lo(v.w), bf16_hi(v.w));
      }
      if (tid < 16) {
        const int g = tid & 7, half = tid >> 3;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c0 + g * 8 < p.C) v = __ldg(reinterpret_cast<const float4*>(p.bias + c0 + g * 8 + half * 4));
        s_b[half * 8 + g] = v;
      }
      __syncthreads();
    }
    mbar_wait(&full[s], (n >> 1) & 1);

    float2 acc[kDwP][4];
    {
      const float4 b0 = s_b[cg], b1 = s_b[8 + cg];
#pragma unroll
      for (int q = 0; q < kDwP; ++q) {
        acc[q][0] = make_float2(b0.x, b0.y); acc[q][1] = make_float2(b0.z, b0.w);
        acc[q][2] = make_float2(b1.x, b1.y); acc[q][3] = make_float2(b1.z, b1.w);
      }
    }
    const uint4* win = reinterpret_cast<const uint4*>(stage_ptr(s));
#pragma unroll
    for (int ky = 0; ky < K; ++ky) {
      const uint4* rowp = win + (((nb * Cfg::WH) + sy * S + ky) * Cfg::WW + sx * S) * 8 + cg;
      float2 in[Cfg::NC][4];
#pragma unroll
      for (int c = 0; c < Cfg::NC; ++c) {
        const uint4 v = rowp[c * 8];
        in[c][0] = make_float2(bf16_lo(v.x), bf16_hi(v.x)); in[c][1] = make_float2(bf16_lo(v.y), bf16_hi(v.y));
        in[c][2] = make_float2(bf16_lo(v.z), bf16_hi(v.z)); in[c][3] = make_float2(bf16_lo(v.w), bf16_hi(v.w));
      }
#pragma unroll
      for (int kx = 0; kx < K; ++kx) {
        const float4 w0 = s_w[(ky * K + kx) * 16 + cg], w1 = s_w[(ky * K + kx) * 16 + 8 + cg];
        const float2 wv[4] = {make_float2(w0.x, w0.y), make_float2(w0.z, w0.w), make_float2(w1.x, w1.y), make_float2(w1.z, w1.w)};
#pragma unroll
        for (int q = 0; q < kDwP; ++q)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[q][j] = ffma2(in[q * S + kx][j], wv[j], acc[q][j]);
      }
    }
    const int oy = ty * TH + sy, ox0 = tx * TW + sx, c = c0 + cg * 8;
    if (oy < p.Ho && c < p.C && b < p.B) {
      __nv_bfloat16* op = p.out + (((long long)b * p.Ho + oy) * p.Wo + ox0) * p.C + c;
#pragma unroll
      for (int q = 0; q < kDwP; ++q) {
        if (ox0 + q >= p.Wo) break;
        float v[8] = {acc[q][0].x, acc[q][0].y, acc[q][1].x, acc[q][1].y, acc[q][2].x, acc[q][2].y, acc[q][3].x, acc[q][3].y};
        if (p.act == DE_ACT_RELU6) {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = fminf(fmaxf(v[j], 0.f), 6.f);
        } else if (p.act == DE_ACT_RELU) {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        *reinterpret_cast<uint4*>(op + (long long)q * p.C) =
            make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
      }
    }
    __syncthreads();                          // stage s and the tap table are free again
  }
}

static int dw_same_pad_before(int in, int k, int s) {
  const int out = (in + s - 1) / s;
  int total = (out - 1) * s + k - in;
  if (total < 0) total = 0;
  return total / 2;
}

template <int K, int S, int TH, int TW>
static int launch_depthwise(const void* x, const DwParams& p0, int H, int W, cudaStream_t st) {
  using Cfg = DwCfg<K, S, TH, TW>;
  DwParams p = p0;
  p.tiles_x = (p.Wo + TW - 1) / TW;
  p.tiles_y = (p.Ho + TH - 1) / TH;
  p.bgroups = (p.B + Cfg::NB - 1) / Cfg::NB;
  const long long nt = (long long)p.chunks * p.bgroups * p.tiles_y * p.tiles_x;
  if (nt > 0x7fffffffLL) return fail(DE_ERR_UNSUPPORTED, "de_depthwise: too many tiles");
  p.ntiles = (int)nt;
  auto fn = get_tensor_map_encoder();
  if (!fn) return DE_ERR_CUDA;
  CUtensorMap tm;
  const uint64_t dims[4] = {(uint64_t)p.C, (uint64_t)W, (uint64_t)H, (uint64_t)p.B};
  const uint64_t strides[3] = {(uint64_t)p.C * 2, (uint64_t)W * p.C * 2, (uint64_t)H * W * p.C * 2};
  const uint32_t box[4] = {kDwCC, (uint32_t)Cfg::WW, (uint32_t)Cfg::WH, (uint32_t)Cfg::NB};
  const uint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(DE_ERR_CUDA, "de_depthwise: cuTensorMapEncodeTiled failed (%d)", (int)r);
  static SmemOptIn optin;
  cudaError_t e = ensure_smem(optin, depthwise_tma_kernel<K, S, TH, TW>, Cfg::kSmem);
  if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(depthwise)");
  static int occ_cache[32] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  int occ = (dev >= 0 && dev < 32) ? occ_cache[dev] : 0;
  if (occ == 0) {
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, depthwise_tma_kernel<K, S, TH, TW>, kDwThreads, Cfg::kSmem);
    if (e != cudaSuccess || occ < 1) occ = 1;
    if (dev >= 0 && dev < 32) occ_cache[dev] = occ;
  }
  int grid = num_sms() * occ;
  if (p.ntiles < grid) grid = p.ntiles;
  e = launch_pdl(8, depthwise_tma_kernel<K, S, TH, TW>, dim3(grid), dim3(kDwThreads), (size_t)Cfg::kSmem, st, tm, p);
  count_launch();
  if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(depthwise_tma_kernel)");
  return cuda_check_launch("depthwise_tma_kernel");
}

}  // namespace deb

using namespace deb;

extern "C" int de_depthwise(const void* x, const void* w_tapmajor, const float* bias, void* out, int B, int H, int W,
                            int C, int k, int stride, int act, void* stream) {
  deb::ProfScope prof_("depthwise", 2.0 * B * ((H + stride - 1) / stride) * ((W + stride - 1) / stride) * (double)C * k * k, stream, (int64_t)B * H * W, C, k, stride);
  if (!x || !w_tapmajor || !bias || !out || C % 8 != 0 || B < 1 || H < 1 || W < 1)
    return fail(DE_ERR_INVALID, "de_depthwise: bad args (C %% 8)");
  if ((k != 3 && k != 5) || (stride != 1 && stride != 2)) return fail(DE_ERR_UNSUPPORTED, "de_depthwise: k=%d stride=%d (k 3|5, stride 1|2)", k, stride);
  if ((reinterpret_cast<uintptr_t>(x) & 15) != 0) return fail(DE_ERR_INVALID, "de_depthwise: x must be 16-byte aligned");
  DwParams p{};
  p.w = reinterpret_cast<const __nv_bfloat16*>(w_tapmajor);
  p.bias = bias;
  p.out = reinterpret_cast<__nv_bfloat16*>(out);
  p.B = B;
  p.C = C;
  p.Ho = (H + stride - 1) / stride;
  p.Wo = (W + stride - 1) / stride;
  p.pad_y = dw_same_pad_before(H, k, stride);
  p.pad_x = dw_same_pad_before(W, k, stride);
  p.act = act;
  p.chunks = (C + kDwCC - 1) / kDwCC;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (stride == 2) {                                          // stride-2 windows are too large to pack frames
    if (k == 3) return launch_depthwise<3, 2, 8, 16>(x, p, H, W, st);
    return launch_depthwise<5, 2, 8, 16>(x, p, H, W, st);
  }
  // stride 1: the tile shape with the least padding; several frames per work item only when the batch has them
  static const int force = getenv("DEB_DW_TILE") ? atoi(getenv("DEB_DW_TILE")) : 0;   // developer A/B: 816 | 808 | 416 | 408
  int th = 8, tw = 16;
  if (p.Wo <= 8) tw = 8;
  if ((p.Ho + 3) / 4 * 4 < (p.Ho + 7) / 8 * 8) th = 4;
  if (kDwPix / (th * tw) > B) {                               // not enough frames to fill the packed shape
    if (B >= 2 && th * tw == 32) { if (p.Wo > 8) tw = 16; else th = 8; }
    else { th = 8; tw = 16; }
  }
  if (force) { th = force / 100; tw = force % 100; }
#define DW_LAUNCH(KK)                                                                         \
  do {                                                                                        \
    if (th == 8 && tw == 16) return launch_depthwise<KK, 1, 8, 16>(x, p, H, W, st);           \
    if (th == 8 && tw == 8) return launch_depthwise<KK, 1, 8, 8>(x, p, H, W, st);             \
    if (th == 4 && tw == 16) return launch_depthwise<KK, 1, 4, 16>(x, p, H, W, st);           \
    if (th == 4 && tw == 8) return launch_depthwise<KK, 1, 4, 8>(x, p, H, W, st);             \
  } while (0)
  if (k == 3) DW_LAUNCH(3);
  else DW_LAUNCH(5);
#undef DW_LAUNCH
  return fail(DE_ERR_INVALID, "de_depthwise: DEB_DW_TILE=%d", force);
}
