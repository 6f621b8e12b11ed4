// Fused depthwise k3 / k5 + pointwise-linear 1x1 of EfficientNet-Lite3's MBConv blocks ([UPSTREAM] geffnet
// InvertedResidual: conv_dw -> bn2 -> act -> conv_pwl -> bn3 [-> + x], reached through the hub model of
// /root/reference/depth_midas.py:10,26 for model_type "MiDaS_small").  Replaces the pair de_depthwise -> de_gemm:
// the depthwise output - the 6x-expanded activation - never exists in global memory.  One CTA owns an 8x16-pixel
// output tile (M = 128 rows of the projection GEMM) and walks 64-channel chunks of the expanded tensor:
//
//   control warp : TMA - the chunk's input window (8x16 outputs + halo, out-of-bounds zero fill = TF-"same" padding)
//                  and the chunk's [cout x 64] slice of the projection weights; tcgen05.mma issue (M 128, N = cout,
//                  K = 64 per chunk), accumulating every chunk into one TMEM accumulator
//   8 compute warps : depthwise taps on the staged window (the arithmetic of csrc/depthwise.cu: fp32, bias first,
//                  taps in (ky, kx) order, packed FFMA2), ReLU6, bf16, written as the 128-byte-swizzled K-major A
//                  operand of the projection; then the epilogue: TMEM -> registers, + folded-BN bias [+ residual], bf16
//
// Low-resolution blocks have few tiles (32 at 8x16 pixels per frame): the chunks of a tile are split over `groups`
// CTAs, each storing its fp32 partial tile into its own slice of a workspace; the last CTA of a tile to arrive
// (counter in the workspace) adds the slices in a fixed order and runs the epilogue - deterministic, no extra launch.
#include <cuda.h>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

constexpr int kMbCC = 64;                     // expanded channels per chunk (one 128-byte K block)
constexpr int kMbTH = 8, kMbTW = 16;          // output tile = 128 GEMM rows
constexpr int kMbP = 4;                       // outputs per thread along x
constexpr int kMbCompute = 256;               // 8 channel groups x 32 strips
constexpr int kMbThreads = kMbCompute + 32;   // + control warp
#ifndef MB_WIN_BUFS
#define MB_WIN_BUFS 2                         // input-window buffers per CTA (1 = ~105 KB, two CTAs per SM: measured slower, 1.78 vs 1.70 ms)
#endif

template <int K, int S>
struct MbCfg {
  static constexpr int WH = (kMbTH - 1) * S + K;
  static constexpr int WW = (kMbTW - 1) * S + K;
  static constexpr int NC = (kMbP - 1) * S + K;
  static constexpr int kWinBytes = (WH * WW * kMbCC * 2 + 1023) / 1024 * 1024;
  static constexpr int kTapBytes = K * K * kMbCC * 4 + kMbCC * 4;     // fp32 taps + bias of one chunk
  static constexpr int kA2Bytes = 128 * kMbCC * 2;                    // 16 KB, 1024-aligned
};

struct MbParams {
  const __nv_bfloat16* dw_w;   // [K*K][C] tap-major
  const float* dw_b;           // [C]
  const float* pw_b;           // [cout]
  const __nv_bfloat16* res;    // [B,Ho,Wo,cout] or null
  __nv_bfloat16* out;          // [B,Ho,Wo,cout]
  float* ws;                   // [tiles][groups][128][cout_pad] fp32 partial tiles (groups > 1)
  unsigned* counters;          // [tiles], zero before the first call, left zero
  int B, C, Ho, Wo, pad_y, pad_x, cout, cout_pad, n1;   // n1: columns of the first MMA part (cout_pad or 192)
  int tiles_x, tiles_y, chunks, groups, cpg, ntiles;
  int wp_bytes, tmem_cols;
};

// smem: [win 0 | win 1 | A2 0 | A2 1 | Wp 0 | Wp 1 | taps | barriers]
template <int K, int S>
__global__ void __launch_bounds__(kMbThreads, MB_WIN_BUFS == 1 ? 2 : 1) dwproject_kernel(const __grid_constant__ CUtensorMap tmX,
                                                                  const __grid_constant__ CUtensorMap tmW, const MbParams p) {
  using Cfg = MbCfg<K, S>;
  extern __shared__ uint8_t mb_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(mb_smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* s_win = smem;
  uint8_t* s_a2 = s_win + MB_WIN_BUFS * Cfg::kWinBytes;
  uint8_t* s_wp = s_a2 + 2 * Cfg::kA2Bytes;
  float4* s_w = reinterpret_cast<float4*>(s_wp + 2 * p.wp_bytes);          // [tap][half][cg] float4, then bias
  float4* s_b = s_w + K * K * 16;
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(s_w) + ((Cfg::kTapBytes + 15) / 16) * 16);
  uint64_t* win_full = bars;          // [2]
  uint64_t* win_empty = bars + 2;     // [2]
  uint64_t* a2_full = bars + 4;       // [2]
  uint64_t* a2_empty = bars + 6;      // [2]
  uint64_t* wp_full = bars + 8;       // [2]
  uint64_t* wp_empty = bars + 10;     // [2]
  uint64_t* acc_full = bars + 12;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);
  __shared__ int s_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = blockIdx.x / p.groups, grp = blockIdx.x - tile * p.groups;
  const int per_img = p.tiles_y * p.tiles_x;
  const int b = tile / per_img, t_in = tile - b * per_img;
  const int ty = t_in / p.tiles_x, tx = t_in - ty * p.tiles_x;
  const int c_begin = grp * p.cpg, c_end = min(c_begin + p.cpg, p.chunks), nch = c_end - c_begin;

  if (warp == 8) {
    if (lane == 0) {
      tma_prefetch_desc(&tmX);
      tma_prefetch_desc(&tmW);
      for (int i = 0; i < 13; ++i) mbar_init(&bars[i], 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, p.tmem_cols);
    tmem_relinquish();
  }
  pdl_launch();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;
  pdl_wait();                                   // x (the expand GEMM's output) and the residual are now readable

  const int win_tx = Cfg::WH * Cfg::WW * kMbCC * 2;
  if (warp == 8) {
    // ------------------------------------------------------------------ control: TMA + MMA issue
    const uint32_t idesc1 = make_idesc_bf16(128, p.n1);
    const int n2 = p.cout_pad - p.n1;
    const uint32_t idesc2 = make_idesc_bf16(128, n2 > 0 ? n2 : 16);
    auto issue_mma = [&](int j) {               // projection of local chunk j: acc (+)= A2[j & 1] x Wp[j & 1]^T
      const int s = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      mbar_wait(&a2_full[s], ph);
      mbar_wait(&wp_full[s], ph);
      tc_fence_after();
      const uint64_t adesc = make_sw128_kmajor_desc(smem_u32(s_a2 + s * Cfg::kA2Bytes));
      const uint64_t bdesc = make_sw128_kmajor_desc(smem_u32(s_wp + s * p.wp_bytes));
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < kMbCC / 16; ++k) {
          umma_bf16(tmem_acc, adesc + 2 * k, bdesc + 2 * k, idesc1, (j | k) != 0);
          if (n2 > 0) umma_bf16(tmem_acc + p.n1, adesc + 2 * k, bdesc + ((p.n1 * 128) >> 4) + 2 * k, idesc2, (j | k) != 0);
        }
        umma_commit(&a2_empty[s]);
        umma_commit(&wp_empty[s]);
        if (j == nch - 1) umma_commit(acc_full);
      }
      __syncwarp();
    };
    for (int i = 0; i < nch; ++i) {
      const int s = i & 1;
      const uint32_t ph = (i >> 1) & 1;
      const int chunk = c_begin + i;
      if (lane == 0) {
        const int ws = MB_WIN_BUFS == 2 ? s : 0;
        const uint32_t wph = MB_WIN_BUFS == 2 ? ph : (i & 1);
        mbar_wait(&win_empty[ws], wph ^ 1);
        mbar_arrive_expect_tx(&win_full[ws], win_tx);
        tma_load_4d(s_win + ws * Cfg::kWinBytes, &tmX, &win_full[ws], chunk * kMbCC, tx * kMbTW * S - p.pad_x, ty * kMbTH * S - p.pad_y, b);
        mbar_wait(&wp_empty[s], ph ^ 1);
        mbar_arrive_expect_tx(&wp_full[s], p.cout_pad * 128);
        tma_load_2d(s_wp + s * p.wp_bytes, &tmW, &wp_full[s], chunk * kMbCC, 0);
        if (n2 > 0) tma_load_2d(s_wp + s * p.wp_bytes + p.n1 * 128, &tmW, &wp_full[s], chunk * kMbCC, p.n1);
      }
      __syncwarp();
      if (i >= 1) issue_mma(i - 1);
    }
    if (nch > 0) issue_mma(nch - 1);
  } else {
    // ------------------------------------------------------------------ compute warps: depthwise -> A2
    const int cg = tid & 7, strip = tid >> 3;
    const int sy = strip >> 2, sx = (strip & 3) * kMbP;
    for (int i = 0; i < nch; ++i) {
      const int s = i & 1;
      const uint32_t ph = (i >> 1) & 1;
      const int c0 = (c_begin + i) * kMbCC;
      for (int e = tid; e < K * K * 8; e += kMbCompute) {      // taps of this chunk -> fp32 in shared memory
        const int tap = e >> 3, g = e & 7;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (c0 + g * 8 < p.C) v = __ldg(reinterpret_cast<const uint4*>(p.dw_w + (long long)tap * p.C + c0 + g * 8));
        s_w[tap * 16 + g] = make_float4(bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y));
        s_w[tap * 16 + 8 + g] = make_float4(bf16_lo(v.z), bf16_hi(v.z), bf16_lo(v.w), bf16_hi(v.w));
      }
      if (tid < 16) {
        const int g = tid & 7, half = tid >> 3;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c0 + g * 8 < p.C) v = __ldg(reinterpret_cast<const float4*>(p.dw_b + c0 + g * 8 + half * 4));
        s_b[half * 8 + g] = v;
      }
      named_bar_sync(1, kMbCompute);
      const int ws = MB_WIN_BUFS == 2 ? s : 0;
      mbar_wait(&win_full[ws], MB_WIN_BUFS == 2 ? ph : (i & 1));
      float2 acc[kMbP][4];
      {
        const float4 b0 = s_b[cg], b1 = s_b[8 + cg];
#pragma unroll
        for (int q = 0; q < kMbP; ++q) {
          acc[q][0] = make_float2(b0.x, b0.y); acc[q][1] = make_float2(b0.z, b0.w);
          acc[q][2] = make_float2(b1.x, b1.y); acc[q][3] = make_float2(b1.z, b1.w);
        }
      }
      const uint4* win = reinterpret_cast<const uint4*>(s_win + ws * Cfg::kWinBytes);
#pragma unroll
      for (int ky = 0; ky < K; ++ky) {
        const uint4* rowp = win + ((sy * S + ky) * Cfg::WW + sx * S) * 8 + cg;
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
          for (int q = 0; q < kMbP; ++q)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[q][j] = ffma2(in[q * S + kx][j], wv[j], acc[q][j]);
        }
      }
      mbar_wait(&a2_empty[s], ph ^ 1);          // the projection MMA of chunk i - 2 has finished reading this buffer
      uint8_t* a2 = s_a2 + s * Cfg::kA2Bytes;
#pragma unroll
      for (int q = 0; q < kMbP; ++q) {
        float v[8] = {acc[q][0].x, acc[q][0].y, acc[q][1].x, acc[q][1].y, acc[q][2].x, acc[q][2].y, acc[q][3].x, acc[q][3].y};
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = fminf(fmaxf(v[j], 0.f), 6.f);          // ReLU6 (bn2 + act of the block)
        const int m = sy * kMbTW + sx + q;      // GEMM row = tile pixel; 16-byte chunk cg of its 128-byte K row, swizzled
        *reinterpret_cast<uint4*>(a2 + m * 128 + ((cg ^ (m & 7)) << 4)) =
            make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
      }
      fence_proxy_async_smem();
      named_bar_sync(1, kMbCompute);            // A2 complete, window and tap table free
      if (tid == 0) {
        mbar_arrive(&a2_full[s]);
        mbar_arrive(&win_empty[ws]);
      }
    }
    // ---------------------------------------------------------------- epilogue
    const int q = warp & 3, half = warp >> 2;
    const int m = q * 32 + lane;                // accumulator row = tile pixel
    const int oy = ty * kMbTH + m / kMbTW, ox = tx * kMbTW + m % kMbTW;
    const bool pix_ok = oy < p.Ho && ox < p.Wo;
    const long long pix = ((long long)b * p.Ho + oy) * p.Wo + ox;
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int ncol_chunks = (p.cout_pad + 31) / 32;
    for (int c = half; c < ncol_chunks; c += 2) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_acc + (static_cast<uint32_t>(q * 32) << 16) + c * 32, r);
      tmem_ld_wait();
      if (p.groups > 1) {                        // fp32 partial tile -> this group's workspace slice
        float* dst = p.ws + (((long long)tile * p.groups + grp) * 128 + m) * p.cout_pad + c * 32;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (c * 32 + j * 4 < p.cout_pad)
            *reinterpret_cast<float4*>(dst + j * 4) = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                                                  __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
      } else if (pix_ok) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {            // 8 columns per 16-byte store
          const int col = c * 32 + j * 8;
          if (col >= p.cout) break;
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[j * 8 + e]) + __ldg(p.pw_b + col + e);
          if (p.res) {
            const uint4 t = __ldg(reinterpret_cast<const uint4*>(p.res + pix * p.cout + col));
            v[0] += bf16_lo(t.x); v[1] += bf16_hi(t.x); v[2] += bf16_lo(t.y); v[3] += bf16_hi(t.y);
            v[4] += bf16_lo(t.z); v[5] += bf16_hi(t.z); v[6] += bf16_lo(t.w); v[7] += bf16_hi(t.w);
          }
          *reinterpret_cast<uint4*>(p.out + pix * p.cout + col) =
              make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
        }
      }
    }
    if (p.groups > 1) {
      // last group of this tile to arrive adds the slices in a fixed order (deterministic) and writes the output
      __threadfence();
      named_bar_sync(1, kMbCompute);
      if (tid == 0) {
        const unsigned prev = atomicAdd(p.counters + tile, 1u);
        s_last = prev == static_cast<unsigned>(p.groups - 1);
      }
      named_bar_sync(1, kMbCompute);
      if (s_last) {
        __threadfence();
        const int q4 = p.cout / 4;               // cout % 8 == 0
        const float* base = p.ws + (long long)tile * p.groups * 128 * p.cout_pad;
        for (int e = tid; e < 128 * q4; e += kMbCompute) {
          const int mm = e / q4, c4 = (e - mm * q4) * 4;
          const int yy = ty * kMbTH + mm / kMbTW, xx = tx * kMbTW + mm % kMbTW;
          if (yy >= p.Ho || xx >= p.Wo) continue;
          float4 sum = *reinterpret_cast<const float4*>(base + (long long)mm * p.cout_pad + c4);
          for (int g = 1; g < p.groups; ++g) {
            const float4 t = *reinterpret_cast<const float4*>(base + ((long long)g * 128 + mm) * p.cout_pad + c4);
            sum.x += t.x; sum.y += t.y; sum.z += t.z; sum.w += t.w;
          }
          const float4 bb = __ldg(reinterpret_cast<const float4*>(p.pw_b + c4));
          sum.x += bb.x; sum.y += bb.y; sum.z += bb.z; sum.w += bb.w;
          const long long o = (((long long)b * p.Ho + yy) * p.Wo + xx) * p.cout + c4;
          if (p.res) {
            const uint2 t = __ldg(reinterpret_cast<const uint2*>(p.res + o));
            sum.x += bf16_lo(t.x); sum.y += bf16_hi(t.x); sum.z += bf16_lo(t.y); sum.w += bf16_hi(t.y);
          }
          *reinterpret_cast<uint2*>(p.out + o) = make_uint2(pack_bf16(sum.x, sum.y), pack_bf16(sum.z, sum.w));
        }
        if (tid == 0) p.counters[tile] = 0;      // ready for the next call
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc(tmem_acc, p.tmem_cols);
  }
}

static int mb_same_pad_before(int in, int k, int s) {
  const int out = (in + s - 1) / s;
  int total = (out - 1) * s + k - in;
  if (total < 0) total = 0;
  return total / 2;
}

template <int K, int S>
static int launch_dwproject(const void* x, const void* pw_w, const MbParams& p0, int H, int W, cudaStream_t st) {
  using Cfg = MbCfg<K, S>;
  MbParams p = p0;
  auto fn = get_tensor_map_encoder();
  if (!fn) return DE_ERR_CUDA;
  CUtensorMap tmx, tmw;
  const uint32_t estr[4] = {1, 1, 1, 1};
  {
    const uint64_t dims[4] = {(uint64_t)p.C, (uint64_t)W, (uint64_t)H, (uint64_t)p.B};
    const uint64_t strides[3] = {(uint64_t)p.C * 2, (uint64_t)W * p.C * 2, (uint64_t)H * W * p.C * 2};
    const uint32_t box[4] = {kMbCC, (uint32_t)Cfg::WW, (uint32_t)Cfg::WH, 1};
    CUresult r = fn(&tmx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(DE_ERR_CUDA, "de_dwproject: cuTensorMapEncodeTiled(x) failed (%d)", (int)r);
  }
  {
    const uint64_t dims[2] = {(uint64_t)p.C, (uint64_t)p.cout};
    const uint64_t strides[1] = {(uint64_t)p.C * 2};
    const uint32_t box[2] = {kMbCC, (uint32_t)p.n1};
    CUresult r = fn(&tmw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(pw_w), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(DE_ERR_CUDA, "de_dwproject: cuTensorMapEncodeTiled(w) failed (%d)", (int)r);
  }
  p.wp_bytes = (p.cout_pad * 128 + 1023) / 1024 * 1024;
  const int smem = 1024 + MB_WIN_BUFS * Cfg::kWinBytes + 2 * Cfg::kA2Bytes + 2 * p.wp_bytes + ((Cfg::kTapBytes + 15) / 16) * 16 + 128;
  if (smem > 231424) return fail(DE_ERR_UNSUPPORTED, "de_dwproject: %d bytes of shared memory needed", smem);
  static SmemOptIn optin;
  cudaError_t e = ensure_smem(optin, dwproject_kernel<K, S>, 231424);   // 227 KB minus the static s_last
  if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(dwproject)");
  e = launch_pdl(8, dwproject_kernel<K, S>, dim3(p.ntiles * p.groups), dim3(kMbThreads), (size_t)smem, st, tmx, tmw, p);
  count_launch();
  if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(dwproject_kernel)");
  return cuda_check_launch("dwproject_kernel");
}

}  // namespace deb

using namespace deb;

static const int64_t kMbCounterBytes = 16384;

extern "C" int64_t de_dwproject_ws_bytes(int B, int H, int W, int cout) {
  if (B < 1 || H < 1 || W < 1 || cout < 8) return 0;
  const int64_t tiles = (int64_t)B * ((H + kMbTH - 1) / kMbTH) * ((W + kMbTW - 1) / kMbTW);
  const int cout_pad = (cout + 15) / 16 * 16;
  return kMbCounterBytes + (tiles <= 4096 && tiles < 100 ? tiles * 16 * 128 * cout_pad * 4 : 0);
}

extern "C" int de_dwproject(const void* x, const void* dw_w, const float* dw_b, const void* pw_w, const float* pw_b,
                            const void* residual, void* out, int B, int H, int W, int C, int cout, int k, void* ws,
                            int64_t ws_bytes, void* stream) {
  deb::ProfScope prof_("dwproject", 2.0 * B * H * W * (double)C * (k * k + cout), stream, (int64_t)B * H * W, C, cout, k);
  if (!x || !dw_w || !dw_b || !pw_w || !pw_b || !out || !ws || B < 1 || H < 1 || W < 1)
    return fail(DE_ERR_INVALID, "de_dwproject: null pointer / bad shape");
  if (C % 8 != 0 || cout % 8 != 0 || cout > 384 || (k != 3 && k != 5))
    return fail(DE_ERR_UNSUPPORTED, "de_dwproject: C=%d cout=%d k=%d (C, cout multiples of 8, cout <= 384, k 3|5, stride 1)", C, cout, k);
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(pw_w) & 15) || (reinterpret_cast<uintptr_t>(ws) & 15))
    return fail(DE_ERR_INVALID, "de_dwproject: x, pw_w and ws must be 16-byte aligned");
  MbParams p{};
  p.dw_w = reinterpret_cast<const __nv_bfloat16*>(dw_w);
  p.dw_b = dw_b;
  p.pw_b = pw_b;
  p.res = reinterpret_cast<const __nv_bfloat16*>(residual);
  p.out = reinterpret_cast<__nv_bfloat16*>(out);
  p.B = B; p.C = C; p.Ho = H; p.Wo = W; p.cout = cout;
  p.pad_y = mb_same_pad_before(H, k, 1);
  p.pad_x = mb_same_pad_before(W, k, 1);
  p.cout_pad = (cout + 15) / 16 * 16;
  p.n1 = p.cout_pad <= 256 ? p.cout_pad : 192;
  if (p.cout_pad > 256 && p.cout_pad - p.n1 > 256) return fail(DE_ERR_UNSUPPORTED, "de_dwproject: cout %d", cout);
  p.tmem_cols = 32;
  while (p.tmem_cols < p.cout_pad) p.tmem_cols <<= 1;
  p.tiles_x = (W + kMbTW - 1) / kMbTW;
  p.tiles_y = (H + kMbTH - 1) / kMbTH;
  p.ntiles = B * p.tiles_y * p.tiles_x;
  p.chunks = (C + kMbCC - 1) / kMbCC;
  int groups = 1;
  if (p.ntiles < 100 && p.ntiles <= 4096) {
    static int target = -1;
    if (target < 0) {
      const char* e = getenv("DEB_MB_CTAS");
      target = e ? atoi(e) : num_sms();
    }
    groups = target / p.ntiles;
    if (groups > 16) groups = 16;
    if (groups > p.chunks) groups = p.chunks;
    if (groups < 1) groups = 1;
  }
  p.cpg = (p.chunks + groups - 1) / groups;
  p.groups = (p.chunks + p.cpg - 1) / p.cpg;
  p.counters = reinterpret_cast<unsigned*>(ws);
  p.ws = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + kMbCounterBytes);
  if (p.groups > 1) {
    const int64_t need = kMbCounterBytes + (int64_t)p.ntiles * p.groups * 128 * p.cout_pad * 4;
    if (ws_bytes < need) return fail(DE_ERR_INVALID, "de_dwproject: workspace %lld < %lld bytes (de_dwproject_ws_bytes)", (long long)ws_bytes, (long long)need);
  } else if (ws_bytes < kMbCounterBytes) {
    return fail(DE_ERR_INVALID, "de_dwproject: workspace smaller than %lld bytes", (long long)kMbCounterBytes);
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (k == 3) return launch_dwproject<3, 1>(x, pw_w, p, H, W, st);
  return launch_dwproject<5, 1>(x, pw_w, p, H, W, st);
}
