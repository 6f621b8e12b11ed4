// Fused DPT head tail for sm_100a: conv3x3(Cin -> 32, pad 1) + bias + ReLU + conv1x1(32 -> 1) + bias + ReLU on an
// NHWC bf16 activation, fp32 [B,H,W] out  ([UPSTREAM] midas/dpt_depth.py head: output_conv[2..5]).
//
// Why not the general implicit-GEMM kernel: with N = 32 the MMAs are short (128x32x16) and the general kernel
// re-fetches the A tile from L2 for each of the 9 filter taps - 85 B/clk/SM of L2->smem traffic, which is what bounds
// it (1.59 ms per 32-frame DPT_Large step, 219 TFLOP/s).  Here
//   * the whole filter bank (9*Cin x 32 bf16 = 72 KB at Cin = 128) is loaded into shared memory once per CTA;
//   * the activation is staged as HALO tiles: for an 8 x 16 pixel output tile, three TMA boxes per 64-channel block,
//     one per horizontal tap dx, each (8+2) rows x 16 pixels x 64 channels.  The three vertical taps of a box are the
//     same shared memory viewed 16 rows (2048 B = two swizzle atoms) further down, so a box feeds 3 taps x 4 K-steps
//     = 12 MMAs and the L2->smem traffic per tile drops from 9 to 3.75 tile-equivalents;
//   * accumulators live in 4 TMEM stages of 32 columns; four epilogue warps apply bias/ReLU, the 32->1 dot and the
//     final ReLU straight from TMEM and write one fp32 per pixel.
// Roles: warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator, warps 4..7 epilogue.  Persistent CTAs, one
// per SM, tiles handed out round-robin (x fastest, so concurrently running CTAs share halo rows in L2).
//
// Three kernels live here: head_conv_kernel (the above, on an activation that already has the output resolution),
// head_up_conv_kernel (round 1: the x2 bilinear upsample fused in front, three dx copies of the halo tile) and
// head_up3_kernel (round 2, what de_head_up_conv runs: the horizontal taps in the MMA's N dimension - see its comment).
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

constexpr int kHcBW = 16, kHcBH = 8;                        // output tile: 8 rows x 16 pixels = 128 accumulator rows
constexpr int kHcBoxRows = kHcBH + 2;
constexpr int kHcStageBytes = kHcBoxRows * kHcBW * 128;     // 20480: (10 x 16) pixels x 64 channels bf16
constexpr int kHcN = 32;
constexpr int kHcWChunk = kHcN * 128;                       // 4096: 32 filters x 64 K bf16
constexpr int kHcAcc = 4;                                   // TMEM accumulator stages
constexpr int kHcThreads = 256;
constexpr int kHcMaxStages = 8;

struct HeadConvParams {
  int B, H, W, cin_blocks, tiles_x, tiles_y, num_tiles, stages;
  const float* bias;
  const float* head_w;
  float head_b;
  float* out;
};

__global__ void __launch_bounds__(kHcThreads, 1)
head_conv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const HeadConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  const int w_chunks = 9 * p.cin_blocks;
  uint8_t* sW = smem;                                        // [tap][cb] chunks of 4096 B
  uint8_t* sA = smem + w_chunks * kHcWChunk;                 // stage ring
  uint8_t* post = sA + p.stages * kHcStageBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(post);
  uint64_t* empty = full + kHcMaxStages;
  uint64_t* tfull = empty + kHcMaxStages;
  uint64_t* tempty = tfull + kHcAcc;
  uint64_t* wfull = tempty + kHcAcc;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 1);
  float* s_bias = reinterpret_cast<float*>(post + 256);      // 32 bias + 32 head weights
  float* s_hw = s_bias + kHcN;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < p.stages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < kHcAcc; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 4);
    }
    mbar_init(wfull, 1);
    mbar_fence_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, kHcAcc * kHcN);
    tmem_relinquish();
  }
  if (warp == 3) {
    s_bias[lane] = p.bias ? __ldg(p.bias + lane) : 0.f;
    s_hw[lane] = __ldg(p.head_w + lane);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int per_img = p.tiles_x * p.tiles_y;
  const int stages_per_tile = 3 * p.cin_blocks;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      mbar_arrive_expect_tx(wfull, w_chunks * kHcWChunk);
      for (int c = 0; c < w_chunks; ++c) tma_load_2d(sW + c * kHcWChunk, &tmW, wfull, c * 64, 0);
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        const int b = tile / per_img, t = tile % per_img;
        const int y0 = (t / p.tiles_x) * kHcBH, x0 = (t % p.tiles_x) * kHcBW;
        for (int s = 0; s < stages_per_tile; ++s) {
          const int cb = s / 3, dx = s % 3;
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], kHcStageBytes);
          tma_load_4d(sA + stage * kHcStageBytes, &tmA, &full[stage], cb * 64, x0 + dx - 1, y0 - 1, b);
          if (++stage == p.stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = make_idesc_bf16(128, kHcN);
    mbar_wait(wfull, 0);
    tc_fence_after();
    int stage = 0;
    uint32_t phase = 0;
    int as = 0;
    uint32_t aphase = 0;
    const uint32_t w_base = smem_u32(sW);
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      mbar_wait(&tempty[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * kHcN;
      for (int s = 0; s < stages_per_tile; ++s) {
        const int cb = s / 3, dx = s % 3;
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        // warp-uniform descriptors + elect_one: back-to-back UTCHMMA issue (see gemm_tcgen05.cu)
        const uint64_t adesc = make_sw128_kmajor_desc(smem_u32(sA + stage * kHcStageBytes));
        const uint64_t bdesc = make_sw128_kmajor_desc(w_base + (dx * p.cin_blocks + cb) * kHcWChunk);
        const uint64_t b_dy = static_cast<uint64_t>((3 * p.cin_blocks * kHcWChunk) >> 4);   // next filter row (dy+1)
        if (elect_one()) {
#pragma unroll
          for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16(d_tmem, adesc + dy * ((kHcBW * 128) >> 4) + 2 * k, bdesc + dy * b_dy + 2 * k, idesc,
                        (s | dy | k) != 0);
          }
          umma_commit(&empty[stage]);
          if (s == stages_per_tile - 1) umma_commit(&tfull[as]);
        }
        __syncwarp();
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (++as == kHcAcc) {
        as = 0;
        aphase ^= 1;
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ epilogue: one accumulator row per thread
    const int q = warp & 3;
    const int r = q * 32 + lane;
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      const int b = tile / per_img, t = tile % per_img;
      const int y = (t / p.tiles_x) * kHcBH + r / kHcBW, x = (t % p.tiles_x) * kHcBW + r % kHcBW;
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
      uint32_t acc[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * kHcN, acc);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);
      float s0 = p.head_b, s1 = 0.f;
#pragma unroll
      for (int n = 0; n < kHcN; n += 2) {
        s0 = fmaf(fmaxf(__uint_as_float(acc[n]) + s_bias[n], 0.f), s_hw[n], s0);
        s1 = fmaf(fmaxf(__uint_as_float(acc[n + 1]) + s_bias[n + 1], 0.f), s_hw[n + 1], s1);
      }
      if (y < p.H && x < p.W) p.out[((long long)b * p.H + y) * p.W + x] = fmaxf(s0 + s1, 0.f);
      if (++as == kHcAcc) {
        as = 0;
        aphase ^= 1;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kHcAcc * kHcN);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Same head tail with the bilinear x2 upsample in front of it fused in ([UPSTREAM] output_conv[1..5]): the kernel
// reads the LOW-resolution activation and never materialises the upsampled one (1.2 GB written + read back per
// 32-frame DPT_Large step; the upsample kernel and the L2->smem feed of the kernel above were the two costs left).
// Eight producer warps replace the TMA producer: per (tile, 64-channel block) they interpolate the 10 x 18 pixel halo
// once (fp32 lerps, rounded to bf16 exactly like de_upsample2x) and write it to shared memory three times - one copy
// per horizontal tap dx, in the 128B-swizzled K-major layout the UMMA descriptors expect - then publish the unit
// with fence.proxy.async + an mbarrier arrive per warp.  MMA issue, TMEM stages and epilogue are as above.
constexpr int kHuProdWarps = 10;                            // two groups of 5 warps (upper / lower 5 halo rows), each 18 halo
constexpr int kHuThreads = 256 + 32 * kHuProdWarps;        // columns x 8 channel groups = 144 of 160 threads busy
constexpr int kHuUnitBytes = 3 * kHcStageBytes;             // three dx copies of one 64-channel block
constexpr int kHuUnits = 2;
// low-resolution source window of one (tile, 64-channel block): the 10 x 18 halo spans <= 4.5 x 8.5 source pixels, i.e.
// <= 7 x 11 taps with the +1 neighbours; staged by TMA (no swizzle, 128 B per pixel) two units ahead of the producers
constexpr int kHuSrcRows = 7, kHuSrcCols = 12;
constexpr int kHuSrcBytes = kHuSrcRows * kHuSrcCols * 128;  // 10752
constexpr int kHuSrcSlotBytes = 11264;                      // 11 KB slots keep everything 1024-aligned
constexpr int kHuSrcSlots = 2;

struct HeadUpParams {
  const __nv_bfloat16* x;     // [B, Hs, Ws, Cin]
  int B, Hs, Ws, H, W, Cin, cin_blocks, tiles_x, tiles_y, num_tiles, align_corners;
  float rh, rw;
  const float* bias;
  const float* head_w;
  float head_b;
  float* out;
};

// first source row / column the halo of a tile touches (same expression in the TMA warp and in the producers)
DEV int hu_src_base(int o0, float r, int align_corners) {
  const int o = o0 < 0 ? 0 : o0;
  const float s = align_corners ? r * o : fmaxf(0.5f * (o + 0.5f) - 0.5f, 0.f);
  return static_cast<int>(s);
}

__global__ void __launch_bounds__(kHuThreads, 1)
head_up_conv_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, const HeadUpParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  const int w_chunks = 9 * p.cin_blocks;
  uint8_t* sW = smem;
  uint8_t* sA = smem + w_chunks * kHcWChunk;                 // kHuUnits units of three 20 KB copies
  uint8_t* sX = sA + kHuUnits * kHuUnitBytes;                // source windows
  uint8_t* post = sX + kHuSrcSlots * kHuSrcSlotBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(post);
  uint64_t* empty = full + kHuUnits;
  uint64_t* tfull = empty + kHuUnits;
  uint64_t* tempty = tfull + kHcAcc;
  uint64_t* wfull = tempty + kHcAcc;
  uint64_t* xfull = wfull + 1;
  uint64_t* xempty = xfull + kHuSrcSlots;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(xempty + kHuSrcSlots);
  float* s_bias = reinterpret_cast<float*>(post + 256);
  float* s_hw = s_bias + kHcN;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmX);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kHuUnits; ++i) {
      mbar_init(&full[i], kHuProdWarps);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < kHuSrcSlots; ++i) {
      mbar_init(&xfull[i], 1);
      mbar_init(&xempty[i], kHuProdWarps);
    }
    for (int i = 0; i < kHcAcc; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 4);
    }
    mbar_init(wfull, 1);
    mbar_fence_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, kHcAcc * kHcN);
    tmem_relinquish();
  }
  if (warp == 3) {
    s_bias[lane] = p.bias ? __ldg(p.bias + lane) : 0.f;
    s_hw[lane] = __ldg(p.head_w + lane);
  }
  pdl_launch();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int per_img = p.tiles_x * p.tiles_y;
  const int units_per_tile = p.cin_blocks;
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {                                         // filter bank once, then the source windows
      mbar_arrive_expect_tx(wfull, w_chunks * kHcWChunk);
      for (int c = 0; c < w_chunks; ++c) tma_load_2d(sW + c * kHcWChunk, &tmW, wfull, c * 64, 0);
      int xs = 0;
      uint32_t xphase = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        const int b = tile / per_img, t = tile % per_img;
        const int ysb = hu_src_base((t / p.tiles_x) * kHcBH - 1, p.rh, p.align_corners);
        const int xsb = hu_src_base((t % p.tiles_x) * kHcBW - 1, p.rw, p.align_corners);
        for (int cb = 0; cb < units_per_tile; ++cb) {
          mbar_wait(&xempty[xs], xphase ^ 1);
          mbar_arrive_expect_tx(&xfull[xs], kHuSrcBytes);
          tma_load_4d(sX + xs * kHuSrcSlotBytes, &tmX, &xfull[xs], cb * 64, xsb, ysb, b);
          if (++xs == kHuSrcSlots) {
            xs = 0;
            xphase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = make_idesc_bf16(128, kHcN);
    mbar_wait(wfull, 0);
    tc_fence_after();
    int u = 0;
    uint32_t phase = 0;
    int as = 0;
    uint32_t aphase = 0;
    const uint32_t w_base = smem_u32(sW);
    const uint64_t b_dy = static_cast<uint64_t>((3 * p.cin_blocks * kHcWChunk) >> 4);
    const uint64_t b_dx = static_cast<uint64_t>((p.cin_blocks * kHcWChunk) >> 4);
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      mbar_wait(&tempty[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * kHcN;
      for (int cb = 0; cb < units_per_tile; ++cb) {
        mbar_wait(&full[u], phase);
        tc_fence_after();
        const uint64_t adesc = make_sw128_kmajor_desc(smem_u32(sA + u * kHuUnitBytes));
        const uint64_t bdesc = make_sw128_kmajor_desc(w_base + cb * kHcWChunk);
        if (elect_one()) {
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) {
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma_bf16(d_tmem, adesc + dx * (kHcStageBytes >> 4) + dy * ((kHcBW * 128) >> 4) + 2 * k,
                          bdesc + dx * b_dx + dy * b_dy + 2 * k, idesc, (cb | dx | dy | k) != 0);
            }
          }
          umma_commit(&empty[u]);
          if (cb == units_per_tile - 1) umma_commit(&tfull[as]);
        }
        __syncwarp();
        if (++u == kHuUnits) {
          u = 0;
          phase ^= 1;
        }
      }
      if (++as == kHcAcc) {
        as = 0;
        aphase ^= 1;
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ------------------------------------------------------------------ epilogue: one accumulator row per thread
    const int q = warp & 3;
    const int r = q * 32 + lane;
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      const int b = tile / per_img, t = tile % per_img;
      const int y = (t / p.tiles_x) * kHcBH + r / kHcBW, x = (t % p.tiles_x) * kHcBW + r % kHcBW;
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
      uint32_t acc[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * kHcN, acc);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);
      float s0 = p.head_b, s1 = 0.f;
#pragma unroll
      for (int n = 0; n < kHcN; n += 2) {
        s0 = fmaf(fmaxf(__uint_as_float(acc[n]) + s_bias[n], 0.f), s_hw[n], s0);
        s1 = fmaf(fmaxf(__uint_as_float(acc[n + 1]) + s_bias[n + 1], 0.f), s_hw[n + 1], s1);
      }
      if (y < p.H && x < p.W) p.out[((long long)b * p.H + y) * p.W + x] = fmaxf(s0 + s1, 0.f);
      if (++as == kHcAcc) {
        as = 0;
        aphase ^= 1;
      }
    }
  } else if (warp >= 8) {
    // ------------------------------------------------------------------ producers: interpolate + swizzled smem writes
    // Thread = (halo column hx, 8-channel group q, upper / lower half of the 10 halo rows).  The column's horizontal
    // taps and weights are fixed for the whole tile; walking down its 5 rows, the horizontally interpolated source row
    // (8 fp32) is kept and reused while the source row index repeats (x2 upsampling: every other step), so a thread
    // reads ~8 instead of 20 16-byte vectors per unit.  Shared-memory bandwidth is what bounds this kernel (the
    // N = 32 MMAs alone read 5 KB each), so every avoided LDS / STS counts.
    const int ptid = threadIdx.x - 256;
    const int half = ptid / 160, pin = ptid - half * 160;    // warp-uniform row half: no divergence in the row walk
    const int q = pin & 7;
    const int hx = pin >> 3;                                 // 0..19, columns >= 18 idle
    const bool active = hx < kHcBW + 2;
    const int hy0 = half * (kHcBoxRows / 2);
    // byte offsets of this thread's 16-byte chunk inside the three dx copies (row = hy*16 + j and 16 % 8 == 0, so the
    // swizzle phase row & 7 is j & 7 for every halo row); -1: this column does not exist in that copy
    int soff[3];
#pragma unroll
    for (int dx = 0; dx < 3; ++dx) {
      const int j = hx - dx;
      soff[dx] = (active && j >= 0 && j < kHcBW) ? dx * kHcStageBytes + j * 128 + ((q ^ (j & 7)) << 4) : -1;
    }
    int u = 0, xs = 0;
    uint32_t phase = 0, xphase = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      const int t = tile % per_img;
      const int y0 = (t / p.tiles_x) * kHcBH - 1, x0 = (t % p.tiles_x) * kHcBW - 1;
      const int ysb = hu_src_base(y0, p.rh, p.align_corners), xsb = hu_src_base(x0, p.rw, p.align_corners);
      const int ox = x0 + hx;
      const bool col_ok = active && ox >= 0 && ox < p.W;
      const float sx = p.align_corners ? p.rw * ox : fmaxf(0.5f * (ox + 0.5f) - 0.5f, 0.f);
      const int xs0 = min(static_cast<int>(sx), p.Ws - 1), xs1 = min(xs0 + 1, p.Ws - 1);
      const float lx = sx - xs0, hxw = 1.f - lx;
      const int c0 = (xs0 - xsb) * 128 + q * 16, c1 = (xs1 - xsb) * 128 + q * 16;
      for (int cb = 0; cb < units_per_tile; ++cb) {
        mbar_wait(&xfull[xs], xphase);                       // source window landed (TMA)
        mbar_wait(&empty[u], phase ^ 1);                     // the MMAs are done with this unit's three copies
        uint8_t* ubase = sA + u * kHuUnitBytes;
        const uint8_t* src = sX + xs * kHuSrcSlotBytes;
        float h0[8], h1[8];
        int cy0 = -1, cy1 = -1;                                // source rows held in h0 / h1
        auto hrow = [&](int ys, float* h) {
          const uint8_t* rp = src + (ys - ysb) * (kHuSrcCols * 128);
          const uint4 a = *reinterpret_cast<const uint4*>(rp + c0);
          const uint4 bq = *reinterpret_cast<const uint4*>(rp + c1);
          const uint32_t* pa = &a.x;
          const uint32_t* pb = &bq.x;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            h[2 * k] = hxw * bf16_lo(pa[k]) + lx * bf16_lo(pb[k]);
            h[2 * k + 1] = hxw * bf16_hi(pa[k]) + lx * bf16_hi(pb[k]);
          }
        };
#pragma unroll
        for (int i = 0; i < kHcBoxRows / 2; ++i) {
          const int hy = hy0 + i;
          const int oy = y0 + hy;
          uint4 val = make_uint4(0u, 0u, 0u, 0u);
          if (col_ok && oy >= 0 && oy < p.H) {
            const float sy = p.align_corners ? p.rh * oy : fmaxf(0.5f * (oy + 0.5f) - 0.5f, 0.f);
            const int ys0 = min(static_cast<int>(sy), p.Hs - 1), ys1 = min(ys0 + 1, p.Hs - 1);
            const float ly = sy - ys0, hyw = 1.f - ly;
            if (ys0 != cy0) {
              if (ys0 == cy1) {
#pragma unroll
                for (int k = 0; k < 8; ++k) h0[k] = h1[k];
              } else {
                hrow(ys0, h0);
              }
              cy0 = ys0;
            }
            if (ys1 != cy1) {
              if (ys1 == cy0) {
#pragma unroll
                for (int k = 0; k < 8; ++k) h1[k] = h0[k];
              } else {
                hrow(ys1, h1);
              }
              cy1 = ys1;
            }
            uint32_t o[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)                        // same expression tree as upsample2x_kernel (misc.cu)
              o[k] = pack_bf16(hyw * h0[2 * k] + ly * h1[2 * k], hyw * h0[2 * k + 1] + ly * h1[2 * k + 1]);
            val = make_uint4(o[0], o[1], o[2], o[3]);
          }
#pragma unroll
          for (int dx = 0; dx < 3; ++dx)
            if (soff[dx] >= 0) *reinterpret_cast<uint4*>(ubase + soff[dx] + hy * (kHcBW * 128)) = val;
        }
        fence_proxy_async_smem();                              // generic-proxy writes -> visible to the UMMA (async proxy)
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&full[u]);
          mbar_arrive(&xempty[xs]);
        }
        if (++u == kHuUnits) {
          u = 0;
          phase ^= 1;
        }
        if (++xs == kHuSrcSlots) {
          xs = 0;
          xphase ^= 1;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kHcAcc * kHcN);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Second formulation of the fused upsample + head tail: the three HORIZONTAL taps ride in the N dimension.
// With N = 32 every 128x32x16 MMA reads 4 KB of A and 1 KB of B from shared memory for 8 clocks of math: shared-memory
// bandwidth (128 B/clk) bounds the kernel above at ~17 % tensor-pipe utilisation, and three dx copies of every halo
// tile have to be written first.  Here the B operand of a (dy, 64-channel block) step is [dx0 | dx1 | dx2] x 32 filters
// = 96 columns, so ONE copy of the halo tile is read once per dy for all three dx taps (7 KB per MMA for 3x the math):
//     P_dx[r, j] = sum_dy sum_c  halo[r + dy, j, c] * w[dy, dx, c, :]             (TMEM, 96 fp32 columns per pixel)
//     out[r, j]  = P_0[r, j - 1] + P_1[r, j] + P_2[r, j + 1]                        (epilogue: two warp shuffles per value)
// The accumulator rows are 4 image rows x 32 columns (a warp = one image row, lane = column), columns 0 and 31 are the
// halo columns: a tile yields 4 x 30 outputs.  Shared-memory traffic per output drops from ~4.1 KB to ~2.3 KB.
constexpr int kH3TR = 4, kH3TC = 32, kH3OutC = kH3TC - 2;
constexpr int kH3HaloRows = kH3TR + 2;
constexpr int kH3UnitBytes = kH3HaloRows * kH3TC * 128;    // 24576: (6 x 32) pixels x 64 channels bf16, one copy
constexpr int kH3Units = 4;
constexpr int kH3N = 3 * kHcN;                              // 96 accumulator columns
constexpr int kH3Acc = 4;                                   // TMEM stages (4 x 96 = 384 of 512 allocated columns)
constexpr int kH3SrcRows = 5, kH3SrcCols = 18;              // source window of the 6 x 32 halo (x2: <= 3+2 rows, 16+2 columns)
constexpr int kH3SrcBytes = kH3SrcRows * kH3SrcCols * 128;  // 11520
constexpr int kH3SrcSlotBytes = 12288;
constexpr int kH3SrcSlots = 4;
#ifndef H3_GROUPS
#define H3_GROUPS 2                                         // producer groups of 8 warps; group g interpolates units g, g + 2, ...
#endif
constexpr int kH3Groups = H3_GROUPS;
constexpr int kH3GroupWarps = 8;
constexpr int kH3ProdWarps = kH3GroupWarps * kH3Groups;
constexpr int kH3Threads = 256 + 32 * kH3ProdWarps;

DEV void tmem_ld_32x16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

__global__ void __launch_bounds__(kH3Threads, 1)
head_up3_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, const HeadUpParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  const int w_chunks = 9 * p.cin_blocks;
  uint8_t* sW = smem;                                        // chunk ((dy * cin_blocks + cb) * 3 + dx): 32 filters x 64 K
  uint8_t* sA = smem + w_chunks * kHcWChunk;                 // kH3Units halo tiles
  uint8_t* sX = sA + kH3Units * kH3UnitBytes;                // source windows
  uint8_t* post = sX + kH3SrcSlots * kH3SrcSlotBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(post);
  uint64_t* empty = full + kH3Units;
  uint64_t* tfull = empty + kH3Units;
  uint64_t* tempty = tfull + kH3Acc;
  uint64_t* wfull = tempty + kH3Acc;
  uint64_t* xfull = wfull + 1;
  uint64_t* xempty = xfull + kH3SrcSlots;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(xempty + kH3SrcSlots);
  float* s_bias = reinterpret_cast<float*>(post + 256);
  float* s_hw = s_bias + kHcN;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmX);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kH3Units; ++i) {
      mbar_init(&full[i], kH3GroupWarps);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < kH3SrcSlots; ++i) {
      mbar_init(&xfull[i], 1);
      mbar_init(&xempty[i], kH3GroupWarps);
    }
    for (int i = 0; i < kH3Acc; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 4);
    }
    mbar_init(wfull, 1);
    mbar_fence_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  if (warp == 3) {
    s_bias[lane] = p.bias ? __ldg(p.bias + lane) : 0.f;
    s_hw[lane] = __ldg(p.head_w + lane);
  }
  pdl_launch();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int per_img = p.tiles_x * p.tiles_y;
  const int units_per_tile = p.cin_blocks;
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {                                         // filter bank once (dx chunks of a (dy, cb) step adjacent), then the source windows
      mbar_arrive_expect_tx(wfull, w_chunks * kHcWChunk);
      for (int dy = 0; dy < 3; ++dy)
        for (int cb = 0; cb < p.cin_blocks; ++cb)
          for (int dx = 0; dx < 3; ++dx)
            tma_load_2d(sW + ((dy * p.cin_blocks + cb) * 3 + dx) * kHcWChunk, &tmW, wfull, ((dy * 3 + dx) * p.cin_blocks + cb) * 64, 0);
      int xs = 0;
      uint32_t xphase = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        const int b = tile / per_img, t = tile % per_img;
        const int ysb = hu_src_base((t / p.tiles_x) * kH3TR - 1, p.rh, p.align_corners);
        const int xsb = hu_src_base((t % p.tiles_x) * kH3OutC - 1, p.rw, p.align_corners);
        for (int cb = 0; cb < units_per_tile; ++cb) {
          mbar_wait(&xempty[xs], xphase ^ 1);
          mbar_arrive_expect_tx(&xfull[xs], kH3SrcBytes);
          tma_load_4d(sX + xs * kH3SrcSlotBytes, &tmX, &xfull[xs], cb * 64, xsb, ysb, b);
          if (++xs == kH3SrcSlots) {
            xs = 0;
            xphase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer: 3 dy x 4 k-steps of 128 x 96 x 16 per unit
    constexpr uint32_t idesc = make_idesc_bf16(128, kH3N);
    mbar_wait(wfull, 0);
    tc_fence_after();
    int u = 0;
    uint32_t phase = 0;
    int as = 0;
    uint32_t aphase = 0;
    const uint32_t w_base = smem_u32(sW);
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      mbar_wait(&tempty[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * kH3N;
      for (int cb = 0; cb < units_per_tile; ++cb) {
        mbar_wait(&full[u], phase);
        tc_fence_after();
        const uint64_t adesc = make_sw128_kmajor_desc(smem_u32(sA + u * kH3UnitBytes));
        if (elect_one()) {
#pragma unroll
          for (int dy = 0; dy < 3; ++dy) {
            const uint64_t bdesc = make_sw128_kmajor_desc(w_base + ((dy * p.cin_blocks + cb) * 3) * kHcWChunk);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16(d_tmem, adesc + dy * ((kH3TC * 128) >> 4) + 2 * k, bdesc + 2 * k, idesc, (cb | dy | k) != 0);
          }
          umma_commit(&empty[u]);
          if (cb == units_per_tile - 1) umma_commit(&tfull[as]);
        }
        __syncwarp();
        if (++u == kH3Units) {
          u = 0;
          phase ^= 1;
        }
      }
      if (++as == kH3Acc) {
        as = 0;
        aphase ^= 1;
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ------------------------------------------------------------------ epilogue: warp = image row of the tile, lane = column
    const int q = warp & 3;
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      const int b = tile / per_img, t = tile % per_img;
      const int y = (t / p.tiles_x) * kH3TR + q, x = (t % p.tiles_x) * kH3OutC - 1 + lane;
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
      const uint32_t tacc = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * kH3N;
      float s0 = p.head_b, s1 = 0.f;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t a0[16], a1[16], a2[16];
        tmem_ld_32x16(tacc + h * 16, a0);
        tmem_ld_32x16(tacc + kHcN + h * 16, a1);
        tmem_ld_32x16(tacc + 2 * kHcN + h * 16, a2);
        tmem_ld_wait();
        if (h == 1) {                                        // the accumulator stage is in registers
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[as]);
        }
#pragma unroll
        for (int n = 0; n < 16; n += 2) {
          const float l0 = __shfl_up_sync(0xffffffffu, __uint_as_float(a0[n]), 1);
          const float l1 = __shfl_up_sync(0xffffffffu, __uint_as_float(a0[n + 1]), 1);
          const float r0 = __shfl_down_sync(0xffffffffu, __uint_as_float(a2[n]), 1);
          const float r1 = __shfl_down_sync(0xffffffffu, __uint_as_float(a2[n + 1]), 1);
          const float v0 = (l0 + __uint_as_float(a1[n])) + r0;
          const float v1 = (l1 + __uint_as_float(a1[n + 1])) + r1;
          s0 = fmaf(fmaxf(v0 + s_bias[h * 16 + n], 0.f), s_hw[h * 16 + n], s0);
          s1 = fmaf(fmaxf(v1 + s_bias[h * 16 + n + 1], 0.f), s_hw[h * 16 + n + 1], s1);
        }
      }
      if (lane >= 1 && lane <= kH3OutC && y < p.H && x < p.W) p.out[((long long)b * p.H + y) * p.W + x] = fmaxf(s0 + s1, 0.f);
      if (++as == kH3Acc) {
        as = 0;
        aphase ^= 1;
      }
    }
  } else if (warp >= 8) {
    // ------------------------------------------------------------------ producers: interpolate the halo tile once, swizzled K-major
    // Thread = (halo column j of 32, 8-channel group q) of one producer GROUP (8 warps = one whole unit); a warp covers
    // 4 columns x 8 channel groups = four full 128-byte rows per store (conflict free).  Unit n of the CTA belongs to
    // group n % kH3Groups, so one group's LDS -> math -> STS -> fence.proxy.async -> arrive chain runs under the other
    // group's arithmetic.  The kernel is issue-bound (ncu: 72 % issue-active, 13 k warp instructions per tile in the
    // first version of this loop), so the loop is built to execute few instructions:
    //   * the <= 5 source rows of the window are interpolated horizontally ONCE per unit into registers (H[0..4]); a
    //     halo row then picks the pair (H[w], H[w+1]) with a warp-uniform switch - no re-loads, no register shuffling;
    //   * the per-row source index / weight is computed by lane hy of every warp once per tile and broadcast by
    //     shuffles (every thread of the CTA needs the same six values); tile coordinates advance incrementally;
    //   * columns outside the image get zero horizontal weights (their bytes come out as +-0) instead of a per-row test.
    // Arithmetic: FMUL2 / FFMA2 in exactly the operation order of upsample2x_kernel (t = l * second; h * first + t), so
    // the interpolated bf16 values stay bit-identical to the stand-alone upsample.
    static_assert(kH3Units % kH3Groups == 0 && kH3SrcSlots == kH3Units, "ring slots are owned by unit index");
    const int ptid = threadIdx.x - 256;
    const int q = ptid & 7;
    const int j = (ptid >> 3) & 31;
    const int grp = ptid >> 8;
    const int soff = j * 128 + ((q ^ (j & 7)) << 4);         // row = hy * 32 + j: swizzle phase (row & 7) = j & 7
    constexpr int kRowBytes = kH3SrcCols * 128;
    int n = 0;                                               // unit counter of the CTA (all groups count every unit)
    int tx, ty;
    {
      const int t = static_cast<int>(blockIdx.x) % per_img;
      ty = t / p.tiles_x;
      tx = t - ty * p.tiles_x;
    }
    const int dtx = static_cast<int>(gridDim.x) % p.tiles_x, dty = (static_cast<int>(gridDim.x) / p.tiles_x) % p.tiles_y;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      const int y0 = ty * kH3TR - 1, x0 = tx * kH3OutC - 1;
      tx += dtx;
      if (tx >= p.tiles_x) {
        tx -= p.tiles_x;
        ++ty;
      }
      ty += dty;
      if (ty >= p.tiles_y) ty -= p.tiles_y;
      const int ysb = hu_src_base(y0, p.rh, p.align_corners), xsb = hu_src_base(x0, p.rw, p.align_corners);
      const int ox = x0 + j;
      const bool col_ok = ox >= 0 && ox < p.W;
      const float sx = p.align_corners ? p.rw * ox : fmaxf(0.5f * (ox + 0.5f) - 0.5f, 0.f);
      const int xs0 = max(min(static_cast<int>(sx), p.Ws - 1), 0), xs1 = min(xs0 + 1, p.Ws - 1);
      const float lx = col_ok ? sx - xs0 : 0.f, hxw = col_ok ? 1.f - lx : 0.f;
      const float2 lx2 = make_float2(lx, lx), hx2 = make_float2(hxw, hxw);
      const int c0 = min(max(xs0 - xsb, 0), kH3SrcCols - 1) * 128 + q * 16, c1 = min(max(xs1 - xsb, 0), kH3SrcCols - 1) * 128 + q * 16;
      // lane hy computes halo row hy: window row of its upper source row (+1; 0: row outside the image) and weight;
      // the six indices travel as one packed word, the weight is fetched from its lane when the row is written
      float ly_lane;
      uint32_t wpack = 0;
      {
        const int oy = y0 + lane;
        const float sy = p.align_corners ? p.rh * oy : fmaxf(0.5f * (oy + 0.5f) - 0.5f, 0.f);
        const int ys0 = min(static_cast<int>(sy), p.Hs - 1);
        ly_lane = sy - ys0;
        const int w1 = (oy >= 0 && oy < p.H) ? min(max(ys0 - ysb, 0), kH3SrcRows - 2) + 1 : 0;
#pragma unroll
        for (int hy = 0; hy < kH3HaloRows; ++hy) wpack |= static_cast<uint32_t>(__shfl_sync(0xffffffffu, w1, hy)) << (4 * hy);
      }
      // any row on window rows (3, 4)?  (w1 == 4 in some nibble)
      const bool need4 = ((wpack & 0x444444u) != 0u);
      // byte offset of window row w: source row min(ysb + w, Hs - 1), i.e. the clamped "+1" neighbour at the bottom border
      int offw[kH3SrcRows];
#pragma unroll
      for (int w = 0; w < kH3SrcRows; ++w) offw[w] = (min(ysb + w, p.Hs - 1) - ysb) * kRowBytes;
      for (int cb = 0; cb < units_per_tile; ++cb, ++n) {
        if (n % kH3Groups != grp) continue;
        const int u = n % kH3Units;
        const uint32_t par = (n / kH3Units) & 1;
        mbar_wait(&xfull[u], par);                           // source window landed (TMA)
        const uint8_t* src = sX + u * kH3SrcSlotBytes;
        float2 H[kH3SrcRows][4];
#pragma unroll
        for (int w = 0; w < kH3SrcRows; ++w) {
          if (w == kH3SrcRows - 1 && !need4) break;
          const uint4 a = *reinterpret_cast<const uint4*>(src + offw[w] + c0);
          const uint4 bq = *reinterpret_cast<const uint4*>(src + offw[w] + c1);
          const uint32_t* pa = &a.x;
          const uint32_t* pb = &bq.x;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            H[w][k] = ffma2(make_float2(bf16_lo(pa[k]), bf16_hi(pa[k])), hx2, fmul2(make_float2(bf16_lo(pb[k]), bf16_hi(pb[k])), lx2));
        }
        mbar_wait(&empty[u], par ^ 1);                       // the MMAs are done with this unit
        uint8_t* ubase = sA + u * kH3UnitBytes + soff;
#pragma unroll
        for (int hy = 0; hy < kH3HaloRows; ++hy) {
          uint4 val = make_uint4(0u, 0u, 0u, 0u);
          const int w1 = static_cast<int>((wpack >> (4 * hy)) & 7u);
          if (w1 != 0) {
            const float ly = __shfl_sync(0xffffffffu, ly_lane, hy);
            const float2 ly2 = make_float2(ly, ly), hy2 = make_float2(1.f - ly, 1.f - ly);
            auto emit = [&](const float2* up, const float2* dn) {
              uint32_t o[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const float2 v = ffma2(up[k], hy2, fmul2(dn[k], ly2));
                o[k] = pack_bf16(v.x, v.y);
              }
              val = make_uint4(o[0], o[1], o[2], o[3]);
            };
            switch (w1) {
              case 1: emit(H[0], H[1]); break;
              case 2: emit(H[1], H[2]); break;
              case 3: emit(H[2], H[3]); break;
              default: emit(H[3], H[4]); break;
            }
          }
          *reinterpret_cast<uint4*>(ubase + hy * (kH3TC * 128)) = val;
        }
        fence_proxy_async_smem();                              // generic-proxy writes -> visible to the UMMA (async proxy)
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&full[u]);
          mbar_arrive(&xempty[u]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

static bool hc_make_map(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                        const uint32_t* box, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  auto fn = get_tensor_map_encoder();
  if (!fn) return false;
  uint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), dims, strides_bytes, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (head conv, rank %d)", (int)r, rank);
    return false;
  }
  return true;
}

}  // namespace deb

extern "C" int de_head_conv(const void* x, const void* w, int64_t ldw, const float* bias, const float* head_w,
                            float head_b, float* out, int B, int H, int W, int Cin, void* stream) {
  deb::ProfScope prof_("head_conv", 2.0 * B * H * W * 32.0 * 9 * Cin, stream, B, H, W, Cin);
  using namespace deb;
  if (!x || !w || !head_w || !out) return fail(DE_ERR_INVALID, "de_head_conv: null pointer");
  if (B < 1 || H < 1 || W < 1 || Cin < 64 || Cin % 64 != 0) return fail(DE_ERR_UNSUPPORTED, "de_head_conv: Cin=%d must be a multiple of 64", Cin);
  if (ldw < 9LL * Cin || ldw % 8 != 0) return fail(DE_ERR_INVALID, "de_head_conv: ldw=%lld too small / unaligned", (long long)ldw);
  HeadConvParams p{};
  p.B = B; p.H = H; p.W = W;
  p.cin_blocks = Cin / 64;
  p.tiles_x = (W + kHcBW - 1) / kHcBW;
  p.tiles_y = (H + kHcBH - 1) / kHcBH;
  const long long tiles = (long long)B * p.tiles_x * p.tiles_y;
  if (tiles > 0x7fffffffLL) return fail(DE_ERR_UNSUPPORTED, "de_head_conv: too many tiles");
  p.num_tiles = (int)tiles;
  p.bias = bias; p.head_w = head_w; p.head_b = head_b; p.out = out;
  const int w_bytes = 9 * p.cin_blocks * kHcWChunk;
  const int post_bytes = 256 + 2 * kHcN * 4;
  int stages = (232448 - 1024 - post_bytes - w_bytes) / kHcStageBytes;
  if (stages > kHcMaxStages) stages = kHcMaxStages;
  if (stages < 3) return fail(DE_ERR_UNSUPPORTED, "de_head_conv: Cin=%d leaves no room for the activation ring", Cin);
  p.stages = stages;
  const int smem_bytes = 1024 + w_bytes + stages * kHcStageBytes + post_bytes;
  CUtensorMap ta, tw;
  {
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)Cin * 2, (uint64_t)W * Cin * 2, (uint64_t)H * W * Cin * 2};
    const uint32_t box[4] = {64, kHcBW, kHcBoxRows, 1};
    if (!hc_make_map(&ta, x, 4, dims, strides, box)) return DE_ERR_CUDA;
  }
  {
    const uint64_t dims[2] = {(uint64_t)9 * Cin, (uint64_t)kHcN};
    const uint64_t strides[1] = {(uint64_t)ldw * 2};
    const uint32_t box[2] = {64, kHcN};
    if (!hc_make_map(&tw, w, 2, dims, strides, box)) return DE_ERR_CUDA;
  }
  static SmemOptIn optin;
  {
    cudaError_t e = ensure_smem(optin, head_conv_kernel, smem_bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(head_conv)");
  }
  int grid = num_sms();
  if (p.num_tiles < grid) grid = p.num_tiles;
  head_conv_kernel<<<grid, kHcThreads, smem_bytes, reinterpret_cast<cudaStream_t>(stream)>>>(ta, tw, p);
  count_launch();
  return cuda_check_launch("head_conv_kernel");
}

extern "C" int de_head_up_conv(const void* x, const void* w, int64_t ldw, const float* bias, const float* head_w,
                               float head_b, float* out, int B, int Hs, int Ws, int Cin, int align_corners, void* stream) {
  deb::ProfScope prof_("head_up_conv", 2.0 * B * 4.0 * Hs * Ws * 32 * 9 * Cin, stream, B, Hs, Ws, Cin);
  using namespace deb;
  if (!x || !w || !head_w || !out) return fail(DE_ERR_INVALID, "de_head_up_conv: null pointer");
  if (B < 1 || Hs < 1 || Ws < 1 || Cin < 64 || Cin % 64 != 0) return fail(DE_ERR_UNSUPPORTED, "de_head_up_conv: Cin=%d must be a multiple of 64", Cin);
  if (ldw < 9LL * Cin || ldw % 8 != 0) return fail(DE_ERR_INVALID, "de_head_up_conv: ldw=%lld too small / unaligned", (long long)ldw);
  HeadUpParams p{};
  p.x = reinterpret_cast<const __nv_bfloat16*>(x);
  p.B = B; p.Hs = Hs; p.Ws = Ws; p.H = 2 * Hs; p.W = 2 * Ws; p.Cin = Cin;
  p.cin_blocks = Cin / 64;
  // DEB_HEAD_UP_V1=1: the first formulation (three dx copies, N = 32 MMAs) for A/B runs
  static const bool v1 = getenv("DEB_HEAD_UP_V1") && atoi(getenv("DEB_HEAD_UP_V1")) != 0;
  p.tiles_x = v1 ? (p.W + kHcBW - 1) / kHcBW : (p.W + kH3OutC - 1) / kH3OutC;
  p.tiles_y = v1 ? (p.H + kHcBH - 1) / kHcBH : (p.H + kH3TR - 1) / kH3TR;
  const long long tiles = (long long)B * p.tiles_x * p.tiles_y;
  if (tiles > 0x7fffffffLL) return fail(DE_ERR_UNSUPPORTED, "de_head_up_conv: too many tiles");
  p.num_tiles = (int)tiles;
  p.align_corners = align_corners;
  // ATen area_pixel_compute_scale (align_corners): (in - 1) / (out - 1) in float
  p.rh = p.H > 1 ? (float)(Hs - 1) / (float)(p.H - 1) : 0.f;
  p.rw = p.W > 1 ? (float)(Ws - 1) / (float)(p.W - 1) : 0.f;
  p.bias = bias; p.head_w = head_w; p.head_b = head_b; p.out = out;
  const int w_bytes = 9 * p.cin_blocks * kHcWChunk;
  const int post_bytes = 256 + 2 * kHcN * 4;
  const int smem_bytes = v1 ? 1024 + w_bytes + kHuUnits * kHuUnitBytes + kHuSrcSlots * kHuSrcSlotBytes + post_bytes
                            : 1024 + w_bytes + kH3Units * kH3UnitBytes + kH3SrcSlots * kH3SrcSlotBytes + post_bytes;
  if (smem_bytes > 232448) return fail(DE_ERR_UNSUPPORTED, "de_head_up_conv: Cin=%d does not fit shared memory", Cin);
  CUtensorMap tw, tx;
  {
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)Ws, (uint64_t)Hs, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)Cin * 2, (uint64_t)Ws * Cin * 2, (uint64_t)Hs * Ws * Cin * 2};
    const uint32_t box1[4] = {64, kHuSrcCols, kHuSrcRows, 1};
    const uint32_t box3[4] = {64, kH3SrcCols, kH3SrcRows, 1};
    if (!hc_make_map(&tx, x, 4, dims, strides, v1 ? box1 : box3, CU_TENSOR_MAP_SWIZZLE_NONE)) return DE_ERR_CUDA;
  }
  {
    const uint64_t dims[2] = {(uint64_t)9 * Cin, (uint64_t)kHcN};
    const uint64_t strides[1] = {(uint64_t)ldw * 2};
    const uint32_t box[2] = {64, kHcN};
    if (!hc_make_map(&tw, w, 2, dims, strides, box)) return DE_ERR_CUDA;
  }
  int grid = num_sms();
  if (p.num_tiles < grid) grid = p.num_tiles;
  cudaError_t le;
  if (v1) {
    static SmemOptIn optin;
    cudaError_t e = ensure_smem(optin, head_up_conv_kernel, smem_bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(head_up_conv)");
    le = launch_pdl(2, head_up_conv_kernel, dim3(grid), dim3(kHuThreads), (size_t)smem_bytes,
                    reinterpret_cast<cudaStream_t>(stream), tw, tx, p);
  } else {
    static SmemOptIn optin3;
    cudaError_t e = ensure_smem(optin3, head_up3_kernel, smem_bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(head_up3)");
    le = launch_pdl(2, head_up3_kernel, dim3(grid), dim3(kH3Threads), (size_t)smem_bytes,
                    reinterpret_cast<cudaStream_t>(stream), tw, tx, p);
  }
  count_launch();
  if (le != cudaSuccess) return cuda_fail(le, "cudaLaunchKernelEx(head_up_conv_kernel)");
  return cuda_check_launch("head_up_conv_kernel");
}
