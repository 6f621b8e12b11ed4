// Persistent, warp-specialised tcgen05 GEMM / implicit-GEMM 3x3 convolution for sm_100a.
//
//   warp 0      : TMA producer (one elected lane) - A and W tiles into a multi-stage smem ring
//   warp 1      : MMA issuer (lane 0) - tcgen05.mma kind::f16, 128 x BN x 16 per instruction,
//                 fp32 accumulators in TMEM, two accumulator stages so the next tile's main loop
//                 overlaps this tile's epilogue
//   warp 2      : TMEM allocator / deallocator
//   warps 4..11 : epilogue - tcgen05.ld TMEM -> registers, fused bias / activation / residual adds /
//                 row remaps (token offsets, pixel shuffle, conv tiles) / dual store / head dot
//
// Covers K2 patch-embed, K3 ViT linears, K4 reassemble (readout, 1x1, ConvT as
// GEMM + pixel-shuffle), K5 3x3 convs as implicit GEMM (A tiles are TMA boxes of the NHWC activation
// shifted per filter tap, halo zero-filled by TMA), K6 1x1 out_conv, K7 head, K9 1x1 MBConv convs.
#include <cuda.h>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

template <int V>
struct IntTag {
  static constexpr int value = V;
};

#ifdef GEMM_TRACE
// developer build only (scripts/gemm_trace.py): clock stamps of one epilogue warp of CTA 0, 16 per column unit
__device__ long long g_gemm_trace[16384];
#define GT_STAMP(slot) do { if (trace_on) g_gemm_trace[(trace_n * 16 + (slot)) & 16383] = clock64(); } while (0)
#define GT_NEXT() do { if (trace_on) ++trace_n; } while (0)
#else
#define GT_STAMP(slot) do { } while (0)
#define GT_NEXT() do { } while (0)
#endif
constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kEpiWarps = 8;
constexpr int kThreads = 128 + 32 * kEpiWarps;
constexpr int kStgStride = 36;                                  // floats per staged row (conflict-free float4 I/O)
constexpr int kEpiWarpBytes = 32 * kStgStride * 4 + 2 * 32 * 8 + 4 * 32 * 4;   // staging tile + two int64 row tables + bias slice
constexpr int kSmemLimit = 232448;                              // 227 KB opt-in maximum per CTA

struct GemmParams {
  int a_mode, M, N, num_k_blocks, m_tiles, n_tiles;
  int H, W, bw, bh, tiles_x, tiles_y, cin_blocks;
  // epilogue
  const float* bias;
  int bias_mod, act;
  const void* add1;
  int add1_dtype, add1_rowmode, add1_pre;
  long long ld_add1;
  const void* add2;
  int add2_dtype;
  void* out;
  int out_dtype;
  long long ldo;
  __nv_bfloat16* out_relu;
  // general epilogue: storage offset of a row = unit * pitch (unit = storage row, or the pixel of a pixel-shuffle store)
  long long out_pitch, relu_pitch, add2_pitch;
  int out_split, relu_split, split_c;   // split3 stores: [hi | hi | lo], each split_c wide
  int out_rowmode, grp, pad, off;
  int ps_gh, ps_gw, ps_s, ps_cout;
  const float* head_w;
  float head_b;
  int accumulate;  // store epilogue only: TMA reduce-add instead of store
  int k_splits, kb_per_split;  // split-K: work item = (tile, split); split s covers k-blocks [s*kb_per_split, ...)
  int split_rows, split_b;     // split-K partial tiles go to slice `split` of the workspace: row / image offset per slice
  // 3x3 convolution, halo mode: the stage ring is carved into `na` A slots of `a_slot` bytes - each one (bh + 2) x bw
  // pixels of one 64-channel block at ONE horizontal tap offset, used by the three vertical taps as views bw rows apart -
  // and `nb` W-tile slots.  The activation is fetched 3 (bh + 2) / bh times instead of 9.
  int halo, na, nb, a_slot;
};

// EPI 0: "store" epilogue - bias + activation only, output tile staged in 128B-swizzled smem and written
//        with TMA stores (no per-row address math, no LSU store traffic).
// EPI 1: "general" epilogue - residual adds, row remaps, pixel shuffle, dual store, head dot.
constexpr int kStoreBufBytes = BM * 128;                        // one 128-row x 128-byte output unit
// staging buffers per epilogue group.  2 lets a group fill one unit while the TMA engine still reads the previous one,
// but costs a pipeline stage (6 -> 5 at 256-wide tiles): measured 1753 against 1778 frames/s on the DPT_Large step, so 1.
#ifndef STORE_BUFS_PER_GROUP
#define STORE_BUFS_PER_GROUP 1
#endif
template <int BN, int CG, int EPI>
struct Cfg {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBRows = BN / CG;                 // rows of the W tile this CTA stages
  static constexpr int kBBytes = kBRows * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // after the stage ring - EPI 0: two 1024-aligned store buffers, mbarriers (256 B), bias slice (BN floats);
  //                        EPI 1: mbarriers (256 B), per-warp staging areas.
  static constexpr int kStoreBufs = EPI == 0 ? 2 * STORE_BUFS_PER_GROUP : 0;  // per epilogue group: STORE_BUFS_PER_GROUP
  static constexpr int kBarOff = kStoreBufs * kStoreBufBytes;                 // relative to the end of the ring
  static constexpr int kPostBytes = EPI == 0 ? kStoreBufs * kStoreBufBytes + 256 + 1024 : 256 + kEpiWarps * kEpiWarpBytes;
  static constexpr int kStagesRaw = (kSmemLimit - 1024 - kPostBytes) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTmemCols = (2 * BN) < 32 ? 32 : 2 * BN;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + kPostBytes;
};

// Exact-erf GELU (timm nn.GELU) without erf: with a = |x|, Phi(-a) = 0.5*erfc(a/sqrt2) = 2^-q(a), and
//   GELU(x) = x*Phi(x) = relu(x) - a*Phi(-a)        (x>0: x - x*Phi(-x);  x<0: x*Phi(x) = -a*Phi(-a)).
// q(a) - 1 is fitted by a degree-6 polynomial without constant term (weighted minimax on [0,6]; beyond that
// 2^-q underflows smoothly).  fp32 evaluation: max abs error 1.9e-6, max relative error 1.3e-4 wherever
// |GELU| > 1e-6 - two orders below the bf16 rounding of the stored value - in 6 FFMA + 1 MUFU.EX2 + FMNMX +
// FFMA, half the instructions (and half the MUFU traffic) of an Abramowitz-Stegun erf.  The epilogue warps are
// issue-bound on this code when K is short (fc1: K=1024), so the count is what matters.
DEV float gelu_fast(float x) {
  const float a = fabsf(x);
  float p = fmaf(-2.771519575617276e-05f, a, 0.0006936242571100593f);
  p = fmaf(p, a, -0.007728741969913244f);
  p = fmaf(p, a, 0.052738506346940994f);
  p = fmaf(p, a, 0.45929643511772156f);
  p = fmaf(p, a, 1.1510614156723022f);
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fmaf(-p, a, -1.0f)));
  return fmaf(-a, e, fmaxf(x, 0.f));
}
// gelu_fast on two values at once, bit for bit: with n = -|x| the Horner steps alternate sign exactly (k0 a + k1 =
// (-k0) n + k1, then -(s a + k2) = s' n - k2, ...), so the negations of the scalar form disappear and every FMA becomes
// one packed FFMA2 - 13 instead of 20 issue slots per pair in the epilogue that bounds the fc1 GEMM (K = 1024).
DEV float2 gelu_fast2(float2 x) {
  const float2 n = make_float2(-fabsf(x.x), -fabsf(x.y));
  float2 p = ffma2(make_float2(2.771519575617276e-05f, 2.771519575617276e-05f), n, make_float2(0.0006936242571100593f, 0.0006936242571100593f));
  p = ffma2(p, n, make_float2(0.007728741969913244f, 0.007728741969913244f));
  p = ffma2(p, n, make_float2(0.052738506346940994f, 0.052738506346940994f));
  p = ffma2(p, n, make_float2(-0.45929643511772156f, -0.45929643511772156f));
  p = ffma2(p, n, make_float2(1.1510614156723022f, 1.1510614156723022f));
  const float2 t = ffma2(p, n, make_float2(-1.0f, -1.0f));
  float2 e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.x) : "f"(t.x));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.y) : "f"(t.y));
  return ffma2(n, e, make_float2(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f)));
}
// ReLU / ReLU6 inline (two FMNMX per value - cheaper than the call), everything else through the out-of-line copy.
__device__ __noinline__ float4 apply_act4(float4 v, int act);
DEV float4 act4_fast(float4 v, int act) {
  if (act == DE_ACT_RELU || act == DE_ACT_RELU6) {
    const float hi = act == DE_ACT_RELU6 ? 6.f : __int_as_float(0x7f800000);
    v.x = fminf(fmaxf(v.x, 0.f), hi); v.y = fminf(fmaxf(v.y, 0.f), hi);
    v.z = fminf(fmaxf(v.z, 0.f), hi); v.w = fminf(fmaxf(v.w, 0.f), hi);
    return v;
  }
  return apply_act4(v, act);
}
// Kept out of line on purpose: one copy of the activation code per kernel instead of one per use.
__device__ __noinline__ float4 apply_act4(float4 v, int act) {
  if (act == DE_ACT_RELU) {
    v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
  } else if (act == DE_ACT_GELU) {
    v.x = gelu_fast(v.x); v.y = gelu_fast(v.y); v.z = gelu_fast(v.z); v.w = gelu_fast(v.w);
  } else if (act == DE_ACT_RELU6) {
    v.x = fminf(fmaxf(v.x, 0.f), 6.f); v.y = fminf(fmaxf(v.y, 0.f), 6.f);
    v.z = fminf(fmaxf(v.z, 0.f), 6.f); v.w = fminf(fmaxf(v.w, 0.f), 6.f);
  }
  return v;
}

// CG = 1: one CTA per 128 x BN tile.  CG = 2: a CTA pair (cluster of 2, cta_group::2) per 256 x BN tile -
// each CTA stages its own 128 A rows and HALF of the W tile, one tcgen05.mma (M=256) drives both SMs, so
// the L2->smem operand traffic per FLOP drops by a third and the W tile is fetched once per pair.
template <int BN, int CG, int EPI>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  using C = Cfg<BN, CG, EPI>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = smem + C::kStages * C::kABytes;
  uint8_t* post = smem + C::kStages * C::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(post + C::kBarOff);
  uint64_t* tfull = bars;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  uint64_t* full = bars + 5;                     // plain mode: kStages full + kStages empty barriers (<= 16 of the 27 left)
  uint64_t* empty = full + C::kStages;
  // halo mode: A ring (na slots) and W ring (nb slots) with their own barrier pairs; na + nb <= 13
  uint64_t* a_full = bars + 5;
  uint64_t* a_empty = a_full + p.na;
  uint64_t* b_full = a_empty + p.na;
  uint64_t* b_empty = b_full + p.nb;
  uint8_t* sAh = smem;
  uint8_t* sBh = smem + p.na * p.a_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = CG == 2 ? cluster_ctarank() : 0u;
  const int unit = blockIdx.x / CG, num_units = gridDim.x / CG;
  const int num_work = ((p.m_tiles + CG - 1) / CG) * p.n_tiles * p.k_splits;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    if (p.halo) {
      for (int i = 0; i < 2 * (p.na + p.nb); ++i) mbar_init(&a_full[i], 1);
    } else {
      for (int i = 0; i < C::kStages; ++i) {
        mbar_init(&full[i], 1);
        mbar_init(&empty[i], 1);
      }
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], kEpiWarps * CG);
    }
    mbar_fence_init();
  }
  if (warp == 2) {
    if (CG == 2) {
      tmem_alloc_cg2(tmem_slot, C::kTmemCols);
      tmem_relinquish_cg2();
    } else {
      tmem_alloc(tmem_slot, C::kTmemCols);
      tmem_relinquish();
    }
  }
  pdl_launch();                       // the next kernel's prologue may overlap this kernel
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                         // predecessor complete: its outputs may be read, its inputs overwritten

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer (one per CTA)
    if (lane == 0 && p.halo) {
      int sa = 0, sb = 0;
      uint32_t pa = 0, pb = 0;
      const uint32_t a_bytes = p.bw * (p.bh + 2) * BK * 2;
      for (int work = unit; work < num_work; work += num_units) {
        const int m_tile = (work / p.n_tiles) * CG + cta_rank, n_tile = work % p.n_tiles;
        const int per_img = p.tiles_x * p.tiles_y;
        const int b = m_tile / per_img, t = m_tile % per_img;      // b >= B for the padding tile of an odd pair: zero fill
        const int y0 = (t / p.tiles_x) * p.bh, x0 = (t % p.tiles_x) * p.bw;
        const int n_row = n_tile * BN + cta_rank * C::kBRows;
        for (int dx = 0; dx < 3; ++dx) {
          for (int cb = 0; cb < p.cin_blocks; ++cb) {
            mbar_wait(&a_empty[sa], pa ^ 1);
            uint8_t* da = sAh + sa * p.a_slot;
            if (CG == 1) {
              mbar_arrive_expect_tx(&a_full[sa], a_bytes);
              tma_load_4d(da, &tmA, &a_full[sa], cb * BK, x0 + dx - 1, y0 - 1, b);
            } else {
              if (cta_rank == 0) mbar_arrive_expect_tx(&a_full[sa], 2 * a_bytes);
              tma_load_4d_cg2(da, &tmA, mapa_u32(&a_full[sa], 0), cb * BK, x0 + dx - 1, y0 - 1, b);
            }
            if (++sa == p.na) { sa = 0; pa ^= 1; }
            for (int dy = 0; dy < 3; ++dy) {
              mbar_wait(&b_empty[sb], pb ^ 1);
              uint8_t* db = sBh + sb * C::kBBytes;
              const int kcol = ((dy * 3 + dx) * p.cin_blocks + cb) * BK;
              if (CG == 1) {
                mbar_arrive_expect_tx(&b_full[sb], C::kBBytes);
                tma_load_2d(db, &tmB, &b_full[sb], kcol, n_row);
              } else {
                if (cta_rank == 0) mbar_arrive_expect_tx(&b_full[sb], 2 * C::kBBytes);
                tma_load_2d_cg2(db, &tmB, mapa_u32(&b_full[sb], 0), kcol, n_row);
              }
              if (++sb == p.nb) { sb = 0; pb ^= 1; }
            }
          }
        }
      }
    } else if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t a_bytes = p.a_mode == 0 ? C::kABytes : p.bw * p.bh * BK * 2;
      for (int work = unit; work < num_work; work += num_units) {
        const int tile = work / p.k_splits, split = work - tile * p.k_splits;
        const int m_tile = (tile / p.n_tiles) * CG + cta_rank, n_tile = tile % p.n_tiles;
        int b = 0, y0 = 0, x0 = 0;
        if (p.a_mode == 1) {
          const int per_img = p.tiles_x * p.tiles_y;
          b = m_tile / per_img;                      // b >= B for the padding tile of an odd pair: TMA zero-fills
          const int t = m_tile % per_img;
          y0 = (t / p.tiles_x) * p.bh;
          x0 = (t % p.tiles_x) * p.bw;
        }
        const int n_row = n_tile * BN + cta_rank * C::kBRows;
        const int kb0 = split * p.kb_per_split, kb1 = min(kb0 + p.kb_per_split, p.num_k_blocks);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* da = sA + stage * C::kABytes;
          uint8_t* db = sB + stage * C::kBBytes;
          if (CG == 1) {
            mbar_arrive_expect_tx(&full[stage], a_bytes + C::kBBytes);
            if (p.a_mode == 0) {
              tma_load_2d(da, &tmA, &full[stage], kb * BK, m_tile * BM);
            } else {
              const int tap = kb / p.cin_blocks, cb = kb % p.cin_blocks;
              tma_load_4d(da, &tmA, &full[stage], cb * BK, x0 + (tap % 3) - 1, y0 + (tap / 3) - 1, b);
            }
            tma_load_2d(db, &tmB, &full[stage], kb * BK, n_row);
          } else {
            // both CTAs' bytes land on the LEADER's full barrier, which only the leader arms
            if (cta_rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * (a_bytes + C::kBBytes));
            const uint32_t lead_full = mapa_u32(&full[stage], 0);
            if (p.a_mode == 0) {
              tma_load_2d_cg2(da, &tmA, lead_full, kb * BK, m_tile * BM);
            } else {
              const int tap = kb / p.cin_blocks, cb = kb % p.cin_blocks;
              tma_load_4d_cg2(da, &tmA, lead_full, cb * BK, x0 + (tap % 3) - 1, y0 + (tap / 3) - 1, b);
            }
            tma_load_2d_cg2(db, &tmB, lead_full, kb * BK, n_row);
          }
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer (leader CTA only)
    if (cta_rank == 0 && p.halo) {
      constexpr uint32_t idesc = make_idesc_bf16(BM * CG, BN);
      int sa = 0, sb = 0;
      uint32_t pa = 0, pb = 0;
      int as = 0;
      uint32_t aphase = 0;
      const uint32_t dy_step = static_cast<uint32_t>(p.bw * BK * 2) >> 4;       // bw pixel rows further down, in 16-byte units
      for (int work = unit; work < num_work; work += num_units) {
        mbar_wait(&tempty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int dx = 0; dx < 3; ++dx) {
          for (int cb = 0; cb < p.cin_blocks; ++cb) {
            mbar_wait(&a_full[sa], pa);
            tc_fence_after();
            const uint64_t adesc0 = make_sw128_kmajor_desc(smem_u32(sAh + sa * p.a_slot));
            for (int dy = 0; dy < 3; ++dy) {
              mbar_wait(&b_full[sb], pb);
              tc_fence_after();
              const uint64_t adesc = adesc0 + dy * dy_step;
              const uint64_t bdesc = make_sw128_kmajor_desc(smem_u32(sBh + sb * C::kBBytes));
              const bool first = (dx | cb | dy) == 0;
              const bool last_a = dy == 2, last_t = last_a && dx == 2 && cb == p.cin_blocks - 1;
              if (elect_one()) {
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                  if (CG == 2) umma_bf16_cg2(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, !(first && k == 0));
                  else umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, !(first && k == 0));
                }
                if (CG == 2) {
                  umma_commit_cg2(&b_empty[sb], 3);
                  if (last_a) umma_commit_cg2(&a_empty[sa], 3);
                  if (last_t) umma_commit_cg2(&tfull[as], 3);
                } else {
                  umma_commit(&b_empty[sb]);
                  if (last_a) umma_commit(&a_empty[sa]);
                  if (last_t) umma_commit(&tfull[as]);
                }
              }
              __syncwarp();
              if (++sb == p.nb) { sb = 0; pb ^= 1; }
            }
            if (++sa == p.na) { sa = 0; pa ^= 1; }
          }
        }
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
    } else if (cta_rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM * CG, BN);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int work = unit; work < num_work; work += num_units) {
        const int split = work % p.k_splits;
        const int kb0 = split * p.kb_per_split, kb1 = min(kb0 + p.kb_per_split, p.num_k_blocks);
        mbar_wait(&tempty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          // elect_one in a converged warp (not `lane == 0`): ptxas then knows the operands are warp-uniform and
          // emits the UTCHMMAs back to back; under a lane test it wraps every one in an ELECT / BRA.U.ANY
          // uniformisation loop (~50-100 cycles per instruction - measured with scripts/umma_probe.cu).
          const uint64_t adesc = make_sw128_kmajor_desc(smem_u32(sA + stage * C::kABytes));
          const uint64_t bdesc = make_sw128_kmajor_desc(smem_u32(sB + stage * C::kBBytes));
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              if (CG == 2) umma_bf16_cg2(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, ((kb - kb0) | k) != 0);
              else umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, ((kb - kb0) | k) != 0);
            }
            if (CG == 2) {
              umma_commit_cg2(&empty[stage], 3);
              if (kb == kb1 - 1) umma_commit_cg2(&tfull[as], 3);
            } else {
              umma_commit(&empty[stage]);
              if (kb == kb1 - 1) umma_commit(&tfull[as]);
            }
          }
          __syncwarp();
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
    }
  } else if (warp >= 4 && EPI == 0) {
    // ------------------------------------------------------------- store epilogue (bias + activation)
    // Two groups of four warps (one warp per TMEM lane quadrant) alternate over 128-byte-wide column units.
    // Per unit: TMEM -> registers, +bias, activation, pack; once the group's smem buffer has been drained by
    // the previous TMA store, the 128 rows are written with the 128B swizzle and one thread issues the store.
    const int q = warp & 3;
    const int grp = (warp - 4) >> 2;
    const int gtid = q * 32 + lane;                            // thread index inside the group == tile row
    const int bar_id = 1 + grp;
    const bool f32out = p.out_dtype == DE_DTYPE_F32;
    const int ucols = f32out ? 32 : 64;                        // columns per 128-byte unit
    const int units_max = BN / ucols;
    uint8_t* const sbuf0 = post + grp * (STORE_BUFS_PER_GROUP * kStoreBufBytes);
    int sb = 0;                                                // staging buffer of the next unit
    float* s_bias = reinterpret_cast<float*>(post + C::kBarOff + 256);
    const int bias_mod = p.bias_mod > 0 ? p.bias_mod : p.N;
    int as = 0;
    uint32_t aphase = 0;
    bool pending = false;                                      // this group has a store in flight on sbuf
#ifdef GEMM_TRACE
    const bool trace_on = blockIdx.x == 0 && warp == 4 && lane == 0;
    int trace_n = 0;
#endif
    int bias_tile = -1;                                        // n_tile whose bias slice s_bias holds
    for (int work = unit; work < num_work; work += num_units) {
      const int tile = work / p.k_splits, split = work - tile * p.k_splits;
      const int m_tile = (tile / p.n_tiles) * CG + cta_rank, n_tile = tile % p.n_tiles;
      // column units that hold at least one valid column (the last n-tile of a ragged N: the rest is skipped whole)
      const int units = min(units_max, (p.N - n_tile * BN + ucols - 1) / ucols);
      // bias slice of this group's units, fetched while the main loop of the tile is still running - only when the
      // CTA moves to another n-tile (its tiles are num_units apart: with few n-tiles that is rarely)
      const bool new_bias = n_tile != bias_tile;
      if (new_bias) {
        if (bias_tile >= 0) named_bar_sync(bar_id, 128);       // the previous tile's readers are done
        bias_tile = n_tile;
        for (int u = grp; u < units_max; u += 2) {
          if (gtid < ucols) {
            const int n = n_tile * BN + u * ucols + gtid;
            s_bias[u * ucols + gtid] = (p.bias && n < p.N) ? __ldg(p.bias + (n % bias_mod)) : 0.f;
          }
        }
      }
      int b = 0, y0 = 0, x0 = 0;
      if (p.a_mode == 1) {
        const int per_img = p.tiles_x * p.tiles_y;
        b = m_tile / per_img;
        const int t = m_tile % per_img;
        y0 = (t / p.tiles_x) * p.bh;
        x0 = (t % p.tiles_x) * p.bw;
      }
      GT_STAMP(0);
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
      GT_STAMP(1);
      if (new_bias) named_bar_sync(bar_id, 128);               // bias slice visible to the whole group
      const int last_u = grp + ((units - 1 - grp) / 2) * 2;    // last unit this group handles (if any)
      if (grp >= units) {                                      // idle group (narrow tiles): just hand TMEM back
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (CG == 2 && cta_rank != 0) mbar_arrive_cluster(mapa_u32(&tempty[as], 0));
          else mbar_arrive_relaxed(&tempty[as]);
        }
      }
#pragma unroll 1
      for (int u = grp; u < units; u += 2) {
        uint32_t ov[32];
        GT_STAMP(2);
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + u * ucols;
        const float* bs = s_bias + u * ucols;
        // The activation is chosen OUTSIDE the column loops (three straight-line variants): with a run-time test - and a
        // possible call - between the bias loads and their uses, ptxas issued every shared-memory load right before its
        // consumer, and the 16 loads of a unit ran one latency after the other: 1220 cycles of "math" for 64 columns in
        // the clock trace (scripts/gemm_trace.py), more than everything else in the unit together.
        const float act_hi = p.act == DE_ACT_RELU6 ? 6.f : __int_as_float(0x7f800000);
        if (f32out) {
          uint32_t acc[32];
          tmem_ld_32x32(taddr, acc);
          tmem_ld_wait();
          GT_STAMP(3);
          auto math = [&](auto tag) {
            constexpr int A = decltype(tag)::value;              // 0 none, 1 ReLU / ReLU6 (clamp), 2 GELU
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4) {
              const float4 bb = *reinterpret_cast<const float4*>(bs + j4 * 4);
              float v[4] = {__uint_as_float(acc[j4 * 4]) + bb.x, __uint_as_float(acc[j4 * 4 + 1]) + bb.y,
                            __uint_as_float(acc[j4 * 4 + 2]) + bb.z, __uint_as_float(acc[j4 * 4 + 3]) + bb.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                if (A == 1) v[e] = fminf(fmaxf(v[e], 0.f), act_hi);
                if (A == 2) v[e] = gelu_fast(v[e]);
                ov[j4 * 4 + e] = __float_as_uint(v[e]);
              }
            }
          };
          if (p.act == DE_ACT_NONE) math(IntTag<0>{});
          else if (p.act == DE_ACT_GELU) math(IntTag<2>{});
          else math(IntTag<1>{});
        } else {
          uint32_t acc2[2][32];                              // both halves in flight: one TMEM round trip per unit
          tmem_ld_32x32(taddr, acc2[0]);
          tmem_ld_32x32(taddr + 32, acc2[1]);
          tmem_ld_wait();
          GT_STAMP(3);
          auto math = [&](auto tag) {
            constexpr int A = decltype(tag)::value;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              const uint32_t* acc = acc2[hh];
#pragma unroll
              for (int j4 = 0; j4 < 8; ++j4) {
                const float4 bb = *reinterpret_cast<const float4*>(bs + hh * 32 + j4 * 4);
                float2 v01 = fadd2(make_float2(__uint_as_float(acc[j4 * 4]), __uint_as_float(acc[j4 * 4 + 1])), make_float2(bb.x, bb.y));
                float2 v23 = fadd2(make_float2(__uint_as_float(acc[j4 * 4 + 2]), __uint_as_float(acc[j4 * 4 + 3])), make_float2(bb.z, bb.w));
                if (A == 1) {
                  v01 = make_float2(fminf(fmaxf(v01.x, 0.f), act_hi), fminf(fmaxf(v01.y, 0.f), act_hi));
                  v23 = make_float2(fminf(fmaxf(v23.x, 0.f), act_hi), fminf(fmaxf(v23.y, 0.f), act_hi));
                }
                if (A == 2) {                                    // hot (ViT fc1): packed GELU
                  v01 = gelu_fast2(v01);
                  v23 = gelu_fast2(v23);
                }
                ov[hh * 16 + j4 * 2] = pack_bf16(v01.x, v01.y);
                ov[hh * 16 + j4 * 2 + 1] = pack_bf16(v23.x, v23.y);
              }
            }
          };
          if (p.act == DE_ACT_NONE) math(IntTag<0>{});
          else if (p.act == DE_ACT_GELU) math(IntTag<2>{});
          else math(IntTag<1>{});
        }
        if (u == last_u) {                                     // accumulator drained: release it early
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (CG == 2 && cta_rank != 0) mbar_arrive_cluster(mapa_u32(&tempty[as], 0));
            else mbar_arrive_relaxed(&tempty[as]);
          }
        }
        // the store that last used this buffer (STORE_BUFS_PER_GROUP units ago) has finished reading it
        GT_STAMP(4);
        if (pending && gtid == 0) tma_store_wait_read<STORE_BUFS_PER_GROUP - 1>();
        GT_STAMP(5);
        uint8_t* const sbuf = sbuf0 + sb * kStoreBufBytes;
        sb = sb + 1 == STORE_BUFS_PER_GROUP ? 0 : sb + 1;
        named_bar_sync(bar_id, 128);
        GT_STAMP(6);
        uint8_t* rowp = sbuf + gtid * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(rowp + ((j ^ (gtid & 7)) << 4)) =
              make_uint4(ov[j * 4], ov[j * 4 + 1], ov[j * 4 + 2], ov[j * 4 + 3]);
        GT_STAMP(7);
        fence_proxy_async_smem();
        GT_STAMP(8);
        named_bar_sync(bar_id, 128);
        GT_STAMP(9);
        if (gtid == 0) {
          if (m_tile < p.m_tiles) {
            const int col0 = n_tile * BN + u * ucols;
            if (p.k_splits > 1) {                // K-slice: raw partial tile into its own slice of the workspace
              if (p.a_mode == 0) tma_store_2d(&tmC, sbuf, col0, split * p.split_rows + m_tile * BM);
              else tma_store_4d(&tmC, sbuf, col0, x0, y0, split * p.split_b + b);
            } else if (p.accumulate) tma_reduce_add_2d(&tmC, sbuf, col0, m_tile * BM);
            else if (p.a_mode == 0) tma_store_2d(&tmC, sbuf, col0, m_tile * BM);
            else tma_store_4d(&tmC, sbuf, col0, x0, y0, b);
          }
          tma_store_commit();            // one (possibly empty) group per unit: wait_read<N> counts units, not stores
        }
        GT_STAMP(10);
        GT_NEXT();
        pending = true;
      }
      if (++as == 2) {
        as = 0;
        aphase ^= 1;
      }
    }
    if (pending && gtid == 0) tma_store_wait_read<0>();        // smem must outlive the last store's read
  } else if (warp >= 4) {
    // ------------------------------------------------------------- general epilogue
    // TMEM -> registers (thread = accumulator row) -> per-warp smem staging tile -> re-read so that 8
    // lanes cover 32 consecutive columns of one row: every global access (residual read, output write)
    // is a fully used 64/128-byte row segment instead of 32 scattered 16-byte pieces.
    const int q = warp & 3;                      // TMEM lane quadrant this warp may access
    const int half = (warp - 4) >> 2;            // column half
    constexpr int kChunks = BN / 32;
    constexpr int kChunksPerHalf = kChunks >= 2 ? kChunks / 2 : 1;
    const int c_begin = kChunks >= 2 ? half * kChunksPerHalf : 0;
    const int c_end = kChunks >= 2 ? c_begin + kChunksPerHalf : (half == 0 ? 1 : 0);
    const int r = q * 32 + lane;
    const int bias_mod = p.bias_mod > 0 ? p.bias_mod : p.N;
    uint8_t* epi = post + 256 + (warp - 4) * kEpiWarpBytes;
    float* stg = reinterpret_cast<float*>(epi);                               // [32][36] fp32
    long long* s_ro = reinterpret_cast<long long*>(epi + 32 * kStgStride * 4);  // out row offset (elements) or -1
    long long* s_ar = s_ro + 32;                                              // add1 row offset (elements)
    float* s_bias = reinterpret_cast<float*>(s_ar + 32);                      // [kChunksPerHalf][32] bias slice
    const int sub = lane >> 3, cc = (lane & 7) * 4;                           // read-back mapping
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = unit; tile < num_work; tile += num_units) {   // k_splits == 1 on this path (checked on the host)
      const int m_tile = (tile / p.n_tiles) * CG + cta_rank, n_tile = tile % p.n_tiles;
      {
        // row-owner phase: logical row index m of accumulator row r, validity, storage offsets
        long long m;
        bool valid;
        if (p.a_mode == 0) {
          m = static_cast<long long>(m_tile) * BM + r;
          valid = m < p.M;
        } else {
          const int per_img = p.tiles_x * p.tiles_y;
          const int b = m_tile / per_img, t = m_tile % per_img;
          const int y = (t / p.tiles_x) * p.bh + r / p.bw, x = (t % p.tiles_x) * p.bw + r % p.bw;
          valid = (m_tile < p.m_tiles) && (r < p.bw * p.bh) && y < p.H && x < p.W;
          m = (static_cast<long long>(b) * p.H + y) * p.W + x;
        }
        long long orow = m;
        if (p.out_rowmode == 1) orow = m + (m / p.grp) * p.pad + p.off;
        long long arow = orow;
        if (p.add1_rowmode == 1) arow = m / p.grp;
        else if (p.add1_rowmode == 2) arow = (m % p.grp) + p.off;
        long long ro = orow;                       // storage unit: row, or the top-left output pixel of a pixel shuffle
        if (p.out_rowmode == 2) {
          const int per = p.ps_gh * p.ps_gw;
          const int ps_b = static_cast<int>(m / per), t = static_cast<int>(m % per);
          const int ps_y = t / p.ps_gw, ps_x = t % p.ps_gw;
          ro = (static_cast<long long>(ps_b) * p.ps_gh * p.ps_s + ps_y * p.ps_s) * (p.ps_gw * p.ps_s) + ps_x * p.ps_s;
        }
        __syncwarp();
        s_ro[lane] = valid ? ro : -1;
        s_ar[lane] = arow * p.ld_add1;
      }
      // The tile's bias slice is fetched into this warp's smem BEFORE waiting for the accumulator: under
      // saturating TMA traffic an L2 round trip costs thousands of cycles, so it overlaps the main loop.
      if (p.bias) {
        for (int ci = 0; ci < c_end - c_begin; ++ci) {
          const int n = n_tile * BN + (c_begin + ci) * 32 + lane;
          s_bias[ci * 32 + lane] = n < p.N ? __ldg(p.bias + (n % bias_mod)) : 0.f;
        }
      }
      // Residual operands of a chunk are fetched one chunk AHEAD of their use (the first chunk's before the wait for
      // the accumulator): under saturating TMA traffic a global round trip costs thousands of cycles, and four of them
      // in series per tile made the epilogue - not the main loop - the bound of the decoder's 3x3 convolutions
      // (347 us against 237 us for the same shape through the store epilogue).  out may alias add1 element for element
      // (in-place updates): a chunk's operands are read before any store to that chunk, other chunks are other columns.
      // The loaded words stay RAW in registers until the compute phase: converting at the load (bf16 -> fp32) makes
      // every load's first consumer follow it directly, and the 16 loads of a chunk then run one round trip after the
      // other (ncu: 70 % of the epilogue warps' samples on those conversions, ~17k cycles per chunk).
      const bool has_res = p.add1 != nullptr || p.add2 != nullptr;
      const bool a1_f32 = p.add1_dtype == DE_DTYPE_F32, a2_f32 = p.add2_dtype == DE_DTYPE_F32;
      uint4 r1_[8], r2_[8];
      auto fetch_res = [&](int c) {
        const int n0 = n_tile * BN + c * 32;
        const bool col_ok = n0 < p.N && cc < p.N - n0;
        long long colpart2 = n0 + cc;
        if (p.out_rowmode == 2) {
          const int g = n0 / p.ps_cout, co = n0 % p.ps_cout;
          const long long px = static_cast<long long>(g / p.ps_s) * (p.ps_gw * p.ps_s) + g % p.ps_s;
          colpart2 = px * p.add2_pitch + co + cc;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int row = i * 4 + sub;
          const long long ro = col_ok ? s_ro[row] : -1;
          r1_[i] = make_uint4(0u, 0u, 0u, 0u);         // zero bits are 0.0 in both dtypes
          r2_[i] = make_uint4(0u, 0u, 0u, 0u);
          if (p.add1 && ro >= 0) {
            const long long ao = s_ar[row] + n0 + cc;
            if (a1_f32) {
              r1_[i] = *reinterpret_cast<const uint4*>(reinterpret_cast<const float*>(p.add1) + ao);
            } else {
              const uint2 t = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(p.add1) + ao);
              r1_[i].x = t.x;
              r1_[i].y = t.y;
            }
          }
          if (p.add2 && ro >= 0) {                     // same rows / pitch as a plain `out`
            const long long o2 = ro * p.add2_pitch + colpart2;
            if (a2_f32) {
              r2_[i] = *reinterpret_cast<const uint4*>(reinterpret_cast<const float*>(p.add2) + o2);
            } else {
              const uint2 t = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(p.add2) + o2);
              r2_[i].x = t.x;
              r2_[i].y = t.y;
            }
          }
        }
      };
      __syncwarp();                                // row tables visible to the whole warp
      if (has_res && c_begin < c_end) fetch_res(c_begin);
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
#pragma unroll 1
      for (int c = c_begin; c < c_end; ++c) {
        uint32_t acc[32];
        __syncwarp();                            // previous chunk's read-back finished
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + c * 32, acc);
        tmem_ld_wait();
        if (c == c_end - 1) {                    // accumulator fully drained into registers: release it early
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (CG == 2 && cta_rank != 0) mbar_arrive_cluster(mapa_u32(&tempty[as], 0));
            else mbar_arrive_relaxed(&tempty[as]);
          }
        }
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4)
          *reinterpret_cast<float4*>(stg + lane * kStgStride + j4 * 4) =
              make_float4(__uint_as_float(acc[j4 * 4]), __uint_as_float(acc[j4 * 4 + 1]),
                          __uint_as_float(acc[j4 * 4 + 2]), __uint_as_float(acc[j4 * 4 + 3]));
        __syncwarp();
        const int n0 = n_tile * BN + c * 32;
        const bool col_ok = n0 < p.N && cc < p.N - n0;       // N % 8 == 0 and cc % 4 == 0: whole float4 in or out
        long long colpart = n0 + cc, colpart_r = n0 + cc;
        if (p.out_rowmode == 2 && n0 < p.N) {
          const int g = n0 / p.ps_cout, co = n0 % p.ps_cout;
          const long long px = static_cast<long long>(g / p.ps_s) * (p.ps_gw * p.ps_s) + g % p.ps_s;
          colpart = px * p.out_pitch + co + cc;
          colpart_r = px * p.relu_pitch + co + cc;
        }
        float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f), hw4 = bias4;
        if (p.bias) bias4 = *reinterpret_cast<const float4*>(s_bias + (c - c_begin) * 32 + cc);
        if (p.head_w && col_ok) hw4 = __ldg(reinterpret_cast<const float4*>(p.head_w + n0 + cc));
        // compute phase: every row's value into registers (consumes this chunk's residual operands) ...  The activation is
        // chosen outside the row loop (straight-line variants, as in the store epilogue): an out-of-line call between the
        // staging loads and their uses serialises them.
        float4 v_[8];
        const float act_hi = p.act == DE_ACT_RELU6 ? 6.f : __int_as_float(0x7f800000);
        auto compute = [&](auto tag) {
          constexpr int A = decltype(tag)::value;                // 0 none, 1 ReLU / ReLU6 (clamp), 2 GELU
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int row = i * 4 + sub;
            float4 v = *reinterpret_cast<const float4*>(stg + row * kStgStride + cc);
            v.x += bias4.x; v.y += bias4.y; v.z += bias4.z; v.w += bias4.w;
            float4 a1 = make_float4(0.f, 0.f, 0.f, 0.f), a2 = a1;
            if (has_res) {
              const uint4 w1 = r1_[i], w2 = r2_[i];
              a1 = a1_f32 ? make_float4(__uint_as_float(w1.x), __uint_as_float(w1.y), __uint_as_float(w1.z), __uint_as_float(w1.w))
                          : make_float4(bf16_lo(w1.x), bf16_hi(w1.x), bf16_lo(w1.y), bf16_hi(w1.y));
              a2 = a2_f32 ? make_float4(__uint_as_float(w2.x), __uint_as_float(w2.y), __uint_as_float(w2.z), __uint_as_float(w2.w))
                          : make_float4(bf16_lo(w2.x), bf16_hi(w2.x), bf16_lo(w2.y), bf16_hi(w2.y));
              if (p.add1_pre) { v.x += a1.x; v.y += a1.y; v.z += a1.z; v.w += a1.w; }
            }
            if (A == 1) {
              v.x = fminf(fmaxf(v.x, 0.f), act_hi); v.y = fminf(fmaxf(v.y, 0.f), act_hi);
              v.z = fminf(fmaxf(v.z, 0.f), act_hi); v.w = fminf(fmaxf(v.w, 0.f), act_hi);
            }
            if (A == 2) { v.x = gelu_fast(v.x); v.y = gelu_fast(v.y); v.z = gelu_fast(v.z); v.w = gelu_fast(v.w); }
            if (has_res) {
              if (!p.add1_pre) { v.x += a1.x; v.y += a1.y; v.z += a1.z; v.w += a1.w; }
              v.x += a2.x; v.y += a2.y; v.z += a2.z; v.w += a2.w;
            }
            v_[i] = v;
          }
        };
        if (p.act == DE_ACT_NONE) compute(IntTag<0>{});
        else if (p.act == DE_ACT_GELU) compute(IntTag<2>{});
        else compute(IntTag<1>{});
        // ... then the NEXT chunk's operand loads go out, and only then this chunk's stores
        if (has_res && c + 1 < c_end) fetch_res(c + 1);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int row = i * 4 + sub;
          const float4 v = v_[i];
          if (p.head_w) {
            float d = v.x * hw4.x + v.y * hw4.y + v.z * hw4.z + v.w * hw4.w;
            d += __shfl_xor_sync(0xffffffffu, d, 1);
            d += __shfl_xor_sync(0xffffffffu, d, 2);
            d += __shfl_xor_sync(0xffffffffu, d, 4);
            const long long hro = s_ro[row];
            if ((lane & 7) == 0 && hro >= 0) reinterpret_cast<float*>(p.out)[hro] = fmaxf(d + p.head_b, 0.f);
            continue;
          }
          const long long ro = col_ok ? s_ro[row] : -1;
          if (ro < 0) continue;
          const long long o = ro * p.out_pitch + colpart;
          if (p.out_dtype == DE_DTYPE_F32) {
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + o) = v;
          } else {
            const uint2 hi = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
            __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(p.out) + o;
            *reinterpret_cast<uint2*>(op) = hi;
            if (p.out_split) {                       // [hi | hi | lo]: the bf16x3 operand layout
              *reinterpret_cast<uint2*>(op + p.split_c) = hi;
              *reinterpret_cast<uint2*>(op + 2 * p.split_c) =
                  make_uint2(pack_bf16(v.x - bf16_lo(hi.x), v.y - bf16_hi(hi.x)), pack_bf16(v.z - bf16_lo(hi.y), v.w - bf16_hi(hi.y)));
            }
          }
          if (p.out_relu) {
            const float4 rv = make_float4(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f), fmaxf(v.z, 0.f), fmaxf(v.w, 0.f));
            const uint2 hi = make_uint2(pack_bf16(rv.x, rv.y), pack_bf16(rv.z, rv.w));
            __nv_bfloat16* rp = p.out_relu + ro * p.relu_pitch + colpart_r;
            *reinterpret_cast<uint2*>(rp) = hi;
            if (p.relu_split) {
              *reinterpret_cast<uint2*>(rp + p.split_c) = hi;
              *reinterpret_cast<uint2*>(rp + 2 * p.split_c) =
                  make_uint2(pack_bf16(rv.x - bf16_lo(hi.x), rv.y - bf16_hi(hi.x)), pack_bf16(rv.z - bf16_lo(hi.y), rv.w - bf16_hi(hi.y)));
            }
          }
        }
      }
      if (c_end <= c_begin) {                    // idle column half (BN == 32): still hand the accumulator back
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (CG == 2 && cta_rank != 0) mbar_arrive_cluster(mapa_u32(&tempty[as], 0));
          else mbar_arrive_relaxed(&tempty[as]);
        }
      }
      if (++as == 2) {
        as = 0;
        aphase ^= 1;
      }
    }
  }

  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if (CG == 2) tmem_dealloc_cg2(tmem_base, C::kTmemCols);
    else tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static bool make_map(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const uint32_t* box, CUtensorMapDataType dtype = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16) {
  auto fn = get_tensor_map_encoder();
  if (!fn) return false;
  uint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(map, dtype, rank, const_cast<void*>(base), dims, strides_bytes, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu %llu, box %u %u)", (int)r, rank,
              (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1]);
    return false;
  }
  return true;
}

// pick the (bw, bh) spatial box with bw*bh <= 128 that wastes the fewest accumulator rows
static void pick_conv_box(int H, int W, int* bw_out, int* bh_out) {
  double best = -1;
  int bbw = 1, bbh = 1;
  for (int bw = 1; bw <= 128 && bw <= W; ++bw) {
    int bh = 128 / bw;
    if (bh > H) bh = H;
    if (bh < 1) continue;
    const long long tiles = (long long)((W + bw - 1) / bw) * ((H + bh - 1) / bh);
    const double eff = (double)H * W / (double)(tiles * 128);
    if (eff > best + 1e-9) {
      best = eff;
      bbw = bw;
      bbh = bh;
    }
  }
  *bw_out = bbw;
  *bh_out = bbh;
}

// Halo mode of the 3x3 convolution: box with bw % 8 == 0 (the three vertical-tap views of a slot must start on 1024-byte
// swizzle atoms) and a (bh + 2) x bw slot of at most 32 KB.  Returns false when no such box keeps the tile efficiency.
static bool pick_conv_box_halo(int H, int W, int plain_bw, int plain_bh, int* bw_out, int* bh_out) {
  const long long plain_tiles = (long long)((W + plain_bw - 1) / plain_bw) * ((H + plain_bh - 1) / plain_bh);
  double best = -1, best_over = 1e9;
  int bbw = 0, bbh = 0;
  for (int bw = 8; bw <= 128; bw += 8) {
    int bh = 128 / bw;
    if (bh > H) bh = H;
    if (bh < 1 || (bh + 2) * bw > 256) continue;
    const long long tiles = (long long)((W + bw - 1) / bw) * ((H + bh - 1) / bh);
    const double eff = (double)H * W / (double)(tiles * 128);
    const double over = (double)(bh + 2) / bh;                // activation bytes fetched per byte used
    if (eff > best + 1e-9 || (eff > best - 1e-9 && over < best_over)) {
      best = eff;
      best_over = over;
      bbw = bw;
      bbh = bh;
    }
  }
  if (bbw == 0) return false;
  const long long tiles = (long long)((W + bbw - 1) / bbw) * ((H + bbh - 1) / bbh);
  if (tiles * 10 > plain_tiles * 11) return false;            // > 10 % more tiles than the free-form box: not worth it
  *bw_out = bbw;
  *bh_out = bbh;
  return true;
}
// bytes of the stage ring of gemm_kernel<bn, cg, epi> (mirrors Cfg)
static int ring_bytes(int bn, int cg, int epi) {
  const int stage = BM * BK * 2 + (bn / cg) * BK * 2;
  const int post = epi == 0 ? 2 * STORE_BUFS_PER_GROUP * kStoreBufBytes + 256 + 1024 : 256 + kEpiWarps * kEpiWarpBytes;
  int stages = (kSmemLimit - 1024 - post) / stage;
  if (stages > 8) stages = 8;
  return stages * stage;
}

template <int BN, int CG, int EPI>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const GemmParams& p,
                  cudaStream_t st) {
  using C = Cfg<BN, CG, EPI>;
  static_assert(C::kStages >= 2, "smem budget leaves too few pipeline stages");
  static SmemOptIn optin;
  {
    cudaError_t e = ensure_smem(optin, gemm_kernel<BN, CG, EPI>, (int)C::kSmemBytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(gemm)");
  }
  const int tiles = ((p.m_tiles + CG - 1) / CG) * p.n_tiles * p.k_splits;
  int units = num_sms() / CG;
  if (tiles < units) units = tiles;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(units * CG);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled(1) ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_kernel<BN, CG, EPI>, ta, tb, tc, p);
  count_launch();
  if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(gemm_kernel)");
  return cuda_check_launch("gemm_kernel");
}

// ------------------------------------------------------------------------------------------------ split-K finalize
// Second half of a split-K call: every K-slice has stored its fp32 partial tiles into its own slice of the workspace
// `acc` [splits][rows_per_slice][N]; this kernel adds the slices in a FIXED order (bitwise reproducible - a reduce-add
// of the slices into one buffer would not be) and applies the call's epilogue (the contract of DeGemm: bias, add1
// before / after the activation, add2, row maps, pixel shuffle, split3 stores, ReLU copy; accumulate: += into the
// fp32 output).  One thread per (row, 4 columns).
struct FinalizeParams {
  const float* acc;
  int splits;
  long long slice_stride;   // elements between slices
  int accumulate;
  int M, N;
  const float* bias;
  int bias_mod, act;
  const void* add1;
  int add1_dtype, add1_rowmode, add1_pre;
  long long ld_add1;
  const void* add2;
  int add2_dtype;
  void* out;
  int out_dtype;
  __nv_bfloat16* out_relu;
  long long out_pitch, relu_pitch, add2_pitch;
  int out_split, relu_split, split_c;
  int out_rowmode, grp, pad, off;
  int ps_gh, ps_gw, ps_s, ps_cout;
};
__global__ void __launch_bounds__(256) splitk_finalize_kernel(const FinalizeParams p) {
  const int n4 = p.N >> 2;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)p.M * n4) return;
  const long long m = idx / n4;
  const int n = static_cast<int>(idx - m * n4) * 4;
  const float* ap = p.acc + m * p.N + n;
  float4 v = *reinterpret_cast<const float4*>(ap);
  for (int sl = 1; sl < p.splits; ++sl) {
    const float4 t = *reinterpret_cast<const float4*>(ap + sl * p.slice_stride);
    v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
  }
  if (p.bias) {
    const int bm = p.bias_mod > 0 ? p.bias_mod : p.N;
    v.x += __ldg(p.bias + (n % bm)); v.y += __ldg(p.bias + ((n + 1) % bm));
    v.z += __ldg(p.bias + ((n + 2) % bm)); v.w += __ldg(p.bias + ((n + 3) % bm));
  }
  long long unit = m;
  if (p.out_rowmode == 1) unit = m + (m / p.grp) * p.pad + p.off;
  long long colpart = n, colpart_r = n, colpart2 = n;
  if (p.out_rowmode == 2) {
    const int per = p.ps_gh * p.ps_gw;
    const int ps_b = static_cast<int>(m / per), t = static_cast<int>(m % per);
    const int ps_y = t / p.ps_gw, ps_x = t % p.ps_gw;
    unit = (static_cast<long long>(ps_b) * p.ps_gh * p.ps_s + ps_y * p.ps_s) * (p.ps_gw * p.ps_s) + ps_x * p.ps_s;
    const int g = n / p.ps_cout, co = n % p.ps_cout;
    const long long px = static_cast<long long>(g / p.ps_s) * (p.ps_gw * p.ps_s) + g % p.ps_s;
    colpart = px * p.out_pitch + co;
    colpart_r = px * p.relu_pitch + co;
    colpart2 = px * p.add2_pitch + co;
  }
  float4 a1 = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.add1) {
    long long arow = unit;
    if (p.add1_rowmode == 1) arow = m / p.grp;
    else if (p.add1_rowmode == 2) arow = (m % p.grp) + p.off;
    const long long ao = arow * p.ld_add1 + n;
    if (p.add1_dtype == DE_DTYPE_F32) {
      a1 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.add1) + ao);
    } else {
      const uint2 t = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(p.add1) + ao);
      a1 = make_float4(bf16_lo(t.x), bf16_hi(t.x), bf16_lo(t.y), bf16_hi(t.y));
    }
  }
  if (p.add1_pre) { v.x += a1.x; v.y += a1.y; v.z += a1.z; v.w += a1.w; }
  if (p.act != DE_ACT_NONE) v = apply_act4(v, p.act);
  if (!p.add1_pre) { v.x += a1.x; v.y += a1.y; v.z += a1.z; v.w += a1.w; }
  if (p.add2) {
    const long long o2 = unit * p.add2_pitch + colpart2;
    if (p.add2_dtype == DE_DTYPE_F32) {
      const float4 t = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.add2) + o2);
      v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
    } else {
      const uint2 t = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(p.add2) + o2);
      v.x += bf16_lo(t.x); v.y += bf16_hi(t.x); v.z += bf16_lo(t.y); v.w += bf16_hi(t.y);
    }
  }
  const long long o = unit * p.out_pitch + colpart;
  if (p.out_dtype == DE_DTYPE_F32) {
    float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + o);
    if (p.accumulate) {
      const float4 old = *op;
      v.x += old.x; v.y += old.y; v.z += old.z; v.w += old.w;
    }
    *op = v;
  } else {
    const uint2 hi = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
    __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(p.out) + o;
    *reinterpret_cast<uint2*>(op) = hi;
    if (p.out_split) {
      *reinterpret_cast<uint2*>(op + p.split_c) = hi;
      *reinterpret_cast<uint2*>(op + 2 * p.split_c) =
          make_uint2(pack_bf16(v.x - bf16_lo(hi.x), v.y - bf16_hi(hi.x)), pack_bf16(v.z - bf16_lo(hi.y), v.w - bf16_hi(hi.y)));
    }
  }
  if (p.out_relu) {
    const float4 rv = make_float4(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f), fmaxf(v.z, 0.f), fmaxf(v.w, 0.f));
    const uint2 hi = make_uint2(pack_bf16(rv.x, rv.y), pack_bf16(rv.z, rv.w));
    __nv_bfloat16* rp = p.out_relu + unit * p.relu_pitch + colpart_r;
    *reinterpret_cast<uint2*>(rp) = hi;
    if (p.relu_split) {
      *reinterpret_cast<uint2*>(rp + p.split_c) = hi;
      *reinterpret_cast<uint2*>(rp + 2 * p.split_c) =
          make_uint2(pack_bf16(rv.x - bf16_lo(hi.x), rv.y - bf16_hi(hi.x)), pack_bf16(rv.z - bf16_lo(hi.y), rv.w - bf16_hi(hi.y)));
    }
  }
}

}  // namespace deb

#ifdef GEMM_TRACE
extern "C" int de_gemm_trace_read(long long* host, int n) {
  return cudaMemcpyFromSymbol(host, deb::g_gemm_trace, sizeof(long long) * (size_t)n) == cudaSuccess ? 0 : -2;
}
#endif
extern "C" int de_gemm(const DeGemm* g, void* stream) {
  deb::ProfScope prof_("gemm", g ? (g->a_mode == 1 ? 2.0 * g->B * g->H * g->W * (double)g->N * 9 * g->Cin : 2.0 * g->M * (double)g->N * g->K) : 0.0, stream, g ? (g->a_mode == 1 ? (int64_t)g->B * g->H * g->W : g->M) : 0, g ? g->N : 0, g ? (g->a_mode == 1 ? 9 * g->Cin : g->K) : 0, g ? g->a_mode * 100 + g->act * 10 + (g->accumulate ? 1 : 0) : 0);
  using namespace deb;
  if (!g || !g->a || !g->w || !g->out) return fail(DE_ERR_INVALID, "de_gemm: null pointer");
  if (g->N % 8 != 0) return fail(DE_ERR_INVALID, "de_gemm: N=%d must be a multiple of 8", g->N);
  GemmParams p{};
  p.a_mode = g->a_mode;
  p.N = g->N;
  int bn = g->block_n;
  if (bn == 0) bn = g->N >= 256 ? 256 : (g->N > 64 ? 128 : (g->N > 32 ? 64 : 32));
  if (g->head_w && (g->N != 32 || bn != 32)) return fail(DE_ERR_INVALID, "de_gemm: head_w needs N == 32");
  if (g->out_rowmode == 2 && (g->ps_cout % 32 != 0)) return fail(DE_ERR_INVALID, "de_gemm: ps_cout %% 32 != 0");
  int cg = g->cta_group;
  CUtensorMap ta, tb;
  int64_t ktot;
  if (g->a_mode == 0) {
    if (g->K % 8 != 0 || g->lda % 8 != 0) return fail(DE_ERR_INVALID, "de_gemm: K and lda must be multiples of 8");
    p.M = g->M;
    p.m_tiles = (g->M + BM - 1) / BM;
    p.num_k_blocks = (g->K + BK - 1) / BK;
    ktot = g->K;
    const uint64_t dims[2] = {(uint64_t)g->K, (uint64_t)g->M};
    const uint64_t strides[1] = {(uint64_t)g->lda * 2};
    const uint32_t box[2] = {BK, BM};
    if (!make_map(&ta, g->a, 2, dims, strides, box)) return DE_ERR_CUDA;
  } else if (g->a_mode == 1) {
    if (g->Cin % 8 != 0) return fail(DE_ERR_INVALID, "de_gemm: Cin must be a multiple of 8");
    p.H = g->H;
    p.W = g->W;
    pick_conv_box(g->H, g->W, &p.bw, &p.bh);
    p.tiles_x = (g->W + p.bw - 1) / p.bw;
    p.tiles_y = (g->H + p.bh - 1) / p.bh;
    p.m_tiles = g->B * p.tiles_x * p.tiles_y;
    p.cin_blocks = (g->Cin + BK - 1) / BK;
    p.num_k_blocks = 9 * p.cin_blocks;
    p.M = g->B * g->H * g->W;
    ktot = (int64_t)9 * p.cin_blocks * BK;
    const uint64_t dims[4] = {(uint64_t)g->Cin, (uint64_t)g->W, (uint64_t)g->H, (uint64_t)g->B};
    const uint64_t strides[3] = {(uint64_t)g->Cin * 2, (uint64_t)g->W * g->Cin * 2, (uint64_t)g->H * g->W * g->Cin * 2};
    const uint32_t box[4] = {BK, (uint32_t)p.bw, (uint32_t)p.bh, 1};
    if (!make_map(&ta, g->a, 4, dims, strides, box)) return DE_ERR_CUDA;
  } else {
    return fail(DE_ERR_INVALID, "de_gemm: bad a_mode %d", g->a_mode);
  }
  // 128 < N < 256 with many row tiles (MiDaS_small's 1x1 expand convs at high resolution, short K): one 256-wide tile
  // per row block instead of a full and a mostly empty 128-wide one - the store epilogue skips all-empty column units,
  // and with a one-k-block main loop the per-tile epilogue is what the kernel costs
  if (g->block_n == 0 && g->N > 128 && g->N < 256 && p.m_tiles >= num_sms() && p.num_k_blocks <= 4) bn = 256;
  // N <= 32 with a plain "bias + activation -> dense output" epilogue (MiDaS_small's depthwise-separable pointwise conv
  // 32 -> 24 and the first projection of a stage): a 64-wide tile runs the TMA-store epilogue - the tensor map clips the
  // columns beyond N - where the 32-wide tile would take the general epilogue with scattered 8-byte stores
  // (scripts/narrow_n_probe.py: M 262144, K 32, N 24: 34.6 -> 21.5 us; equal or better on every shape measured)
  if (g->block_n == 0 && bn == 32 && !g->add1 && !g->add2 && !g->out_relu && !g->head_w && g->out_rowmode == 0 && !g->out_split &&
      g->epilogue != 1 && (g->ldo * (g->out_dtype == DE_DTYPE_F32 ? 4 : 2)) % 16 == 0 && (reinterpret_cast<uintptr_t>(g->out) & 15) == 0 &&
      (g->a_mode == 0 || g->ldo == g->N))
    bn = 64;
  // Split-K for small grids (the batch-1 latency path, the coarse decoder levels): when the tiles cover a fraction of the
  // SMs and K is long, K-slices of a tile run on different SMs and store their fp32 partial tiles into per-slice
  // regions of the caller's workspace; splitk_finalize_kernel adds the slices in a fixed order and applies the epilogue.
  int k_splits = 1;
  if (g->k_splits != 1 && !g->head_w && g->epilogue != 1 && g->splitk_ws && g->N % 4 == 0 &&
      (reinterpret_cast<uintptr_t>(g->splitk_ws) & 15) == 0 && (!g->accumulate || (g->act == DE_ACT_NONE && g->out_dtype == DE_DTYPE_F32))) {
    const int bn_s = g->block_n ? bn : (g->N >= 256 ? 256 : bn);
    const int tiles = p.m_tiles * ((g->N + bn_s - 1) / bn_s);
    const long long slice_bytes = (g->a_mode == 0 ? (long long)p.m_tiles * BM : (long long)p.M) * g->N * 4;
    int want = g->k_splits > 1 ? g->k_splits : num_sms() / (tiles > 0 ? tiles : 1);
    if (want > p.num_k_blocks / 2) want = p.num_k_blocks / 2;         // at least two k-blocks per slice
    if (want > 16) want = 16;
    if (slice_bytes > 0 && want > g->splitk_ws_bytes / slice_bytes) want = (int)(g->splitk_ws_bytes / slice_bytes);
    // the finalize launch must pay for itself: a third of the SMs busy and a medium K, or half of them and a long K
    // (DPT_Hybrid batch 1, bf16x3 decoder convolutions at 96x96: 72 tiles, 108 k-blocks)
    static const int k2 = getenv("DEB_SPLITK_HALF") ? atoi(getenv("DEB_SPLITK_HALF")) : 1;
    const bool worth = (tiles * 3 <= num_sms() && p.num_k_blocks >= 24) || (k2 && tiles * 2 <= num_sms() && p.num_k_blocks >= 64);
    if (want >= 2 && bn_s >= 64 && (g->k_splits > 1 || worth)) {
      k_splits = want;
      bn = bn_s;                                                      // wide tiles + K-slices instead of narrow tiles
    }
  }
  const bool split_ws = k_splits > 1;
  // small problems (the batch-1 latency path: 5 row tiles at T=577): halve the tile width until the grid covers at
  // least half of the SMs - a 45-CTA launch of 256-wide tiles leaves 100 SMs idle and every CTA streams 2x the weights
  if (g->block_n == 0 && !g->head_w && k_splits == 1) {
    const int half_sms = num_sms() / 2;
    // 32-wide tiles exist only with the general epilogue: go that far only when it is what runs anyway
    const bool general = g->add1 || g->add2 || g->out_relu || g->out_rowmode != 0 || g->out_split || g->epilogue == 1;
    const int min_bn = general ? 32 : 64;
    while (bn > min_bn && p.m_tiles * ((g->N + bn - 1) / bn) < half_sms) bn >>= 1;
  }
  if (g->ldw % 8 != 0 || g->ldw < ktot) return fail(DE_ERR_INVALID, "de_gemm: ldw=%lld too small / unaligned for K=%lld", (long long)g->ldw, (long long)ktot);
  {
    const uint64_t dims[2] = {(uint64_t)ktot, (uint64_t)g->N};
    const uint64_t strides[1] = {(uint64_t)g->ldw * 2};
    // pair mode: N a multiple of 256 and enough 256-row tile pairs to occupy the 74 SM pairs
    if (cg == 0) cg = (k_splits == 1 && (bn == 256 || bn == 128) && g->N % bn == 0 && ((p.m_tiles + 1) / 2) * (g->N / bn) >= 64) ? 2 : 1;
    if (cg == 2 && ((bn != 256 && bn != 128) || g->N % bn != 0))
      return fail(DE_ERR_INVALID, "de_gemm: cta_group 2 needs block_n 128 / 256 and N %% block_n == 0");
    const uint32_t box[2] = {BK, (uint32_t)(bn / cg)};
    if (!make_map(&tb, g->w, 2, dims, strides, box)) return DE_ERR_CUDA;
  }
  p.n_tiles = (g->N + bn - 1) / bn;
  p.bias = g->bias;
  p.bias_mod = g->bias_mod;
  p.act = g->act;
  p.add1 = g->add1;
  p.add1_dtype = g->add1_dtype;
  p.add1_rowmode = g->add1_rowmode;
  p.add1_pre = g->add1_pre;
  p.ld_add1 = g->ld_add1;
  p.add2 = g->add2;
  p.add2_dtype = g->add2_dtype;
  if (g->add2 && g->add2_dtype == DE_DTYPE_F32 && g->add1_pre) return fail(DE_ERR_INVALID, "de_gemm: fp32 add2 with add1_pre is not supported");
  p.out = g->out;
  p.out_dtype = g->out_dtype;
  p.ldo = g->ldo;
  p.out_relu = reinterpret_cast<__nv_bfloat16*>(g->out_relu);
  // split3 stores ([hi | hi | lo] rows, the bf16x3 operand layout): out and / or out_relu
  p.out_split = g->out_split ? 1 : 0;
  p.relu_split = g->relu_split ? 1 : 0;
  p.split_c = g->out_rowmode == 2 ? g->ps_cout : g->N;
  if (p.out_split && (g->out_dtype != DE_DTYPE_BF16 || g->head_w || g->accumulate || (g->out_rowmode != 2 && g->ldo < 3LL * g->N)))
    return fail(DE_ERR_INVALID, "de_gemm: out_split needs a bf16 output with ldo >= 3N");
  if (p.relu_split && !g->out_relu) return fail(DE_ERR_INVALID, "de_gemm: relu_split without out_relu");
  {
    const long long ldr = g->ld_relu > 0 ? g->ld_relu : g->ldo;
    if (g->out_relu && g->out_rowmode != 2 && ldr < (p.relu_split ? 3LL : 1LL) * g->N) return fail(DE_ERR_INVALID, "de_gemm: ld_relu too small");
    if (g->out_rowmode == 2) {
      p.out_pitch = (long long)(g->ps_cout > 0 ? g->ps_cout : 1) * (p.out_split ? 3 : 1);
      p.relu_pitch = (long long)(g->ps_cout > 0 ? g->ps_cout : 1) * (p.relu_split ? 3 : 1);
    } else {
      p.out_pitch = g->head_w ? 1 : g->ldo;
      p.relu_pitch = ldr;
    }
    p.add2_pitch = g->out_rowmode == 2 ? (long long)(g->ps_cout > 0 ? g->ps_cout : 1) : (g->ld_add2 > 0 ? g->ld_add2 : g->ldo);
  }
  p.out_rowmode = g->out_rowmode;
  p.grp = g->grp > 0 ? g->grp : 1;
  p.pad = g->pad;
  p.off = g->off;
  p.ps_gh = g->ps_gh; p.ps_gw = g->ps_gw; p.ps_s = g->ps_s; p.ps_cout = g->ps_cout > 0 ? g->ps_cout : 1;
  p.head_w = g->head_w;
  p.head_b = g->head_b;
  p.accumulate = g->accumulate;
  p.k_splits = k_splits;
  p.kb_per_split = (p.num_k_blocks + k_splits - 1) / k_splits;
  p.k_splits = (p.num_k_blocks + p.kb_per_split - 1) / p.kb_per_split;   // no empty slice
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (split_ws) {
    // phase 1 stores raw partial tiles into workspace slices; the epilogue proper runs in splitk_finalize_kernel
    const long long slice_rows = g->a_mode == 0 ? (long long)p.m_tiles * BM : (long long)p.M;
    FinalizeParams f{};
    f.acc = reinterpret_cast<const float*>(g->splitk_ws);
    f.splits = p.k_splits;
    f.slice_stride = slice_rows * g->N;
    f.accumulate = g->accumulate;
    f.M = p.M; f.N = g->N;
    f.bias = p.bias; f.bias_mod = p.bias_mod; f.act = p.act;
    f.add1 = p.add1; f.add1_dtype = p.add1_dtype; f.add1_rowmode = p.add1_rowmode; f.add1_pre = p.add1_pre; f.ld_add1 = p.ld_add1;
    f.add2 = p.add2; f.add2_dtype = p.add2_dtype;
    f.out = p.out; f.out_dtype = p.out_dtype; f.out_relu = p.out_relu;
    f.out_pitch = p.out_pitch; f.relu_pitch = p.relu_pitch; f.add2_pitch = p.add2_pitch;
    f.out_split = p.out_split; f.relu_split = p.relu_split; f.split_c = p.split_c;
    f.out_rowmode = p.out_rowmode; f.grp = p.grp; f.pad = p.pad; f.off = p.off;
    f.ps_gh = p.ps_gh; f.ps_gw = p.ps_gw; f.ps_s = p.ps_s; f.ps_cout = p.ps_cout;
    p.bias = nullptr; p.act = DE_ACT_NONE; p.add1 = nullptr; p.add2 = nullptr; p.out_relu = nullptr;
    p.out = g->splitk_ws; p.out_dtype = DE_DTYPE_F32; p.ldo = g->N; p.out_split = 0; p.relu_split = 0; p.out_rowmode = 0;
    p.accumulate = 0;
    p.split_rows = (int)slice_rows;
    p.split_b = g->B;
    CUtensorMap tcw;
    const uint32_t ucols = 32;
    if (g->a_mode == 0) {
      const uint64_t dims[2] = {(uint64_t)g->N, (uint64_t)(slice_rows * p.k_splits)};
      const uint64_t strides[1] = {(uint64_t)g->N * 4};
      const uint32_t box[2] = {ucols, BM};
      if (!make_map(&tcw, g->splitk_ws, 2, dims, strides, box, CU_TENSOR_MAP_DATA_TYPE_FLOAT32)) return DE_ERR_CUDA;
    } else {
      const uint64_t dims[4] = {(uint64_t)g->N, (uint64_t)g->W, (uint64_t)g->H, (uint64_t)g->B * p.k_splits};
      const uint64_t strides[3] = {(uint64_t)g->N * 4, (uint64_t)g->W * g->N * 4, (uint64_t)g->H * g->W * g->N * 4};
      const uint32_t box[4] = {ucols, (uint32_t)p.bw, (uint32_t)p.bh, 1};
      if (!make_map(&tcw, g->splitk_ws, 4, dims, strides, box, CU_TENSOR_MAP_DATA_TYPE_FLOAT32)) return DE_ERR_CUDA;
    }
    int rc;
    switch (bn) {
      case 256: rc = launch<256, 1, 0>(ta, tb, tcw, p, st); break;
      case 128: rc = launch<128, 1, 0>(ta, tb, tcw, p, st); break;
      case 64: rc = launch<64, 1, 0>(ta, tb, tcw, p, st); break;
      default: return fail(DE_ERR_INVALID, "de_gemm: split-K needs block_n >= 64");
    }
    if (rc != DE_OK) return rc;
    const long long nthr = (long long)f.M * (f.N / 4);
    splitk_finalize_kernel<<<(unsigned)((nthr + 255) / 256), 256, 0, st>>>(f);
    count_launch();
    return cuda_check_launch("splitk_finalize_kernel");
  }
  // epilogue flavour: the TMA-store epilogue covers "bias + activation -> dense output"
  const int es = g->out_dtype == DE_DTYPE_F32 ? 4 : 2;
  bool store_epi = !g->add1 && !g->add2 && !g->out_relu && !g->head_w && g->out_rowmode == 0 && bn >= 64 && !g->out_split &&
                   (g->ldo * es) % 16 == 0 && (reinterpret_cast<uintptr_t>(g->out) & 15) == 0 &&
                   (g->a_mode == 0 || g->ldo == g->N) && g->epilogue != 1;
  if (g->accumulate && (!store_epi || g->a_mode != 0 || g->out_dtype != DE_DTYPE_F32))
    return fail(DE_ERR_INVALID, "de_gemm: accumulate needs the dense fp32 store epilogue (a_mode 0, no add1/add2/out_relu/row maps)");
  // 3x3 convolution in halo mode (see GemmParams): decided here, where tile width, CTA pairing and the epilogue - hence
  // the bytes of the stage ring - are known.  DEB_CONV_HALO=0 keeps the nine-box main loop.
  if (g->a_mode == 1 && bn >= 64) {
    static int halo_on = -1;
    if (halo_on < 0) {
      const char* e = getenv("DEB_CONV_HALO");
      halo_on = (e && e[0] == '0') ? 0 : 1;
    }
    int hbw = 0, hbh = 0;
    if (halo_on && pick_conv_box_halo(g->H, g->W, p.bw, p.bh, &hbw, &hbh)) {
      const int ring = ring_bytes(bn, cg, store_epi ? 0 : 1), b_bytes = (bn / cg) * BK * 2;
      const int a_slot = ((hbh + 2) * hbw * BK * 2 + 1023) / 1024 * 1024;
      int na = 3, nb = (ring - na * a_slot) / b_bytes;
      if (nb < 5) {
        na = 2;
        nb = (ring - na * a_slot) / b_bytes;
      }
      if (nb > 13 - na) nb = 13 - na;
      if (nb >= 4) {
        p.halo = 1;
        p.na = na;
        p.nb = nb;
        p.a_slot = a_slot;
        p.bw = hbw;
        p.bh = hbh;
        p.tiles_x = (g->W + p.bw - 1) / p.bw;
        p.tiles_y = (g->H + p.bh - 1) / p.bh;
        p.m_tiles = g->B * p.tiles_x * p.tiles_y;
        const uint64_t dims[4] = {(uint64_t)g->Cin, (uint64_t)g->W, (uint64_t)g->H, (uint64_t)g->B};
        const uint64_t strides[3] = {(uint64_t)g->Cin * 2, (uint64_t)g->W * g->Cin * 2, (uint64_t)g->H * g->W * g->Cin * 2};
        const uint32_t box[4] = {BK, (uint32_t)p.bw, (uint32_t)(p.bh + 2), 1};
        if (!make_map(&ta, g->a, 4, dims, strides, box)) return DE_ERR_CUDA;
      }
    }
  }
  CUtensorMap tc = tb;
  if (store_epi) {
    const CUtensorMapDataType dt = es == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const uint32_t ucols = 128 / es;
    if (g->a_mode == 0) {
      const uint64_t dims[2] = {(uint64_t)g->N, (uint64_t)g->M};
      const uint64_t strides[1] = {(uint64_t)g->ldo * es};
      const uint32_t box[2] = {ucols, BM};
      if (!make_map(&tc, g->out, 2, dims, strides, box, dt)) return DE_ERR_CUDA;
    } else {
      const uint64_t dims[4] = {(uint64_t)g->N, (uint64_t)g->W, (uint64_t)g->H, (uint64_t)g->B};
      const uint64_t strides[3] = {(uint64_t)g->N * es, (uint64_t)g->W * g->N * es, (uint64_t)g->H * g->W * g->N * es};
      const uint32_t box[4] = {ucols, (uint32_t)p.bw, (uint32_t)p.bh, 1};
      if (!make_map(&tc, g->out, 4, dims, strides, box, dt)) return DE_ERR_CUDA;
    }
  }
  switch (bn) {
    case 256:
      if (cg == 2) return store_epi ? launch<256, 2, 0>(ta, tb, tc, p, st) : launch<256, 2, 1>(ta, tb, tc, p, st);
      return store_epi ? launch<256, 1, 0>(ta, tb, tc, p, st) : launch<256, 1, 1>(ta, tb, tc, p, st);
    case 128:
      if (cg == 2) return store_epi ? launch<128, 2, 0>(ta, tb, tc, p, st) : launch<128, 2, 1>(ta, tb, tc, p, st);
      return store_epi ? launch<128, 1, 0>(ta, tb, tc, p, st) : launch<128, 1, 1>(ta, tb, tc, p, st);
    case 64: return store_epi ? launch<64, 1, 0>(ta, tb, tc, p, st) : launch<64, 1, 1>(ta, tb, tc, p, st);
    case 32: return launch<32, 1, 1>(ta, tb, tc, p, st);
  }
  return fail(DE_ERR_INVALID, "de_gemm: block_n %d not in {32,64,128,256}", bn);
}
