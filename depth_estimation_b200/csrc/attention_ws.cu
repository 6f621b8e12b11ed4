// Multi-head self-attention forward, warp-specialised on tcgen05 + TMEM (replaces timm Attention.forward inside
// the ViT the reference runs at /root/reference/depth_midas.py:26).
//
// One persistent CTA per SM carries TWO independent flash-attention streams (each a sequence of 128-query
// tiles); twelve warps (three warpgroups, registers re-balanced with setmaxnreg):
//   warps 0-3 / 4-7 : softmax of stream 0 / 1 - one query row per thread (thread == TMEM lane)
//   warp 8          : TMA producer - double-buffered Q, K and V tiles of both streams
//   warps 9 / 10    : MMA issuer of stream 0 / 1 - S = Q K^T (128x128x64, SS) and O += P V (128x64x128, P read
//                     from TMEM); one issuing thread could not keep both streams fed (~100 cycles per tcgen05.mma)
// Measured facts that shaped it (clock64 traces on B200, DEB_ATTN_TRACE=1):
//   * These MMAs are small; with both operands in shared memory each one occupies the tensor pipe for
//     ~130 (PV) / ~210 (S) cycles - several times its math floor - so the tensor pipe, not MUFU.EX2, was the
//     limiter.  P therefore goes to TMEM (tcgen05.st, bf16 pairs) and the PV MMA takes its A operand from there.
//   * S of block j+1 is issued as soon as the softmax threads have pulled S of block j into registers (its own
//     barrier), so it runs under the exponentials instead of after them.
//   * O accumulates in TMEM across key blocks (no per-block read-back); the running max used as the exponent
//     reference only advances when it grew by more than 2^8 (any row of the warp), and only then is O rescaled
//     in place (tcgen05.ld / st).  P <= 256 and the fp32 row sum stay exact enough; the final O / l uses the
//     same reference, so the result is the ordinary softmax.
//   * the last (partial) key block skips the 32-column chunks that hold no key, and its PV MMA is cut to
//     ceil(keys/16) k-steps.
// TMEM columns (512): S0 0-127 | S1 128-255 | O0 256-319 | O1 320-383 | P0 384-447 | P1 448-511.
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

constexpr int kWsThreads = 384;                          // 3 warpgroups: softmax 0, softmax 1, {TMA, MMA 0, MMA 1, idle}
constexpr int kWsTile = 128 * 128;                       // [128 rows x 64 bf16], 128B-swizzled = 16 KB
constexpr int kWsStreamBytes = 6 * kWsTile;              // Q0 Q1 | K0 K1 | V0 V1
constexpr int kWsSmem = 2 * kWsStreamBytes + 1024 /*align slack*/ + 512 /*barriers*/;
constexpr float kRescaleThreshold = 8.0f;                // log2 units
#ifndef WS_POLY_MASK
#define WS_POLY_MASK 0x24                                // pairs 2 and 5 of every 8: a quarter of the exponentials
#endif
constexpr unsigned kWsPolyMask = WS_POLY_MASK;

DEV void tmem_st_32x32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
      "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
      "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
DEV void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
DEV float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
// tcgen05.mma, both operands from shared-memory descriptors, accumulate flag known at compile time
template <bool kAcc>
DEV void umma_bf16_c(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "n"(kAcc ? 1 : 0)
      : "memory");
}
// tcgen05.mma with the A operand in tensor memory: lane = row, each 32-bit column holds two consecutive bf16
// K elements (low half = even k), so one K=16 instruction reads 8 columns starting at a_tmem.
DEV void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

struct WsBars {                    // one set per stream
  uint64_t q_full[2], q_free[2], k_full[2], k_free[2], v_full[2], v_free[2];
  uint64_t s_full, s_free, p_full, pv_done, o_full, o_free;
  uint64_t kick;                   // stream 0 -> stream 1: first S tile consumed (stream 1 starts about half a period late)
};

struct WsParams {
  __nv_bfloat16* out;
  int T, heads, qtiles, nblk, ntiles;
  int out_split;    // 1: out rows are split3 operands [hi | hi | lo] (3*heads*64 wide)
  long long* dbg;   // DEB_ATTN_TRACE: per-block clock64 stamps of CTA 0 (debug only)
  float scale_log2e;
};

// trace slots: [stream][g < 64][stamp < 16]
#define WS_STAMP(k) do { if (p.dbg && blockIdx.x == 0 && w4 == 0 && lane == 0 && g < 64) p.dbg[(s * 64 + g) * 16 + (k)] = clock64(); } while (0)
#define MMA_STAMP(k) do { if (p.dbg && blockIdx.x == 0 && lane == 0 && g < 64) p.dbg[(s * 64 + g) * 16 + (k)] = clock64(); } while (0)

__global__ void __launch_bounds__(kWsThreads, 1)
attention_ws_kernel(const __grid_constant__ CUtensorMap tmQKV, const WsParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  WsBars* bars = reinterpret_cast<WsBars*>(smem + 2 * kWsStreamBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int D = p.heads * 64;
  const int nblk = p.nblk;
  const int grid = gridDim.x;
  const int slot1 = grid + blockIdx.x;
  const int ntile0 = (p.ntiles - (int)blockIdx.x + 2 * grid - 1) / (2 * grid);          // blockIdx.x < ntiles always
  const int ntile1 = slot1 < p.ntiles ? (p.ntiles - slot1 + 2 * grid - 1) / (2 * grid) : 0;
  const int last_valid = p.T - (nblk - 1) * 128;            // keys in the last block (1..128)

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQKV);
    for (int s = 0; s < 2; ++s) {
      WsBars& b = bars[s];
      for (int i = 0; i < 2; ++i) {
        mbar_init(&b.q_full[i], 1);
        mbar_init(&b.q_free[i], 1);
        mbar_init(&b.k_full[i], 1);
        mbar_init(&b.k_free[i], 1);
        mbar_init(&b.v_full[i], 1);
        mbar_init(&b.v_free[i], 1);
      }
      mbar_init(&b.s_full, 1);
      mbar_init(&b.s_free, 4);
      mbar_init(&b.p_full, 4);
      mbar_init(&b.pv_done, 1);
      mbar_init(&b.o_full, 1);
      mbar_init(&b.o_free, 4);
      mbar_init(&b.kick, 4);
    }
    mbar_fence_init();
  }
  if (warp == 9) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  pdl_launch();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  // tile index -> (row of the first query, row of the first key of the image, head column)
  auto tile_of = [&](int s, int it, int& q_row, int& kv_row, int& hcol) {
    const int t = s * grid + blockIdx.x + it * 2 * grid;
    const int qt = t % p.qtiles, bh = t / p.qtiles;
    const int h = bh % p.heads, b = bh / p.heads;
    kv_row = b * p.T;
    q_row = kv_row + qt * 128;
    hcol = h * 64;
  };

  // register budget: 384 x 168 at launch -> 2 x 128 x 216 (softmax) + 128 x 72 (service warpgroup)
  if (warp >= 8) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
  }
  if (warp == 8) {
    // ------------------------------------------------------------------ TMA producer
    // Converged warp, elect_one around the issue, tile coordinates computed once per tile (not per block): this
    // warp feeds two streams and must stay far below one block period per step.
    const int max_it = ntile0 > ntile1 ? ntile0 : ntile1;
    for (int it = 0; it < max_it; ++it) {
      int q_row[2], kv_row[2], hcol[2];
      tile_of(0, it, q_row[0], kv_row[0], hcol[0]);
      tile_of(1, it, q_row[1], kv_row[1], hcol[1]);
      for (int j = 0; j < nblk; ++j) {
        const int g = it * nblk + j;
        const int kb = g & 1;
        const uint32_t free_parity = ((g >> 1) - 1) & 1;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          if (it >= (s == 0 ? ntile0 : ntile1)) continue;
          WsBars& b = bars[s];
          uint8_t* base = smem + s * kWsStreamBytes;
          if (j == 0) {
            const int qb = it & 1;
            if (it >= 2) mbar_wait(&b.q_free[qb], ((it >> 1) - 1) & 1);
            if (elect_one()) {
              mbar_arrive_expect_tx(&b.q_full[qb], kWsTile);
              tma_load_2d(base + qb * kWsTile, &tmQKV, &b.q_full[qb], hcol[s], q_row[s]);
            }
            __syncwarp();
          }
          if (g >= 2) mbar_wait(&b.k_free[kb], free_parity);
          if (elect_one()) {
            mbar_arrive_expect_tx(&b.k_full[kb], kWsTile);
            tma_load_2d(base + (2 + kb) * kWsTile, &tmQKV, &b.k_full[kb], D + hcol[s], kv_row[s] + j * 128);
          }
          __syncwarp();
        }
#pragma unroll
        for (int s = 0; s < 2; ++s) {                           // V after both K: K gates the early S issue
          if (it >= (s == 0 ? ntile0 : ntile1)) continue;
          WsBars& b = bars[s];
          uint8_t* base = smem + s * kWsStreamBytes;
          if (g >= 2) mbar_wait(&b.v_free[kb], free_parity);
          if (elect_one()) {
            mbar_arrive_expect_tx(&b.v_full[kb], kWsTile);
            tma_load_2d(base + (4 + kb) * kWsTile, &tmQKV, &b.v_full[kb], 2 * D + hcol[s], kv_row[s] + j * 128);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 9 || warp == 10) {
    // ------------------------------------------------------------------ MMA issuers (one warp per stream)
    // The whole warp runs the (uniform) control flow; lane 0 fires the tcgen05 instructions.  Barrier waits that
    // belong together are taken by different lanes at once (one try_wait round trip instead of three).
    const int s = warp - 9;
    constexpr uint32_t idesc_s = make_idesc_bf16(128, 128);
    constexpr uint32_t idesc_o = make_idesc_bf16_bmn(128, 64);
    const int total = (s == 0 ? ntile0 : ntile1) * nblk;
    const int last_ksteps = (last_valid + 15) >> 4;
    WsBars& b = bars[s];
    const uint32_t base = smem_u32(smem + s * kWsStreamBytes);
    const uint64_t dQ0 = make_sw128_kmajor_desc(base), dQ1 = make_sw128_kmajor_desc(base + kWsTile);
    const uint64_t dK0 = make_sw128_kmajor_desc(base + 2 * kWsTile), dK1 = make_sw128_kmajor_desc(base + 3 * kWsTile);
    const uint64_t dV0 = make_sw128_mnmajor_desc(base + 4 * kWsTile), dV1 = make_sw128_mnmajor_desc(base + 5 * kWsTile);
    const uint32_t tS = tmem_base + s * 128, tO = tmem_base + 256 + s * 64, tP = tmem_base + 384 + s * 64;
    // converged waits, most-likely-late barrier first (a divergent "one lane per barrier" wait measured slower)
    auto wait3 = [&](uint64_t* b0, uint32_t p0, uint64_t* b1, uint32_t p1, uint64_t* b2, uint32_t p2) {
      mbar_wait(b0, p0);
      if (b1 != nullptr) mbar_wait(b1, p1);
      if (b2 != nullptr) mbar_wait(b2, p2);
    };
    int jn = 0, itn = 0;                                        // (block, tile) of the NEXT S to issue
    auto issue_s = [&](int g, bool first) {                     // S(g); `first`: no S(g-1) to wait for
      const int qb = itn & 1, kb = g & 1;
      wait3(&b.k_full[kb], (g >> 1) & 1, jn == 0 ? &b.q_full[qb] : nullptr, (itn >> 1) & 1,
            first ? nullptr : &b.s_free, (g - 1) & 1);
      tc_fence_after();
      const uint64_t adesc = qb ? dQ1 : dQ0, bdesc = kb ? dK1 : dK0;
      if (elect_one()) {                                        // converged warp + elect: UTCHMMAs issue back to back
        umma_bf16_c<false>(tS, adesc, bdesc, idesc_s);
        umma_bf16_c<true>(tS, adesc + 2, bdesc + 2, idesc_s);
        umma_bf16_c<true>(tS, adesc + 4, bdesc + 4, idesc_s);
        umma_bf16_c<true>(tS, adesc + 6, bdesc + 6, idesc_s);
        umma_commit(&b.s_full);
        umma_commit(&b.k_free[kb]);
        if (jn == nblk - 1) umma_commit(&b.q_free[qb]);
      }
      __syncwarp();
      if (++jn == nblk) {
        jn = 0;
        ++itn;
      }
    };
    // Both streams share the MUFU pipe.  Started together they run in lockstep - exponentials at the same time,
    // then max / P store / barrier latencies at the same time with the MUFU idle.  Stream 1 therefore starts when
    // stream 0 has pulled its first S tile (the Q/K loads and the first S MMA of stream 1 then add up to about half
    // a block period), so one stream's max / P-store / barrier latencies sit under the other's exp2.
    if (s == 1 && total > 0) mbar_wait(&bars[0].kick, 0);
    if (total > 0) issue_s(0, true);
    int j = 0, it = 0;                                          // (block, tile) of step g
    for (int g = 0; g < total; ++g) {
      if (g + 1 < total) {                                      // S(g+1) as soon as S(g) sits in registers
        MMA_STAMP(7);
        issue_s(g + 1, false);
        MMA_STAMP(8);
      }
      MMA_STAMP(5);
      wait3(&b.p_full, g & 1, &b.v_full[g & 1], (g >> 1) & 1, (j == 0 && it > 0) ? &b.o_free : nullptr, (it - 1) & 1);
      tc_fence_after();
      MMA_STAMP(6);
      const uint64_t bdesc = (g & 1) ? dV1 : dV0;
      constexpr uint64_t kVStep = 16 * 128 >> 4;                // 16 key rows of 128 bytes
      const int ksteps = j == nblk - 1 ? last_ksteps : 8;
      if (elect_one()) {
        umma_bf16_ts(tO, tP, bdesc, idesc_o, j != 0);
#pragma unroll
        for (int k = 1; k < 8; ++k)
          if (k < ksteps) umma_bf16_ts(tO, tP + k * 8, bdesc + k * kVStep, idesc_o, 1u);
        umma_commit(&b.v_free[g & 1]);
        umma_commit(&b.pv_done);
        if (j == nblk - 1) umma_commit(&b.o_full);
      }
      __syncwarp();
      MMA_STAMP(9);
      if (++j == nblk) {
        j = 0;
        ++it;
      }
    }
  } else if (warp < 8) {
    // ------------------------------------------------------------------ softmax warps
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int s = warp >> 2, w4 = warp & 3;
    const int r = w4 * 32 + lane;                               // query row in the tile == TMEM lane
    WsBars& b = bars[s];
    const uint32_t lane_sel = static_cast<uint32_t>(w4 * 32) << 16;
    const uint32_t tS = tmem_base + lane_sel + s * 128;
    const uint32_t tO = tmem_base + lane_sel + 256 + s * 64;
    const uint32_t tP = tmem_base + lane_sel + 384 + s * 64;
    const float sl2e = p.scale_log2e;
    const int ntile = s == 0 ? ntile0 : ntile1;
    for (int it = 0; it < ntile; ++it) {
      int q_row, kv_row, hcol;
      tile_of(s, it, q_row, kv_row, hcol);
      float m_run = -INFINITY, m_ref = -INFINITY, l_run = 0.f;
      for (int j = 0; j < nblk; ++j) {
        const int g = it * nblk + j;
        const int nvalid = p.T - j * 128;                       // keys of this block that exist (may exceed 128)
        const int nch = nvalid >= 128 ? 4 : (nvalid + 31) >> 5;
        WS_STAMP(0);
        mbar_wait(&b.s_full, g & 1);
        tc_fence_after();
        WS_STAMP(1);
        uint32_t sv[128];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld_32x32(tS + c * 32, sv + c * 32);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&b.s_free);                               // S(g) is in registers: S(g+1) may overwrite it
          if (s == 0 && g == 0) mbar_arrive(&b.kick);           // releases stream 1 (its start-up adds ~half a period)
        }
        if (nvalid < 128) {                                     // last block only (uniform): -inf -> exp2 gives exactly 0
#pragma unroll
          for (int i = 0; i < 128; ++i)
            if (i >= nvalid) sv[i] = 0xff800000u;
        }
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int i = 0; i < 128; i += 8) {                      // 3-input max (FMNMX3): 64 instructions, not 128
          mx4[0] = fmax3(mx4[0], __uint_as_float(sv[i]), __uint_as_float(sv[i + 1]));
          mx4[1] = fmax3(mx4[1], __uint_as_float(sv[i + 2]), __uint_as_float(sv[i + 3]));
          mx4[2] = fmax3(mx4[2], __uint_as_float(sv[i + 4]), __uint_as_float(sv[i + 5]));
          mx4[3] = fmax3(mx4[3], __uint_as_float(sv[i + 6]), __uint_as_float(sv[i + 7]));
        }
        WS_STAMP(2);
        const float m_new = fmaxf(m_run, fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])));
        m_run = m_new;
        bool pv_waited = false;
        if (j == 0) {
          m_ref = m_new;                                        // O is overwritten by PV_0: nothing to rescale
        } else if (__any_sync(0xffffffffu, (m_new - m_ref) * sl2e > kRescaleThreshold)) {
          mbar_wait(&b.pv_done, (g - 1) & 1);                   // O complete and idle: rescale it in place
          tc_fence_after();
          pv_waited = true;
          const float alpha = ex2_approx((m_ref - m_new) * sl2e);
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            uint32_t ov[32];
            tmem_ld_32x32(tO + c * 32, ov);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) ov[i] = __float_as_uint(__uint_as_float(ov[i]) * alpha);
            tmem_st_32x32(tO + c * 32, ov);
          }
          tmem_st_wait();
          l_run *= alpha;
          m_ref = m_new;
        }
        const float msc = m_ref * sl2e;
        const float2 sl2 = make_float2(sl2e, sl2e), nmsc2 = make_float2(-msc, -msc);
        float2 rs2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
        uint32_t pk[64];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (c < nch) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              // packed scale / subtract and row sums (FFMA2 / FADD2); kWsPolyMask selects the pairs of every 8 whose
              // exponentials run as a polynomial on the FMA pipe instead of MUFU.EX2 (the softmax bottleneck)
              const float2 x = ffma2(make_float2(__uint_as_float(sv[c * 32 + 2 * i]), __uint_as_float(sv[c * 32 + 2 * i + 1])), sl2, nmsc2);
              float2 pp;
              if ((kWsPolyMask >> (i & 7)) & 1) pp = ex2_poly2(x);
              else pp = make_float2(ex2_approx(x.x), ex2_approx(x.y));
              rs2[i & 1] = fadd2(rs2[i & 1], pp);
              pk[c * 16 + i] = pack_bf16(pp.x, pp.y);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) pk[c * 16 + i] = 0u;
          }
          if (c == 1) {                                         // first half of P leaves while the second is computed
            if (g > 0 && !pv_waited) {                          // PV(g-1) must be done reading P(g-1)
              mbar_wait(&b.pv_done, (g - 1) & 1);
              tc_fence_after();
            }
            tmem_st_32x32(tP, pk);
          }
        }
        l_run += (rs2[0].x + rs2[0].y) + (rs2[1].x + rs2[1].y);
        WS_STAMP(3);
        if (nch > 2) tmem_st_32x32(tP + 32, pk + 32);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&b.p_full);
        }
        WS_STAMP(4);
      }
      // ---- tile epilogue: O / l -> bf16 -> global (one 128-byte row segment per thread)
      mbar_wait(&b.o_full, it & 1);
      tc_fence_after();
      const float inv = 1.f / l_run;
      const int qrow_in_img = q_row - kv_row + r;
      // out_split: the row is written as the bf16x3 operand [hi | hi | lo] (each D wide) the proj GEMM consumes
      __nv_bfloat16* op = p.out + (static_cast<long long>(q_row) + r) * (p.out_split ? 3 * D : D) + hcol;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t ov[32];
        tmem_ld_32x32(tO + c * 32, ov);
        tmem_ld_wait();
        if (qrow_in_img < p.T) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float f[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(ov[8 * i + j]) * inv;
            uint4 v;
            v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]); v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
            *reinterpret_cast<uint4*>(op + c * 32 + 8 * i) = v;
            if (p.out_split) {
              uint4 lo;
              lo.x = pack_bf16(f[0] - bf16_lo(v.x), f[1] - bf16_hi(v.x)); lo.y = pack_bf16(f[2] - bf16_lo(v.y), f[3] - bf16_hi(v.y));
              lo.z = pack_bf16(f[4] - bf16_lo(v.z), f[5] - bf16_hi(v.z)); lo.w = pack_bf16(f[6] - bf16_lo(v.w), f[7] - bf16_hi(v.w));
              *reinterpret_cast<uint4*>(op + D + c * 32 + 8 * i) = v;
              *reinterpret_cast<uint4*>(op + 2 * D + c * 32 + 8 * i) = lo;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&b.o_free);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace deb

namespace deb {
int launch_attention_ws2(const CUtensorMap& tm, __nv_bfloat16* out, int T, int heads, int qtiles, int ntiles, int out_split,
                         float scale_log2e, cudaStream_t stream);   // attention_ws2.cu: two softmax threads per row
}
static int attention_impl(const void* qkv, void* out, int B, int T, int heads, float scale, int out_split, void* stream);
extern "C" int de_attention(const void* qkv, void* out, int B, int T, int heads, float scale, void* stream) {
  return attention_impl(qkv, out, B, T, heads, scale, 0, stream);
}
extern "C" int de_attention_split3(const void* qkv, void* out_split3, int B, int T, int heads, float scale, void* stream) {
  return attention_impl(qkv, out_split3, B, T, heads, scale, 1, stream);
}
static int attention_impl(const void* qkv, void* out, int B, int T, int heads, float scale, int out_split, void* stream) {
  deb::ProfScope prof_("attention", 4.0 * B * heads * (double)T * T * 64, stream, B, T, heads);
  using namespace deb;
  if (!qkv || !out || B <= 0 || T <= 0 || heads <= 0) return fail(DE_ERR_INVALID, "de_attention: bad args");
  if (!(scale > 0.f)) return fail(DE_ERR_INVALID, "de_attention: scale must be positive");
  auto fn = get_tensor_map_encoder();
  if (!fn) return DE_ERR_CUDA;
  CUtensorMap tm;
  const uint64_t dims[2] = {(uint64_t)3 * heads * 64, (uint64_t)B * T};
  const uint64_t strides[1] = {(uint64_t)3 * heads * 64 * 2};
  const uint32_t box[2] = {64, 128};
  const uint32_t estr[2] = {1, 1};
  CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(qkv), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(DE_ERR_CUDA, "de_attention: cuTensorMapEncodeTiled failed (%d)", (int)r);
  static const int v2 = getenv("DEB_ATTN_V2") ? atoi(getenv("DEB_ATTN_V2")) : 1;   // 0: one softmax thread per row (the kernel below)
  if (v2)
    return launch_attention_ws2(tm, reinterpret_cast<__nv_bfloat16*>(out), T, heads, (T + 127) / 128, B * heads * ((T + 127) / 128),
                                out_split, scale * 1.4426950408889634f, reinterpret_cast<cudaStream_t>(stream));
  static SmemOptIn optin;
  {
    cudaError_t e = ensure_smem(optin, attention_ws_kernel, (int)kWsSmem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(attention_ws)");
  }
  WsParams p{};
  p.out = reinterpret_cast<__nv_bfloat16*>(out);
  p.out_split = out_split;
  p.T = T;
  p.heads = heads;
  p.qtiles = (T + 127) / 128;
  p.nblk = (T + 127) / 128;
  p.ntiles = B * heads * p.qtiles;
  {
    static long long* dbg = nullptr;
    static int want = -1;
    if (want < 0) want = getenv("DEB_ATTN_TRACE") ? 1 : 0;
    if (want && !dbg) {
      cudaMalloc(&dbg, 2 * 64 * 16 * sizeof(long long));
      cudaMemset(dbg, 0, 2 * 64 * 16 * sizeof(long long));
      fprintf(stderr, "DEB_ATTN_TRACE buffer %p\n", (void*)dbg);
    }
    p.dbg = dbg;
  }
  p.scale_log2e = scale * 1.4426950408889634f;
  const int sms = num_sms();
  const int grid = p.ntiles < sms ? p.ntiles : sms;
  cudaError_t le = launch_pdl(2, attention_ws_kernel, dim3(grid), dim3(kWsThreads), kWsSmem, reinterpret_cast<cudaStream_t>(stream), tm, p);
  count_launch();
  if (le != cudaSuccess) return cuda_fail(le, "cudaLaunchKernelEx(attention_ws_kernel)");
  return cuda_check_launch("attention_ws_kernel");
}
