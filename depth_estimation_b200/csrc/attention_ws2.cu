// Second softmax organisation of the warp-specialised attention kernel (attention_ws.cu describes the pipeline and
// the measurements behind it): TWO threads per query row.  Sixteen softmax warps instead of eight - warp (stream s,
// quadrant q, half h) owns the 32 rows of TMEM quadrant q and the key columns [64h, 64h + 64) of the stream's S tile -
// so four instead of two softmax warps sit on every scheduler and one row's 128 exponentials are two dependent chains of
// 64 (the softmax is bound by the latency of that chain, not by MUFU throughput: DESIGN.md section 4, "What the
// measurements taught").  The two halves of a row agree on the block maximum through shared memory (one float per row,
// a 64-thread named barrier per (stream, quadrant)), therefore take the same lazy-rescale decisions, write their own 32
// packed P columns, rescale / normalise their own 32 O columns and add their partial row sums at the end of the tile.
// S is read from TMEM twice per block in 32-column chunks (once for the maximum, once for the exponentials) so that a
// thread never holds more than 32 scores: 640 threads leave 96 registers per thread, no setmaxnreg.
// TMEM columns (512): S0 0-127 | S1 128-255 | O0 256-319 | O1 320-383 | P0 384-447 | P1 448-511 (as attention_ws.cu).
// The default since its measurement (profiles/r02_attention_ws2.md); DEB_ATTN_V2=0 selects attention_ws_kernel.
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

constexpr int kW2Threads = 640;                          // 16 softmax warps + {TMA, MMA 0, MMA 1, idle}
constexpr int kW2Tile = 128 * 128;                       // [128 rows x 64 bf16], 128B-swizzled = 16 KB
constexpr int kW2StreamBytes = 6 * kW2Tile;              // Q0 Q1 | K0 K1 | V0 V1
constexpr int kW2XchBytes = 2 * 2 * 2 * 128 * 4 + 2 * 2 * 128 * 4;   // block maxima [stream][parity][half][row] + row sums [stream][half][row]
constexpr int kW2Smem = 2 * kW2StreamBytes + 1024 /*align slack*/ + 512 /*barriers*/ + kW2XchBytes;
constexpr float kW2RescaleThreshold = 8.0f;              // log2 units
#ifndef WS_POLY_MASK
#define WS_POLY_MASK 0x24                                // pairs 2 and 5 of every 8: a quarter of the exponentials
#endif
constexpr unsigned kW2PolyMask = WS_POLY_MASK;

DEV void w2_tmem_st_32x16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
DEV void w2_tmem_st_32x32(uint32_t taddr, const uint32_t* r) {
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
DEV void w2_tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
DEV float w2_fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
template <bool kAcc>
DEV void w2_umma_c(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "n"(kAcc ? 1 : 0)
      : "memory");
}
DEV void w2_umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

struct W2Bars {                    // one set per stream
  uint64_t q_full[2], q_free[2], k_full[2], k_free[2], v_full[2], v_free[2];
  uint64_t s_full, s_free, p_full, pv_done, o_full, o_free;
  uint64_t kick;                   // stream 0 -> stream 1: first S tile consumed
};

struct W2Params {
  __nv_bfloat16* out;
  int T, heads, qtiles, nblk, ntiles;
  int out_split;    // 1: out rows are split3 operands [hi | hi | lo] (3*heads*64 wide)
  float scale_log2e;
};

__global__ void __launch_bounds__(kW2Threads, 1)
attention_ws2_kernel(const __grid_constant__ CUtensorMap tmQKV, const W2Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  W2Bars* bars = reinterpret_cast<W2Bars*>(smem + 2 * kW2StreamBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  float* xch_max = reinterpret_cast<float*>(smem + 2 * kW2StreamBytes + 512);   // [stream][parity][half][128]
  float* xch_sum = xch_max + 2 * 2 * 2 * 128;                                    // [stream][half][128]

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
      W2Bars& b = bars[s];
      for (int i = 0; i < 2; ++i) {
        mbar_init(&b.q_full[i], 1);
        mbar_init(&b.q_free[i], 1);
        mbar_init(&b.k_full[i], 1);
        mbar_init(&b.k_free[i], 1);
        mbar_init(&b.v_full[i], 1);
        mbar_init(&b.v_free[i], 1);
      }
      mbar_init(&b.s_full, 1);
      mbar_init(&b.s_free, 8);
      mbar_init(&b.p_full, 8);
      mbar_init(&b.pv_done, 1);
      mbar_init(&b.o_full, 1);
      mbar_init(&b.o_free, 8);
      mbar_init(&b.kick, 8);
    }
    mbar_fence_init();
  }
  if (warp == 17) {
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

  if (warp == 16) {
    // ------------------------------------------------------------------ TMA producer (as attention_ws.cu)
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
          W2Bars& b = bars[s];
          uint8_t* base = smem + s * kW2StreamBytes;
          if (j == 0) {
            const int qb = it & 1;
            if (it >= 2) mbar_wait(&b.q_free[qb], ((it >> 1) - 1) & 1);
            if (elect_one()) {
              mbar_arrive_expect_tx(&b.q_full[qb], kW2Tile);
              tma_load_2d(base + qb * kW2Tile, &tmQKV, &b.q_full[qb], hcol[s], q_row[s]);
            }
            __syncwarp();
          }
          if (g >= 2) mbar_wait(&b.k_free[kb], free_parity);
          if (elect_one()) {
            mbar_arrive_expect_tx(&b.k_full[kb], kW2Tile);
            tma_load_2d(base + (2 + kb) * kW2Tile, &tmQKV, &b.k_full[kb], D + hcol[s], kv_row[s] + j * 128);
          }
          __syncwarp();
        }
#pragma unroll
        for (int s = 0; s < 2; ++s) {                           // V after both K: K gates the early S issue
          if (it >= (s == 0 ? ntile0 : ntile1)) continue;
          W2Bars& b = bars[s];
          uint8_t* base = smem + s * kW2StreamBytes;
          if (g >= 2) mbar_wait(&b.v_free[kb], free_parity);
          if (elect_one()) {
            mbar_arrive_expect_tx(&b.v_full[kb], kW2Tile);
            tma_load_2d(base + (4 + kb) * kW2Tile, &tmQKV, &b.v_full[kb], 2 * D + hcol[s], kv_row[s] + j * 128);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 17 || warp == 18) {
    // ------------------------------------------------------------------ MMA issuers (one warp per stream; as attention_ws.cu)
    const int s = warp - 17;
    constexpr uint32_t idesc_s = make_idesc_bf16(128, 128);
    constexpr uint32_t idesc_o = make_idesc_bf16_bmn(128, 64);
    const int total = (s == 0 ? ntile0 : ntile1) * nblk;
    const int last_ksteps = (last_valid + 15) >> 4;
    W2Bars& b = bars[s];
    const uint32_t base = smem_u32(smem + s * kW2StreamBytes);
    const uint64_t dQ0 = make_sw128_kmajor_desc(base), dQ1 = make_sw128_kmajor_desc(base + kW2Tile);
    const uint64_t dK0 = make_sw128_kmajor_desc(base + 2 * kW2Tile), dK1 = make_sw128_kmajor_desc(base + 3 * kW2Tile);
    const uint64_t dV0 = make_sw128_mnmajor_desc(base + 4 * kW2Tile), dV1 = make_sw128_mnmajor_desc(base + 5 * kW2Tile);
    const uint32_t tS = tmem_base + s * 128, tO = tmem_base + 256 + s * 64, tP = tmem_base + 384 + s * 64;
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
      if (elect_one()) {
        w2_umma_c<false>(tS, adesc, bdesc, idesc_s);
        w2_umma_c<true>(tS, adesc + 2, bdesc + 2, idesc_s);
        w2_umma_c<true>(tS, adesc + 4, bdesc + 4, idesc_s);
        w2_umma_c<true>(tS, adesc + 6, bdesc + 6, idesc_s);
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
    if (s == 1 && total > 0) mbar_wait(&bars[0].kick, 0);
    if (total > 0) issue_s(0, true);
    int j = 0, it = 0;                                          // (block, tile) of step g
    for (int g = 0; g < total; ++g) {
      if (g + 1 < total) issue_s(g + 1, false);                 // S(g+1) as soon as S(g) has been read
      wait3(&b.p_full, g & 1, &b.v_full[g & 1], (g >> 1) & 1, (j == 0 && it > 0) ? &b.o_free : nullptr, (it - 1) & 1);
      tc_fence_after();
      const uint64_t bdesc = (g & 1) ? dV1 : dV0;
      constexpr uint64_t kVStep = 16 * 128 >> 4;                // 16 key rows of 128 bytes
      const int ksteps = j == nblk - 1 ? last_ksteps : 8;
      if (elect_one()) {
        w2_umma_ts(tO, tP, bdesc, idesc_o, j != 0);
#pragma unroll
        for (int k = 1; k < 8; ++k)
          if (k < ksteps) w2_umma_ts(tO, tP + k * 8, bdesc + k * kVStep, idesc_o, 1u);
        umma_commit(&b.v_free[g & 1]);
        umma_commit(&b.pv_done);
        if (j == nblk - 1) umma_commit(&b.o_full);
      }
      __syncwarp();
      if (++j == nblk) {
        j = 0;
        ++it;
      }
    }
  } else if (warp < 16) {
    // ------------------------------------------------------------------ softmax warps: (stream, quadrant, column half)
    const int s = warp >> 3, q = warp & 3, half = (warp >> 2) & 1;
    const int r = q * 32 + lane;                                // query row in the tile == TMEM lane
    W2Bars& b = bars[s];
    const uint32_t lane_sel = static_cast<uint32_t>(q * 32) << 16;
    const uint32_t tS = tmem_base + lane_sel + s * 128 + half * 64;
    const uint32_t tO = tmem_base + lane_sel + 256 + s * 64 + half * 32;
    const uint32_t tP = tmem_base + lane_sel + 384 + s * 64 + half * 32;
    const int pair_bar = 1 + s * 4 + q;                         // named barrier of the two warps that share these rows
    float* my_max = xch_max + (s * 4 + half) * 128 + r;         // + parity * 256
    const float* his_max = xch_max + (s * 4 + (half ^ 1)) * 128 + r;
    const float sl2e = p.scale_log2e;
    const int ntile = s == 0 ? ntile0 : ntile1;
    for (int it = 0; it < ntile; ++it) {
      int q_row, kv_row, hcol;
      tile_of(s, it, q_row, kv_row, hcol);
      float m_run = -INFINITY, m_ref = -INFINITY, l_run = 0.f;  // m_*: identical in both halves; l_run: this half's share
      for (int j = 0; j < nblk; ++j) {
        const int g = it * nblk + j;
        const int nv = min(max(p.T - j * 128 - half * 64, 0), 64);   // keys of this thread's 64 columns that exist
        mbar_wait(&b.s_full, g & 1);
        tc_fence_after();
        // ---- pass 1: the maximum of this half's columns
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t sv[32];
          tmem_ld_32x32(tS + c * 32, sv);
          tmem_ld_wait();
          if (nv < 64) {                                        // last block only (warp-uniform)
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (c * 32 + i >= nv) sv[i] = 0xff800000u;
          }
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            mx4[0] = w2_fmax3(mx4[0], __uint_as_float(sv[i]), __uint_as_float(sv[i + 1]));
            mx4[1] = w2_fmax3(mx4[1], __uint_as_float(sv[i + 2]), __uint_as_float(sv[i + 3]));
            mx4[2] = w2_fmax3(mx4[2], __uint_as_float(sv[i + 4]), __uint_as_float(sv[i + 5]));
            mx4[3] = w2_fmax3(mx4[3], __uint_as_float(sv[i + 6]), __uint_as_float(sv[i + 7]));
          }
        }
        const float m_loc = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        my_max[(g & 1) * 256] = m_loc;
        named_bar_sync(pair_bar, 64);
        const float m_new = fmaxf(m_run, fmaxf(m_loc, his_max[(g & 1) * 256]));
        m_run = m_new;
        bool pv_waited = false;
        if (j == 0) {
          m_ref = m_new;                                        // O is overwritten by PV_0: nothing to rescale
        } else if (__any_sync(0xffffffffu, (m_new - m_ref) * sl2e > kW2RescaleThreshold)) {
          mbar_wait(&b.pv_done, (g - 1) & 1);                   // O complete and idle: rescale this half's 32 columns
          tc_fence_after();
          pv_waited = true;
          const float alpha = ex2_approx((m_ref - m_new) * sl2e);
          uint32_t ov[32];
          tmem_ld_32x32(tO, ov);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) ov[i] = __float_as_uint(__uint_as_float(ov[i]) * alpha);
          w2_tmem_st_32x32(tO, ov);
          w2_tmem_st_wait();
          l_run *= alpha;
          m_ref = m_new;
        }
        // ---- pass 2: exponentials, 32 columns at a time
        const float msc = m_ref * sl2e;
        const float2 sl2 = make_float2(sl2e, sl2e), nmsc2 = make_float2(-msc, -msc);
        float2 rs2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t sv[32];
          tmem_ld_32x32(tS + c * 32, sv);
          tmem_ld_wait();
          if (c == 1) {                                         // S(g) has been read for the last time
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              mbar_arrive(&b.s_free);
              if (s == 0 && g == 0) mbar_arrive(&b.kick);       // releases stream 1
            }
          }
          if (nv < 64) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (c * 32 + i >= nv) sv[i] = 0xff800000u;        // exp2(-inf) = 0 exactly
          }
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float2 x = ffma2(make_float2(__uint_as_float(sv[2 * i]), __uint_as_float(sv[2 * i + 1])), sl2, nmsc2);
            float2 pp;
            if ((kW2PolyMask >> (i & 7)) & 1) pp = ex2_poly2(x);
            else pp = make_float2(ex2_approx(x.x), ex2_approx(x.y));
            rs2[i & 1] = fadd2(rs2[i & 1], pp);
            pk[i] = pack_bf16(pp.x, pp.y);
          }
          if (c == 0 && g > 0 && !pv_waited) {                  // PV(g-1) must be done reading P(g-1)
            mbar_wait(&b.pv_done, (g - 1) & 1);
            tc_fence_after();
          }
          w2_tmem_st_32x16(tP + c * 16, pk);
        }
        l_run += (rs2[0].x + rs2[0].y) + (rs2[1].x + rs2[1].y);
        w2_tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&b.p_full);
      }
      // ---- tile epilogue: O / l -> bf16 -> global; this thread owns 32 of the row's 64 output columns
      xch_sum[(s * 2 + half) * 128 + r] = l_run;
      named_bar_sync(pair_bar, 64);
      const float l_tot = half == 0 ? l_run + xch_sum[(s * 2 + 1) * 128 + r] : xch_sum[(s * 2) * 128 + r] + l_run;
      named_bar_sync(pair_bar, 64);                             // both sums read before the next tile overwrites them
      mbar_wait(&b.o_full, it & 1);
      tc_fence_after();
      const float inv = 1.f / l_tot;
      const int qrow_in_img = q_row - kv_row + r;
      __nv_bfloat16* op = p.out + (static_cast<long long>(q_row) + r) * (p.out_split ? 3 * D : D) + hcol + half * 32;
      {
        uint32_t ov[32];
        tmem_ld_32x32(tO, ov);
        tmem_ld_wait();
        if (qrow_in_img < p.T) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float f[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] = __uint_as_float(ov[8 * i + k]) * inv;
            uint4 v;
            v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]); v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
            *reinterpret_cast<uint4*>(op + 8 * i) = v;
            if (p.out_split) {
              uint4 lo;
              lo.x = pack_bf16(f[0] - bf16_lo(v.x), f[1] - bf16_hi(v.x)); lo.y = pack_bf16(f[2] - bf16_lo(v.y), f[3] - bf16_hi(v.y));
              lo.z = pack_bf16(f[4] - bf16_lo(v.z), f[5] - bf16_hi(v.z)); lo.w = pack_bf16(f[6] - bf16_lo(v.w), f[7] - bf16_hi(v.w));
              *reinterpret_cast<uint4*>(op + D + 8 * i) = v;
              *reinterpret_cast<uint4*>(op + 2 * D + 8 * i) = lo;
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
  if (warp == 17) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// launched by attention_impl (attention_ws.cu) unless DEB_ATTN_V2=0
int launch_attention_ws2(const CUtensorMap& tm, __nv_bfloat16* out, int T, int heads, int qtiles, int ntiles, int out_split,
                         float scale_log2e, cudaStream_t stream) {
  static SmemOptIn optin;
  cudaError_t e = ensure_smem(optin, attention_ws2_kernel, (int)kW2Smem);
  if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(attention_ws2)");
  W2Params p{};
  p.out = out;
  p.out_split = out_split;
  p.T = T;
  p.heads = heads;
  p.qtiles = qtiles;
  p.nblk = qtiles;
  p.ntiles = ntiles;
  p.scale_log2e = scale_log2e;
  const int sms = num_sms();
  const int grid = ntiles < sms ? ntiles : sms;
  cudaError_t le = launch_pdl(2, attention_ws2_kernel, dim3(grid), dim3(kW2Threads), kW2Smem, stream, tm, p);
  count_launch();
  if (le != cudaSuccess) return cuda_fail(le, "cudaLaunchKernelEx(attention_ws2_kernel)");
  return cuda_check_launch("attention_ws2_kernel");
}

}  // namespace deb
