// Multi-head self-attention forward for the ViT encoder (timm Attention.forward, reached through
// /root/reference/depth_midas.py:26).  Non-causal, head_dim 64, sequence 577..1009 tokens.
// v1: flash-attention-2 style kernel - 64 query rows per CTA (16 per warp), K/V streamed in 64-key
// blocks through a cp.async double buffer, QK^T and PV on mma.sync m16n8k16 bf16 with fp32
// accumulate, online softmax in fp32 registers.  The score matrix is never materialised.
#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

constexpr int AT_BM = 64, AT_BN = 64, AT_THREADS = 128;

DEV uint32_t swz(int row, int chunk) { return static_cast<uint32_t>(row * 128 + ((chunk ^ (row & 7)) << 4)); }
DEV void cp_async16(uint32_t saddr, const void* g, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(saddr), "l"(g), "r"(src_bytes) : "memory");
}
DEV void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
DEV void cp_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
DEV void ldsm_x4(uint32_t* r, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
DEV void ldsm_x4_t(uint32_t* r, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
DEV void mma_bf16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// load a [64 rows][64 bf16] tile (rows row0.. of the token matrix, clamped/zero-filled past T)
DEV void load_tile(uint32_t s_tile, const __nv_bfloat16* g, long long ld, int row0, int T, int tid) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = tid + i * AT_THREADS;
    const int row = c >> 3, ch = c & 7;
    const int gr = row0 + row;
    const bool ok = gr < T;
    const __nv_bfloat16* src = g + (long long)(ok ? gr : T - 1) * ld + ch * 8;
    cp_async16(s_tile + swz(row, ch), src, ok ? 16 : 0);
  }
}

__global__ void __launch_bounds__(AT_THREADS) attention_kernel(const __nv_bfloat16* __restrict__ qkv,
                                                               __nv_bfloat16* __restrict__ out, int T, int heads,
                                                               float scale_log2e) {
  __shared__ __align__(128) uint8_t smem[AT_BM * 128 + 4 * AT_BN * 128];
  const uint32_t sQ = smem_u32(smem);
  const uint32_t sK = sQ + AT_BM * 128;
  const uint32_t sV = sK + 2 * AT_BN * 128;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q0 = blockIdx.x * AT_BM, h = blockIdx.y, b = blockIdx.z;
  const long long ld = 3LL * heads * 64;
  const __nv_bfloat16* base = qkv + (long long)b * T * ld;
  const __nv_bfloat16* gq = base + h * 64;
  const __nv_bfloat16* gk = base + (heads + h) * 64;
  const __nv_bfloat16* gv = base + (2 * heads + h) * 64;
  const int nblocks = (T + AT_BN - 1) / AT_BN;

  load_tile(sQ, gq, ld, q0, T, tid);
  load_tile(sK, gk, ld, 0, T, tid);
  load_tile(sV, gv, ld, 0, T, tid);
  cp_commit();

  uint32_t qf[4][4];
  float o[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  const int mi = lane >> 3, lr = lane & 7;

  for (int j = 0; j < nblocks; ++j) {
    const int buf = j & 1;
    if (j + 1 < nblocks) {
      load_tile(sK + (buf ^ 1) * AT_BN * 128, gk, ld, (j + 1) * AT_BN, T, tid);
      load_tile(sV + (buf ^ 1) * AT_BN * 128, gv, ld, (j + 1) * AT_BN, T, tid);
      cp_commit();
      cp_wait<1>();
    } else {
      cp_wait<0>();
    }
    __syncthreads();
    if (j == 0) {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks)
        ldsm_x4(qf[ks], sQ + swz(warp * 16 + (mi & 1) * 8 + lr, ks * 2 + (mi >> 1)));
    }
    const uint32_t kb = sK + buf * AT_BN * 128, vb = sV + buf * AT_BN * 128;
    float s[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int kp = 0; kp < 2; ++kp) {
        uint32_t bf[4];
        ldsm_x4(bf, kb + swz(nt * 8 + lr, kp * 4 + mi));
        mma_bf16(s[nt], qf[2 * kp], bf[0], bf[1]);
        mma_bf16(s[nt], qf[2 * kp + 1], bf[2], bf[3]);
      }
    }
    if ((j + 1) * AT_BN > T) {
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const int key = j * AT_BN + nt * 8 + (lane & 3) * 2;
        if (key >= T) s[nt][0] = s[nt][2] = -INFINITY;
        if (key + 1 >= T) s[nt][1] = s[nt][3] = -INFINITY;
      }
    }
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      mx[0] = fmaxf(mx[0], fmaxf(s[nt][0], s[nt][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[nt][2], s[nt][3]));
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
    }
    float alpha[2], msc[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const float m_new = fmaxf(m_run[r], mx[r]);
      alpha[r] = exp2f((m_run[r] - m_new) * scale_log2e);
      m_run[r] = m_new;
      msc[r] = m_new * scale_log2e;
      l_run[r] *= alpha[r];
    }
    uint32_t pa[4][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const float p0 = exp2f(fmaf(s[nt][0], scale_log2e, -msc[0]));
      const float p1 = exp2f(fmaf(s[nt][1], scale_log2e, -msc[0]));
      const float p2 = exp2f(fmaf(s[nt][2], scale_log2e, -msc[1]));
      const float p3 = exp2f(fmaf(s[nt][3], scale_log2e, -msc[1]));
      l_run[0] += p0 + p1;
      l_run[1] += p2 + p3;
      pa[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16(p0, p1);
      pa[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
      o[nt][0] *= alpha[0];
      o[nt][1] *= alpha[0];
      o[nt][2] *= alpha[1];
      o[nt][3] *= alpha[1];
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int dp = 0; dp < 4; ++dp) {
        uint32_t bf[4];
        ldsm_x4_t(bf, vb + swz(kk * 16 + (mi & 1) * 8 + lr, dp * 2 + (mi >> 1)));
        mma_bf16(o[2 * dp], pa[kk], bf[0], bf[1]);
        mma_bf16(o[2 * dp + 1], pa[kk], bf[2], bf[3]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  const float inv0 = 1.f / l_run[0], inv1 = 1.f / l_run[1];
  const int row0 = q0 + warp * 16 + (lane >> 2), row1 = row0 + 8;
  const long long ldo = (long long)heads * 64;
  __nv_bfloat16* ob = out + (long long)b * T * ldo + h * 64 + (lane & 3) * 2;
#pragma unroll
  for (int dt = 0; dt < 8; ++dt) {
    if (row0 < T) *reinterpret_cast<uint32_t*>(ob + row0 * ldo + dt * 8) = pack_bf16(o[dt][0] * inv0, o[dt][1] * inv0);
    if (row1 < T) *reinterpret_cast<uint32_t*>(ob + row1 * ldo + dt * 8) = pack_bf16(o[dt][2] * inv1, o[dt][3] * inv1);
  }
}

}  // namespace deb

extern "C" int de_attention_mma(const void* qkv, void* out, int B, int T, int heads, float scale, void* stream) {
  using namespace deb;
  if (!qkv || !out || B <= 0 || T <= 0 || heads <= 0) return fail(DE_ERR_INVALID, "de_attention_mma: bad args");
  dim3 grid((T + AT_BM - 1) / AT_BM, heads, B);
  attention_kernel<<<grid, AT_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<__nv_bfloat16*>(out), T, heads,
      scale * 1.4426950408889634f);
  count_launch();
  return cuda_check_launch("attention_kernel");
}
