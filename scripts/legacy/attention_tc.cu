// Multi-head self-attention forward on the 5th-gen tensor cores (tcgen05 + TMEM), flash style.
//
// One CTA = one (128-query tile, head, image); two CTAs share an SM so one tile's softmax overlaps the other's
// MMAs.  Per 128-key block:  S = Q K^T  (tcgen05.mma 128x128x64, accumulator in TMEM)  ->  the 128 threads
// (one per query row) read their S row from TMEM, update the running max / sum in fp32 and write
// P = exp2(..) as bf16 into a 128B-swizzled smem tile  ->  O_j = P V  (tcgen05.mma 128x64x128 with V taken
// straight from the qkv buffer as an MN-major operand - no transpose)  ->  O_j is read back and folded into
// the per-thread fp32 output row (O = O*alpha + O_j), so TMEM never needs a read-modify-write rescale.
// Q/K/V tiles arrive by TMA from the packed qkv matrix; K/V blocks are double-buffered one block ahead and
// the next S MMA is issued right behind the current PV so the tensor pipe works through the softmax.
#include <cuda.h>

#include "common.cuh"
#include "depth_b200.h"
#include "host_util.h"

namespace deb {

constexpr int AQ = 128;       // queries per CTA
constexpr int AK = 128;       // keys per block
constexpr int AD = 64;        // head dim
constexpr int kAttThreads = 128;
constexpr int kTileBytes = 128 * 128;   // one [128 rows x 64 bf16] swizzled tile = 16 KB
// smem: Q | K0 K1 | V0 V1 | P(lo) P(hi) | barriers
constexpr int kAttSmem = kTileBytes * 7 + 64;

__global__ void __launch_bounds__(kAttThreads, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, __nv_bfloat16* __restrict__ out, int T, int heads,
                    float scale_log2e) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if (smem_u32(smem) & 1023u) __trap();          // the swizzled tiles need 1024-byte alignment
  uint8_t* sQ = smem;
  uint8_t* sK = smem + kTileBytes;
  uint8_t* sV = smem + 3 * kTileBytes;
  uint8_t* sP = smem + 5 * kTileBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 7 * kTileBytes);
  uint64_t* bar_q = bars;
  uint64_t* bar_kv = bars + 1;   // [2]
  uint64_t* bar_s = bars + 3;
  uint64_t* bar_o = bars + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q0 = blockIdx.x * AQ, h = blockIdx.y, b = blockIdx.z;
  const int D = heads * AD;
  const int nblk = (T + AK - 1) / AK;
  const int row_base = b * T;

  if (tid == 0) {
    tma_prefetch_desc(&tmQKV);
    mbar_init(bar_q, 1);
    mbar_init(&bar_kv[0], 1);
    mbar_init(&bar_kv[1], 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    mbar_fence_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base;          // 128 columns
  const uint32_t tO = tmem_base + 128;    // 64 columns
  const uint32_t lane_sel = static_cast<uint32_t>(warp * 32) << 16;

  constexpr uint32_t idesc_s = make_idesc_bf16(128, AK);
  constexpr uint32_t idesc_o = make_idesc_bf16_bmn(128, AD);

  auto load_kv = [&](int j) {
    const int buf = j & 1;
    mbar_arrive_expect_tx(&bar_kv[buf], 2 * kTileBytes);
    tma_load_2d(sK + buf * kTileBytes, &tmQKV, &bar_kv[buf], D + h * AD, row_base + j * AK);
    tma_load_2d(sV + buf * kTileBytes, &tmQKV, &bar_kv[buf], 2 * D + h * AD, row_base + j * AK);
  };
  auto issue_s = [&](int j) {               // S = Q K_j^T
    const int buf = j & 1;
    mbar_wait(&bar_kv[buf], (j >> 1) & 1);
    tc_fence_after();
    const uint64_t adesc = make_sw128_kmajor_desc(smem_u32(sQ));
    const uint64_t bdesc = make_sw128_kmajor_desc(smem_u32(sK + buf * kTileBytes));
#pragma unroll
    for (int k = 0; k < AD / 16; ++k) umma_bf16(tS, adesc + 2 * k, bdesc + 2 * k, idesc_s, k != 0);
    umma_commit(bar_s);
  };

  if (tid == 0) {
    mbar_arrive_expect_tx(bar_q, kTileBytes);
    tma_load_2d(sQ, &tmQKV, bar_q, h * AD, row_base + q0);
    load_kv(0);
    if (nblk > 1) load_kv(1);
    mbar_wait(bar_q, 0);
    issue_s(0);
  }

  float o[AD];
#pragma unroll
  for (int i = 0; i < AD; ++i) o[i] = 0.f;
  float m_run = -INFINITY, l_run = 0.f;
  const int r = tid;                                   // query row inside the tile == TMEM lane
  uint8_t* prow = sP + r * 128;

  for (int j = 0; j < nblk; ++j) {
    const int buf = j & 1;
    const int nvalid = T - j * AK;                     // keys of this block that exist (may exceed AK)
    mbar_wait(bar_s, j & 1);
    tc_fence_after();
    // ---- the whole 128-key score row of this thread lives in registers: one TMEM read per block
    uint32_t sv[AK];
#pragma unroll
    for (int c = 0; c < AK / 32; ++c) tmem_ld_32x32(tS + lane_sel + c * 32, sv + c * 32);
    tmem_ld_wait();
    if (nvalid < AK) {                                 // last block only (CTA-uniform): -inf -> exp2 gives exactly 0
#pragma unroll
      for (int i = 0; i < AK; ++i)
        if (i >= nvalid) sv[i] = 0xff800000u;
    }
    float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int i = 0; i < AK; i += 4) {
      mx4[0] = fmaxf(mx4[0], __uint_as_float(sv[i]));
      mx4[1] = fmaxf(mx4[1], __uint_as_float(sv[i + 1]));
      mx4[2] = fmaxf(mx4[2], __uint_as_float(sv[i + 2]));
      mx4[3] = fmaxf(mx4[3], __uint_as_float(sv[i + 3]));
    }
    const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
    const float m_new = fmaxf(m_run, mx);
    const float alpha = ex2_approx((m_run - m_new) * scale_log2e);
    const float msc = m_new * scale_log2e;
    m_run = m_new;
    // ---- P = exp2(S*scale - max*scale) -> bf16 -> swizzled smem (A operand of the PV MMA)
    float rs4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < AK / 32; ++c) {
      uint32_t pk[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float p0 = ex2_approx(fmaf(__uint_as_float(sv[c * 32 + 2 * i]), scale_log2e, -msc));
        const float p1 = ex2_approx(fmaf(__uint_as_float(sv[c * 32 + 2 * i + 1]), scale_log2e, -msc));
        rs4[i & 3] += p0 + p1;
        pk[i] = pack_bf16(p0, p1);
      }
      uint8_t* base = prow + (c >> 1) * kTileBytes;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int jj = (c & 1) * 4 + i;
        *reinterpret_cast<uint4*>(base + ((jj ^ (r & 7)) << 4)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
      }
    }
    const float rowsum = (rs4[0] + rs4[1]) + (rs4[2] + rs4[3]);
    l_run = l_run * alpha + rowsum;
    if (j > 0) {                                       // O_{j-1}: complete (S_j was committed behind it)
#pragma unroll
      for (int c = 0; c < AD / 32; ++c) {
        uint32_t ov[32];
        tmem_ld_32x32(tO + lane_sel + c * 32, ov);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c * 32 + i] += __uint_as_float(ov[i]);
      }
    }
    if (alpha != 1.f) {
#pragma unroll
      for (int i = 0; i < AD; ++i) o[i] *= alpha;
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();                                   // P complete, everyone is done reading S
    if (tid == 0) {
      tc_fence_after();
      const uint64_t bdesc = make_sw128_mnmajor_desc(smem_u32(sV + buf * kTileBytes));
#pragma unroll
      for (int k = 0; k < AK / 16; ++k) {
        const uint64_t adesc = make_sw128_kmajor_desc(smem_u32(sP + (k >> 2) * kTileBytes)) + 2 * (k & 3);
        umma_bf16(tO, adesc, bdesc + static_cast<uint64_t>(k) * (16 * 128 >> 4), idesc_o, k != 0);
      }
      umma_commit(bar_o);
      if (j + 1 < nblk) issue_s(j + 1);                // tensor pipe keeps going during the O read-back
    }
    // K/V buffer `buf` and P are free once PV_j has completed (bar_o); prefetch block j+2 into it
    if (tid == 0 && j + 2 < nblk) {
      mbar_wait(bar_o, j & 1);
      load_kv(j + 2);
    }
  }

  mbar_wait(bar_o, (nblk - 1) & 1);
  tc_fence_after();
#pragma unroll
  for (int c = 0; c < AD / 32; ++c) {
    uint32_t ov[32];
    tmem_ld_32x32(tO + lane_sel + c * 32, ov);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) o[c * 32 + i] += __uint_as_float(ov[i]);
  }
  const int qrow = q0 + r;
  if (qrow < T) {
    const float inv = 1.f / l_run;
    __nv_bfloat16* op = out + (static_cast<long long>(row_base) + qrow) * D + h * AD;
#pragma unroll
    for (int i = 0; i < AD / 8; ++i) {
      uint4 v;
      v.x = pack_bf16(o[8 * i] * inv, o[8 * i + 1] * inv);
      v.y = pack_bf16(o[8 * i + 2] * inv, o[8 * i + 3] * inv);
      v.z = pack_bf16(o[8 * i + 4] * inv, o[8 * i + 5] * inv);
      v.w = pack_bf16(o[8 * i + 6] * inv, o[8 * i + 7] * inv);
      *reinterpret_cast<uint4*>(op + 8 * i) = v;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

}  // namespace deb

// First tcgen05 attention kernel (two CTAs per SM, O folded in registers).  Superseded by attention_ws.cu; reachable
// with DEB_ATTENTION=1 as a cross-check.
extern "C" int de_attention_v1(const void* qkv, void* out, int B, int T, int heads, float scale, void* stream) {
  using namespace deb;
  if (!qkv || !out || B <= 0 || T <= 0 || heads <= 0) return fail(DE_ERR_INVALID, "de_attention: bad args");
  auto fn = get_tensor_map_encoder();
  if (!fn) return DE_ERR_CUDA;
  CUtensorMap tm;
  const uint64_t dims[2] = {(uint64_t)3 * heads * AD, (uint64_t)B * T};
  const uint64_t strides[1] = {(uint64_t)3 * heads * AD * 2};
  const uint32_t box[2] = {AD, 128};
  const uint32_t estr[2] = {1, 1};
  CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(qkv), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(DE_ERR_CUDA, "de_attention: cuTensorMapEncodeTiled failed (%d)", (int)r);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttSmem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(attention_tc)");
    attr_set = true;
  }
  dim3 grid((T + AQ - 1) / AQ, heads, B);
  attention_tc_kernel<<<grid, kAttThreads, kAttSmem, reinterpret_cast<cudaStream_t>(stream)>>>(
      tm, reinterpret_cast<__nv_bfloat16*>(out), T, heads, scale * 1.4426950408889634f);
  count_launch();
  return cuda_check_launch("attention_tc_kernel");
}
