// Launch sequences of the three hub models behind /root/reference/depth_midas.py:26 (`self.estimator(input_batch)`),
// over the operator entry points of this library.  Restates:
//   [UPSTREAM] intel-isl/MiDaS v3.0 DPTDepthModel.forward (midas/dpt_depth.py) with forward_vit / forward_flex /
//   ProjectReadout (midas/vit.py) over timm 0.6.7 VisionTransformer / vit_base_resnet50_384 (ResNetV2 3/4/9, StdConv2dSame +
//   GroupNorm(32) + ReLU); MidasNet_small (midas/midas_net_custom.py) over geffnet tf_efficientnet_lite3; the "scratch"
//   decoder of midas/blocks.py (FeatureFusionBlock_custom, ResidualConvUnit_custom) and the two heads.
// Numerics: bf16 GEMM operands, fp32 accumulation (TMEM), fp32 ViT residual stream / LayerNorm / softmax / GroupNorm
// statistics; DPT_Hybrid's ResNetV2 stages (and by default its decoder) as bf16x3 "split3" operands.
// Fusions that change the op graph (all parity-tested against the oracle): the RCU's pre-activation ReLU is a second
// output of the previous conv; `path + RCU1(l)` is folded into RCU1.conv2's epilogue; the 1x1 out_conv runs BEFORE the
// bilinear x2 (linear ops commute, interpolation weights sum to one); the readout Linear(2D->D) is split into
// W_tok x + (W_cls cls + b) broadcast; ConvTranspose(k == s) is a GEMM with a pixel-shuffle store; the ViT residual
// adds are TMA reduce-adds of the proj / fc2 GEMMs into the fp32 stream; the DPT head (up x2, conv3x3+ReLU, conv1x1+ReLU)
// is one launch.
#include <stdio.h>

#include "engine.h"

namespace deb {
namespace eng {

namespace {

struct Gemm : DeGemm {
  Gemm() { memset(static_cast<DeGemm*>(this), 0, sizeof(DeGemm)); }
};
inline Gemm dense(const void* a, int64_t lda, int M, int K, const void* w, int64_t ldw, int N, void* out, int out_dtype, int64_t ldo) {
  Gemm g;
  g.a_mode = 0;
  g.a = a;
  g.lda = lda;
  g.M = M;
  g.K = K;
  g.w = w;
  g.ldw = ldw;
  g.N = N;
  g.out = out;
  g.out_dtype = out_dtype;
  g.ldo = ldo;
  return g;
}
// 3x3 stride-1 pad-1 convolution on an NHWC activation with `cin` stored channels (3C for a split3 operand)
inline Gemm conv3(const void* a, int B, int H, int W, int cin, const void* w, int N, void* out, int out_dtype, int64_t ldo) {
  Gemm g;
  g.a_mode = 1;
  g.a = a;
  g.B = B;
  g.H = H;
  g.W = W;
  g.Cin = cin;
  g.w = w;
  g.ldw = (int64_t)9 * ((cin + 63) / 64 * 64);
  g.N = N;
  g.out = out;
  g.out_dtype = out_dtype;
  g.ldo = ldo;
  return g;
}
// One launch sequence over `B` frames of the plan: the whole batch, or - MiDaS_small - one "lane" of it (a contiguous
// group of frames with its own work buffers, enqueued on its own stream so that the latency-bound kernels of several
// lanes overlap; see forward_small).
struct Ctx {
  Engine& e;
  Plan& p;
  cudaStream_t st;
  int B;                                     // frames of this sequence
  float* pred;                               // its slice of the plan's output
  int lane = 0;
  void* splitk_ws = nullptr;                 // lanes run concurrently: each needs its own K-slice scratch
  int64_t splitk_ws_bytes = 0;
  Ctx(Engine& e_, Plan& p_, cudaStream_t st_) : e(e_), p(p_), st(st_), B(p_.B), pred(p_.pred), splitk_ws(p_.splitk_ws), splitk_ws_bytes(p_.splitk_ws_bytes) {}
  template <typename T = void>
  T* buf(const char* fmt, size_t bytes, int a = 0, int b = 0, int c = 0) {
    char key[112];
    int n = lane ? snprintf(key, sizeof(key), "L%d/", lane) : 0;
    snprintf(key + n, sizeof(key) - n, fmt, a, b, c);
    return reinterpret_cast<T*>(p.ws.get(key, bytes));
  }
};

// every GEMM may use the plan's split-K workspace (de_gemm decides; small grids only)
inline void run(Gemm g, Ctx& c) {
  g.splitk_ws = c.splitk_ws;
  g.splitk_ws_bytes = c.splitk_ws_bytes;
  ENG_CK(de_gemm(&g, c.st));
}

struct Feat {
  void* p = nullptr;
  int H = 0, W = 0, C = 0;                 // C = stored channels per pixel (3x for split3)
};

// ------------------------------------------------------------------------------------------------ ViT
// timm VisionTransformer blocks on the fp32 token stream x [B*T, D]; on_hook(i) after each hooked block
template <typename Hook>
void vit_blocks(Ctx& c, float* x, int B, int T, Hook on_hook) {
  const VitW& v = c.e.vit;
  const int D = v.dim, M = B * T;
  const int kx = v.x3 ? 3 : 1;                  // bf16x3: GEMM operands are split3 rows [hi | hi | lo], K is 3x
  void* xn = c.buf("vit.xn", (size_t)M * kx * D * 2);
  void* qkv = c.buf("vit.qkv", (size_t)M * 3 * D * 2);
  void* att = c.buf("vit.att", (size_t)M * kx * D * 2);
  void* hid = c.buf("vit.hid", (size_t)M * kx * 4 * D * 2);
  const float scale = 1.0f / sqrtf((float)(D / v.heads));
  auto ln = [&](const float* g, const float* b) {
    if (v.x3) ENG_CK(de_layernorm_split3(x, D, g, b, M, D, 1e-6f, xn, 3 * D, c.st));
    else ENG_CK(de_layernorm(x, D, nullptr, 0, 0, g, b, M, D, 1e-6f, xn, D, c.st));
  };
  for (int i = 0; i < v.depth; ++i) {
    const VitBlockW& w = v.blocks[i];
    ln(w.n1w, w.n1b);
    {
      Gemm g = dense(xn, kx * D, M, kx * D, w.qkv_w, kx * D, 3 * D, qkv, DE_DTYPE_BF16, 3 * D);
      g.bias = w.qkv_b;
      run(g, c);
    }
    if (v.x3) ENG_CK(de_attention_split3(qkv, att, B, T, v.heads, scale, c.st));
    else ENG_CK(de_attention(qkv, att, B, T, v.heads, scale, c.st));
    {
      // the branch output is ADDED into the fp32 stream by the GEMM itself (TMA reduce-add epilogue)
      Gemm g = dense(att, kx * D, M, kx * D, w.proj_w, kx * D, D, x, DE_DTYPE_F32, D);
      g.bias = w.proj_b;
      g.accumulate = 1;
      run(g, c);
    }
    ln(w.n2w, w.n2b);
    {
      Gemm g = dense(xn, kx * D, M, kx * D, w.fc1_w, kx * D, 4 * D, hid, DE_DTYPE_BF16, kx * 4 * D);
      g.bias = w.fc1_b;
      g.act = DE_ACT_GELU;
      g.out_split = v.x3 ? 1 : 0;
      run(g, c);
    }
    {
      Gemm g = dense(hid, kx * 4 * D, M, kx * 4 * D, w.fc2_w, kx * 4 * D, D, x, DE_DTYPE_F32, D);
      g.bias = w.fc2_b;
      g.accumulate = 1;
      run(g, c);
    }
    on_hook(i);
  }
}

// [UPSTREAM] act_postprocessN: ProjectReadout -> 1x1 conv -> {ConvT k4s4, ConvT k2s2, id, conv3x3 s2}
Feat reassemble(Ctx& c, const ReasmW& r, int idx, const float* x, int B, int gh, int gw) {
  const int D = r.dim, N = gh * gw, T = N + 1, co = r.cout;
  const int kx = r.x3 ? 3 : 1;                  // bf16x3: every operand row is [hi | hi | lo], every output a split3 row
  void* tok = c.buf("ra%d.tok", (size_t)B * N * kx * D * 2, idx);
  void* cls = c.buf("ra%d.cls", (size_t)B * kx * D * 2, idx);
  if (r.x3) {
    ENG_CK(de_cast_tokens_split3(x, D, B * T, D, T, 1, tok, c.st));
    ENG_CK(de_cast_tokens_split3(x, (int64_t)T * D, B, D, 0, 0, cls, c.st));
  } else {
    ENG_CK(de_cast_tokens(x, D, B * T, D, T, 1, tok, c.st));
    ENG_CK(de_cast_tokens(x, (int64_t)T * D, B, D, 0, 0, cls, c.st));          // row stride T*D: the cls rows
  }
  float* cproj = c.buf<float>("ra%d.cproj", (size_t)B * D * 4, idx);
  {
    Gemm g = r.x3 ? dense(cls, 3 * D, B, 3 * D, r.w_cls3, 3 * D, D, cproj, DE_DTYPE_F32, D)
                  : dense(cls, D, B, D, reinterpret_cast<const uint16_t*>(r.w_read) + D, 2 * D, D, cproj, DE_DTYPE_F32, D);
    g.bias = r.b_read;
    run(g, c);
  }
  void* rr = c.buf("ra%d.r", (size_t)B * N * kx * D * 2, idx);
  {
    Gemm g = r.x3 ? dense(tok, 3 * D, B * N, 3 * D, r.w_tok3, 3 * D, D, rr, DE_DTYPE_BF16, 3 * D)
                  : dense(tok, D, B * N, D, r.w_read, 2 * D, D, rr, DE_DTYPE_BF16, D);
    g.act = DE_ACT_GELU;
    g.add1 = cproj;
    g.add1_dtype = DE_DTYPE_F32;
    g.ld_add1 = D;
    g.add1_rowmode = 1;
    g.add1_pre = 1;
    g.grp = N;
    g.out_split = r.x3 ? 1 : 0;
    run(g, c);
  }
  void* pp = c.buf("ra%d.p", (size_t)B * N * kx * co * 2, idx);
  {
    Gemm g = dense(rr, kx * D, B * N, kx * D, r.w1, kx * D, co, pp, DE_DTYPE_BF16, kx * co);
    g.bias = r.b1;
    g.out_split = r.x3 ? 1 : 0;
    run(g, c);
  }
  Feat f;
  f.C = kx * co;
  if (r.kind == 0) {
    f.p = pp;
    f.H = gh;
    f.W = gw;
    return f;
  }
  if (r.kind == 1) {
    const int s = r.s;
    void* out = c.buf("ra%d.o", (size_t)B * N * s * s * kx * co * 2, idx);
    Gemm g = dense(pp, kx * co, B * N, kx * co, r.w4, kx * co, s * s * co, out, DE_DTYPE_BF16, kx * co);
    g.bias = r.b4;
    g.bias_mod = co;
    g.out_rowmode = 2;
    g.ps_gh = gh;
    g.ps_gw = gw;
    g.ps_s = s;
    g.ps_cout = co;
    g.out_split = r.x3 ? 1 : 0;
    run(g, c);
    f.p = out;
    f.H = gh * s;
    f.W = gw * s;
    return f;
  }
  const int ho = (gh - 1) / 2 + 1, wo = (gw - 1) / 2 + 1;
  void* col = c.buf("ra%d.col", (size_t)B * ho * wo * 9 * kx * co * 2, idx);
  ENG_CK(de_im2col3x3(pp, col, B, gh, gw, kx * co, 2, 1, 1, ho, wo, c.st));
  void* out = c.buf("ra%d.o", (size_t)B * ho * wo * kx * co * 2, idx);
  {
    Gemm g = dense(col, 9 * kx * co, B * ho * wo, 9 * kx * co, r.w4, 9 * kx * co, co, out, DE_DTYPE_BF16, kx * co);
    g.bias = r.b4;
    g.out_split = r.x3 ? 1 : 0;
    run(g, c);
  }
  f.p = out;
  f.H = ho;
  f.W = wo;
  return f;
}

// ------------------------------------------------------------------------------------------------ decoder
// layerN_rn -> refinenet4..1 -> output_conv, NHWC bf16, writes the plan's `pred` [B, 2*H1, 2*W1] (fp32, >= 0)
void decoder_plain(Ctx& c, const Feat* feats) {
  const DecoderW& d = c.e.dec;
  const int B = c.B;
  Feat raw[4], rel[4];
  for (int i = 0; i < 4; ++i) {
    const Feat& x = feats[i];
    const int f = d.feat[i];
    const size_t n = (size_t)B * x.H * x.W * f * 2;
    raw[i] = rel[i] = Feat{nullptr, x.H, x.W, f};
    raw[i].p = c.buf("dec.rn%d", n, i);
    rel[i].p = c.buf("dec.rnr%d", n, i);
    Gemm g = conv3(x.p, B, x.H, x.W, x.C, d.rn_w[i], f, raw[i].p, DE_DTYPE_BF16, f);
    g.out_relu = rel[i].p;
    run(g, c);
  }
  // conv2(relu(conv1(relu(x)))) + x [+ extra]; returns the raw sum, optionally its ReLU copy
  auto rcu = [&](int lvl, int unit, const void* x_raw, const void* x_relu, int H, int W, int C, const void* extra, void** out_relu) {
    const RefineW& r = d.ref[lvl];
    const size_t n = (size_t)B * H * W * C * 2;
    void* t = c.buf("dec.l%d.u%d.t", n, lvl, unit);
    {
      Gemm g = conv3(x_relu, B, H, W, C, r.w[unit][0], C, t, DE_DTYPE_BF16, C);
      g.bias = r.b[unit][0];
      g.act = DE_ACT_RELU;
      run(g, c);
    }
    void* o = c.buf("dec.l%d.u%d.o", n, lvl, unit);
    void* orl = out_relu ? c.buf("dec.l%d.u%d.or", n, lvl, unit) : nullptr;
    {
      Gemm g = conv3(t, B, H, W, C, r.w[unit][1], C, o, DE_DTYPE_BF16, C);
      g.bias = r.b[unit][1];
      g.add1 = x_raw;
      g.add1_dtype = DE_DTYPE_BF16;
      g.ld_add1 = C;
      g.add2 = extra;
      g.add2_dtype = DE_DTYPE_BF16;
      g.out_relu = orl;
      run(g, c);
    }
    if (out_relu) *out_relu = orl;
    return o;
  };
  Feat path;
  for (int i = 3; i >= 0; --i) {
    const RefineW& r = d.ref[i];
    const int H = raw[i].H, W = raw[i].W, C = raw[i].C;
    const void *s_raw = raw[i].p, *s_relu = rel[i].p;
    if (path.p) {
      void* srl = nullptr;
      s_raw = rcu(i, 0, raw[i].p, rel[i].p, H, W, C, path.p, &srl);
      s_relu = srl;
    }
    void* u = rcu(i, 1, s_raw, s_relu, H, W, C, nullptr, nullptr);
    const int co = r.cout;
    void* v = c.buf("dec.l%d.oc", (size_t)B * H * W * co * 2, i);
    {
      Gemm g = dense(u, C, B * H * W, C, r.out_w, C, co, v, DE_DTYPE_BF16, co);
      g.bias = r.out_b;
      run(g, c);
    }
    path = Feat{c.buf("dec.l%d.up", (size_t)B * 4 * H * W * co * 2, i), 2 * H, 2 * W, co};
    ENG_CK(de_upsample2x(v, path.p, B, H, W, co, 1, c.st));
  }
  const int H = path.H, W = path.W;
  void* h0 = c.buf("dec.h0", (size_t)B * H * W * d.h0_c * 2);
  {
    Gemm g = conv3(path.p, B, H, W, path.C, d.h0_w, d.h0_c, h0, DE_DTYPE_BF16, d.h0_c);
    g.bias = d.h0_b;
    run(g, c);
  }
  if (d.h0_c % 64 == 0 && d.h0_c <= 128) {      // DPT: upsample + head tail in one halo-tiled kernel
    ENG_CK(de_head_up_conv(h0, d.h2_w, (int64_t)9 * d.h0_c, d.h2_b, d.h4_w, d.h4_b, c.pred, B, H, W, d.h0_c, d.head_ac ? 1 : 0, c.st));
  } else {                                      // MiDaS_small (32 channels): general implicit-GEMM kernel with the head-dot epilogue
    void* h1 = c.buf("dec.h1", (size_t)B * 4 * H * W * d.h0_c * 2);
    ENG_CK(de_upsample2x(h0, h1, B, H, W, d.h0_c, d.head_ac ? 1 : 0, c.st));
    Gemm g = conv3(h1, B, 2 * H, 2 * W, d.h0_c, d.h2_w, 32, c.pred, DE_DTYPE_F32, 1);
    g.bias = d.h2_b;
    g.act = DE_ACT_RELU;
    g.head_w = d.h4_w;
    g.head_b = d.h4_b;
    run(g, c);
  }
}

// The same decoder at bf16x3 precision.  Conv operands are split3 bf16 tensors [B,H,W,3C] against [Wh | Wl | Wh]
// filters; everything a skip connection reads (layerN_rn outputs, RCU sums, the upsampled path) stays fp32; the
// pre-activation ReLU copies are produced as split3 second outputs.  feats[i].C is the stored channel count (3*Ci
// for the levels that arrive as split3 operands, Ci for plain bf16).
void decoder_split3(Ctx& c, const Feat* feats) {
  const DecoderW& d = c.e.dec;
  const int B = c.B;
  Feat raw[4], rel[4];
  for (int i = 0; i < 4; ++i) {
    const Feat& x = feats[i];
    const int f = d.feat[i];
    const size_t px = (size_t)B * x.H * x.W;
    raw[i] = Feat{c.buf("dec3.rn%d", px * f * 4, i), x.H, x.W, f};
    rel[i] = Feat{c.buf("dec3.rnr%d", px * 3 * f * 2, i), x.H, x.W, 3 * f};
    Gemm g = conv3(x.p, B, x.H, x.W, x.C, d.rn_w[i], f, raw[i].p, DE_DTYPE_F32, f);
    g.out_relu = rel[i].p;
    g.relu_split = 1;
    g.ld_relu = 3 * f;
    run(g, c);
  }
  // last == false: (fp32 sum, split3 relu(sum)); last == true: the sum as a split3 operand
  auto rcu = [&](int lvl, int unit, const void* x_raw, const void* x_relu, int H, int W, int C, const void* extra, bool last,
                 void** out_relu) {
    const RefineW& r = d.ref[lvl];
    const size_t px = (size_t)B * H * W;
    void* t = c.buf("dec3.l%d.u%d.t", px * 3 * C * 2, lvl, unit);
    {
      Gemm g = conv3(x_relu, B, H, W, 3 * C, r.w[unit][0], C, t, DE_DTYPE_BF16, 3 * C);
      g.bias = r.b[unit][0];
      g.act = DE_ACT_RELU;
      g.out_split = 1;
      run(g, c);
    }
    Gemm g = conv3(t, B, H, W, 3 * C, r.w[unit][1], C, nullptr, 0, 0);
    g.bias = r.b[unit][1];
    g.add1 = x_raw;
    g.add1_dtype = DE_DTYPE_F32;
    g.ld_add1 = C;
    void* o;
    if (last) {
      o = c.buf("dec3.l%d.u%d.o3", px * 3 * C * 2, lvl, unit);
      g.out = o;
      g.out_dtype = DE_DTYPE_BF16;
      g.ldo = 3 * C;
      g.out_split = 1;
    } else {
      o = c.buf("dec3.l%d.u%d.o", px * C * 4, lvl, unit);
      void* orl = c.buf("dec3.l%d.u%d.or", px * 3 * C * 2, lvl, unit);
      g.out = o;
      g.out_dtype = DE_DTYPE_F32;
      g.ldo = C;
      g.add2 = extra;
      g.add2_dtype = DE_DTYPE_F32;
      g.out_relu = orl;
      g.relu_split = 1;
      g.ld_relu = 3 * C;
      *out_relu = orl;
    }
    run(g, c);
    return o;
  };
  Feat path;
  for (int i = 3; i >= 0; --i) {
    const RefineW& r = d.ref[i];
    const int H = raw[i].H, W = raw[i].W, C = raw[i].C;
    const void *s_raw = raw[i].p, *s_relu = rel[i].p;
    if (path.p) {
      void* srl = nullptr;
      s_raw = rcu(i, 0, raw[i].p, rel[i].p, H, W, C, path.p, false, &srl);
      s_relu = srl;
    }
    void* u = rcu(i, 1, s_raw, s_relu, H, W, C, nullptr, true, nullptr);
    const int co = r.cout;
    float* v = c.buf<float>("dec3.l%d.oc", (size_t)B * H * W * co * 4, i);
    {
      Gemm g = dense(u, 3 * C, B * H * W, 3 * C, r.out_w, 3 * C, co, v, DE_DTYPE_F32, co);
      g.bias = r.out_b;
      run(g, c);
    }
    if (i > 0) {
      path = Feat{c.buf("dec3.l%d.up", (size_t)B * 4 * H * W * co * 4, i), 2 * H, 2 * W, co};
      ENG_CK(de_upsample2x_ex(v, DE_DTYPE_F32, path.p, 2, B, H, W, co, 1, c.st));
    } else {
      path = Feat{c.buf("dec3.l%d.up", (size_t)B * 4 * H * W * 3 * co * 2, i), 2 * H, 2 * W, 3 * co};
      ENG_CK(de_upsample2x_ex(v, DE_DTYPE_F32, path.p, 1, B, H, W, co, 1, c.st));
    }
  }
  const int H = path.H, W = path.W;
  float* h0 = c.buf<float>("dec3.h0", (size_t)B * H * W * d.h0_c * 4);
  {
    Gemm g = conv3(path.p, B, H, W, path.C, d.h0_w, d.h0_c, h0, DE_DTYPE_F32, d.h0_c);
    g.bias = d.h0_b;
    run(g, c);
  }
  void* h1 = c.buf("dec3.h1", (size_t)B * 4 * H * W * 3 * d.h0_c * 2);
  ENG_CK(de_upsample2x_ex(h0, DE_DTYPE_F32, h1, 1, B, H, W, d.h0_c, d.head_ac ? 1 : 0, c.st));
  Gemm g = conv3(h1, B, 2 * H, 2 * W, 3 * d.h0_c, d.h2_w, 32, c.pred, DE_DTYPE_F32, 1);
  g.bias = d.h2_b;
  g.act = DE_ACT_RELU;
  g.head_w = d.h4_w;
  g.head_b = d.h4_b;
  run(g, c);
}

void decoder(Ctx& c, const Feat* feats) {
  if (c.e.dec.split3) decoder_split3(c, feats);
  else decoder_plain(c, feats);
}

int same_pad_before(int n, int k, int s) {
  const int out = (n + s - 1) / s;
  int total = (out - 1) * s + k - n;
  if (total < 0) total = 0;
  return total / 2;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ DPT_Large
// ViT-L/16 (24 blocks, D = 1024, 16 heads, hooks 5/11/17/23) + reassemble + decoder.  x_in: the bf16 patch matrix
// [B*gh*gw, 768] K1 writes (layout 1).
void forward_large(Engine& e, Plan& p, cudaStream_t st) {
  Ctx c(e, p, st);
  const int B = p.B, gh = p.net_h / 16, gw = p.net_w / 16, N = gh * gw, D = e.vit.dim, T = N + 1;
  float* x = c.buf<float>("x", (size_t)B * T * D * 4);
  float* pos = e.pos_embed(gh, gw);
  {
    const void* patches = p.x_in;
    int kp = 768;
    if (e.x3) {                                   // K1 wrote fp32 NCHW: re-tile it into the split3 patch matrix
      void* p3 = c.buf("patch3", (size_t)B * N * 3 * 768 * 2);
      ENG_CK(de_patchify_split3(reinterpret_cast<const float*>(p.x_in), p3, B, p.net_h, p.net_w, st));
      patches = p3;
      kp = 3 * 768;
    }
    Gemm g = dense(patches, kp, B * N, kp, e.pe_w, kp, D, x, DE_DTYPE_F32, D);
    g.bias = e.pe_b;
    g.add1 = pos;
    g.add1_dtype = DE_DTYPE_F32;
    g.ld_add1 = D;
    g.add1_rowmode = 2;
    g.out_rowmode = 1;
    g.grp = N;
    g.pad = 1;
    g.off = 1;
    run(g, c);
  }
  ENG_CK(de_fill_cls(x, D, B, T, D, e.vit.cls, pos, st));
  Feat feats[4];
  vit_blocks(c, x, B, T, [&](int i) {
    const int hooks[4] = {5, 11, 17, 23};
    for (int j = 0; j < 4; ++j)
      if (hooks[j] == i) feats[j] = reassemble(c, e.reasm[j], j, x, B, gh, gw);
  });
  decoder(c, feats);
}

// ------------------------------------------------------------------------------------------------ DPT_Hybrid
namespace {

struct HybridRun {
  Ctx& c;
  void* gn_ws;
  int64_t gn_ws_bytes;
  int B;

  // x: fp32 conv output [B,H,W,C] -> split3 bf16 [B,H,W,3C] (+ fp32 copy, + plain bf16 copy)
  void gn(const float* x, const float* gamma, const float* beta, int HW, int C, int act, const float* add, void* out3,
          float* out32, void* out_plain) {
    ENG_CK(de_groupnorm(x, DE_DTYPE_F32, gamma, beta, add, DE_DTYPE_F32, out3, 1, out32, out_plain, B, HW, C, 32, 1e-5f, act,
                        gn_ws, gn_ws_bytes, c.st));
  }
};

}  // namespace

void forward_hybrid(Engine& e, Plan& p, cudaStream_t st) {
  Ctx c(e, p, st);
  const HybridW& hw = e.hyb;
  const int B = p.B;
  int H = (p.net_h + 1) / 2, W = (p.net_w + 1) / 2;
  // GroupNorm workspace: an upper bound over every call of this geometry (16 KB of arrival counters, which start at
  // zero, + stats + at most ~148 partial-sum blocks per sample: de_groupnorm_ws_bytes)
  const int64_t need = 16384 + 256LL * B + 16LL * B * 160 * 32 + 4096;
  HybridRun hr{c, p.ws.get("gn.ws", (size_t)need, /*zero=*/true), need, B};

  // stem: StdConv 7x7/2 (TF-same) as split3 im2col rows x [Wh | Wl | Wh] -> GroupNorm + ReLU -> MaxPool 3x3/2 (TF-same)
  float* s0 = c.buf<float>("stem", (size_t)B * H * W * 64 * 4);
  {
    void* col = c.buf("stemcol", (size_t)B * H * W * 3 * 152 * 2);
    ENG_CK(de_stem_im2col_split3(reinterpret_cast<const float*>(p.x_in), col, B, p.net_h, p.net_w, 7, 2, 152, st));
    Gemm g = dense(col, 3 * 152, B * H * W, 3 * 152, hw.stem_w3, 3 * 152, 64, s0, DE_DTYPE_F32, 64);
    run(g, c);
  }
  float* s1 = c.buf<float>("stemn", (size_t)B * H * W * 64 * 4);
  ENG_CK(de_groupnorm(s0, DE_DTYPE_F32, hw.stem_g, hw.stem_b, nullptr, 0, nullptr, 0, s1, nullptr, B, H * W, 64, 32, 1e-5f,
                      DE_ACT_RELU, hr.gn_ws, hr.gn_ws_bytes, st));
  const int Hp = (H + 1) / 2, Wp = (W + 1) / 2;
  void* x = c.buf("pool", (size_t)B * Hp * Wp * 3 * 64 * 2);
  ENG_CK(de_maxpool3x3s2(s1, DE_DTYPE_F32, x, 1, B, H, W, 64, st));
  H = Hp;
  W = Wp;
  const float* x32 = nullptr;
  Feat feats[4];
  for (int si = 0; si < 3; ++si) {
    const auto& blocks = hw.stages[si];
    for (int bi = 0; bi < (int)blocks.size(); ++bi) {
      const BottleneckW& k = blocks[bi];
      const int s = k.stride, mid = k.mid, cout = k.cout, cin = k.cin;
      const int Ho = (H + s - 1) / s, Wo = (W + s - 1) / s;
      const size_t pin = (size_t)B * H * W, pout = (size_t)B * Ho * Wo;
      const float* shortcut = x32;
      if (k.has_proj) {
        const void* xs = x;
        if (s > 1) {
          void* sub = c.buf("s%d.b%d.sub", pout * 3 * cin * 2, si, bi);
          ENG_CK(de_subsample(x, sub, B, H, W, 3 * cin, s, st));
          xs = sub;
        }
        float* dconv = c.buf<float>("s%d.b%d.dc", pout * cout * 4, si, bi);
        run(dense(xs, 3 * cin, (int)pout, 3 * cin, k.wd, 3 * cin, cout, dconv, DE_DTYPE_F32, cout), c);
        float* sc = c.buf<float>("s%d.b%d.dn", pout * cout * 4, si, bi);
        ENG_CK(de_groupnorm(dconv, DE_DTYPE_F32, k.ndg, k.ndb, nullptr, 0, nullptr, 0, sc, nullptr, B, Ho * Wo, cout, 32, 1e-5f,
                            DE_ACT_NONE, hr.gn_ws, hr.gn_ws_bytes, st));
        shortcut = sc;
      }
      float* t = c.buf<float>("s%d.b%d.c1", pin * mid * 4, si, bi);
      run(dense(x, 3 * cin, (int)pin, 3 * cin, k.w1, 3 * cin, mid, t, DE_DTYPE_F32, mid), c);
      void* t3 = c.buf("s%d.b%d.n1", pin * 3 * mid * 2, si, bi);
      hr.gn(t, k.ng[0], k.nb[0], H * W, mid, DE_ACT_RELU, nullptr, t3, nullptr, nullptr);
      float* u = c.buf<float>("s%d.b%d.c2", pout * mid * 4, si, bi);
      if (s == 1) {
        run(conv3(t3, B, H, W, 3 * mid, k.w2, mid, u, DE_DTYPE_F32, mid), c);
      } else {
        void* col = c.buf("s%d.b%d.col", pout * 27 * mid * 2, si, bi);
        ENG_CK(de_im2col3x3(t3, col, B, H, W, 3 * mid, s, same_pad_before(H, 3, s), same_pad_before(W, 3, s), Ho, Wo, st));
        run(dense(col, 27 * mid, (int)pout, 27 * mid, k.w2, 27 * mid, mid, u, DE_DTYPE_F32, mid), c);
      }
      void* u3 = c.buf("s%d.b%d.n2", pout * 3 * mid * 2, si, bi);
      hr.gn(u, k.ng[1], k.nb[1], Ho * Wo, mid, DE_ACT_RELU, nullptr, u3, nullptr, nullptr);
      float* v = c.buf<float>("s%d.b%d.c3", pout * cout * 4, si, bi);
      run(dense(u3, 3 * mid, (int)pout, 3 * mid, k.w3, 3 * mid, cout, v, DE_DTYPE_F32, cout), c);
      void* y3 = c.buf("s%d.b%d.n3", pout * 3 * cout * 2, si, bi);
      float* y32 = c.buf<float>("s%d.b%d.n3f", pout * cout * 4, si, bi);
      // stages 0 / 1 also feed the decoder: as the split3 operand they already are, or as a plain bf16 copy
      const bool last = bi == (int)blocks.size() - 1 && si < 2;
      void* ypl = (last && !hw.decoder_split3) ? c.buf("s%d.b%d.n3p", pout * cout * 2, si, bi) : nullptr;
      hr.gn(v, k.ng[2], k.nb[2], Ho * Wo, cout, DE_ACT_RELU, shortcut, y3, y32, ypl);
      x = y3;
      x32 = y32;
      H = Ho;
      W = Wo;
      if (last) feats[si] = hw.decoder_split3 ? Feat{y3, H, W, 3 * cout} : Feat{ypl, H, W, cout};
    }
  }
  const int gh = p.net_h / 16, gw = p.net_w / 16;
  if (H != gh || W != gw) raise(DE_ERR_INVALID, "DPT_Hybrid: backbone grid %dx%d != token grid %dx%d", H, W, gh, gw);
  const int N = gh * gw, D = e.vit.dim, T = N + 1;
  float* tok = c.buf<float>("x", (size_t)B * T * D * 4);
  float* pos = e.pos_embed(gh, gw);
  {
    Gemm g = dense(x, 3 * 1024, B * N, 3 * 1024, hw.pe_w, 3 * 1024, D, tok, DE_DTYPE_F32, D);
    g.bias = hw.pe_b;
    g.add1 = pos;
    g.add1_dtype = DE_DTYPE_F32;
    g.ld_add1 = D;
    g.add1_rowmode = 2;
    g.out_rowmode = 1;
    g.grp = N;
    g.pad = 1;
    g.off = 1;
    run(g, c);
  }
  ENG_CK(de_fill_cls(tok, D, B, T, D, e.vit.cls, pos, st));
  vit_blocks(c, tok, B, T, [&](int i) {
    if (i == 8) feats[2] = reassemble(c, e.reasm[0], 3, tok, B, gh, gw);
    if (i == 11) feats[3] = reassemble(c, e.reasm[1], 4, tok, B, gh, gw);
  });
  decoder(c, feats);
}

// ------------------------------------------------------------------------------------------------ MiDaS_small
// tf_efficientnet_lite3 (BatchNorm folded, ReLU6, TF-same padding, no SE) + expand decoder 64/128/256/512.
// x_in: bf16 NCHW [B,3,net_h,net_w] (K1 layout 0).  1x1 convs on the tcgen05 GEMM with bias / ReLU6 / residual
// epilogues; depthwise convs on the TMA-staged bandwidth kernel; the 3-channel stem on the direct kernel.
// DEB_MBCONV_FUSED=1 runs the stride-1 blocks' depthwise + projection as ONE kernel (de_dwproject, csrc/mbconv.cu).
// Off by default - MEASURED SLOWER on B200 (32 frames 720p: 1.70 against 1.46 ms; 640x480 batch 1: 0.91 against 0.68 ms):
// with ~190 KB of shared memory per CTA the fused kernel runs one CTA of 8 compute warps per SM where the stand-alone
// depthwise kernel runs three, and at 8x16 / 4x8 pixels per frame the few output tiles leave the depthwise arithmetic
// on a quarter of the warps; the launch it saves (3-6 us) does not pay for that (profiles/r02_mbconv_fused.md).
static bool small_fused() {
  static int on = -1;
  if (on < 0) {
    const char* v = getenv("DEB_MBCONV_FUSED");
    on = (v && v[0] == '1') ? 1 : 0;
  }
  return on != 0;
}

// MiDaS_small's launch sequence over the frames of one Ctx (the whole batch or one lane), network input at `x_in`
static void small_sequence(Ctx& c, const void* x_in) {
  Engine& e = c.e;
  Plan& p = c.p;
  cudaStream_t st = c.st;
  const int B = c.B;
  int H = (p.net_h + 1) / 2, W = (p.net_w + 1) / 2;
  void* x = c.buf("stem", (size_t)B * H * W * 32 * 2);
  ENG_CK(de_stem_conv(x_in, DE_DTYPE_BF16, e.small.stem_w, e.small.stem_b, x, DE_DTYPE_BF16, B, p.net_h, p.net_w, 32, 3, 2,
                      DE_ACT_RELU6, st));
  Feat feats[4];
  int bi = 0;
  for (const MbBlockW& m : e.small.blocks) {
    const int s = m.stride, Ho = (H + s - 1) / s, Wo = (W + s - 1) / s;
    const size_t pin = (size_t)B * H * W, pout = (size_t)B * Ho * Wo;
    const void* t = x;
    if (m.ir) {
      void* ex = c.buf("mb%d.e", pin * m.mid * 2, bi);
      Gemm g = dense(x, m.cin, (int)pin, m.cin, m.exp_w, m.cin, m.mid, ex, DE_DTYPE_BF16, m.mid);
      g.bias = m.exp_b;
      g.act = DE_ACT_RELU6;
      run(g, c);
      t = ex;
    }
    void* y = c.buf("mb%d.y", pout * m.cout * 2, bi);
    if (s == 1 && small_fused()) {
      // depthwise + projection in one kernel (csrc/mbconv.cu): the 6x-expanded depthwise output stays on the SM
      const int64_t wsb = de_dwproject_ws_bytes(B, H, W, m.cout);
      void* ws = c.p.ws.get(std::string(c.lane ? "L" + std::to_string(c.lane) + "/" : "") + "mb.ws" + std::to_string(bi), (size_t)wsb, true);
      ENG_CK(de_dwproject(t, m.dw_w, m.dw_b, m.pw_w, m.pw_b, m.residual ? x : nullptr, y, B, H, W, m.mid, m.cout, m.k, ws, wsb, st));
      x = y;
      H = Ho;
      W = Wo;
      if (m.last) {
        const int lvl = m.stage == 1 ? 0 : m.stage == 2 ? 1 : m.stage == 4 ? 2 : m.stage == 6 ? 3 : -1;
        if (lvl >= 0) feats[lvl] = Feat{x, H, W, m.cout};
      }
      ++bi;
      continue;
    }
    void* dwo = c.buf("mb%d.d", pout * m.mid * 2, bi);
    ENG_CK(de_depthwise(t, m.dw_w, m.dw_b, dwo, B, H, W, m.mid, m.k, s, DE_ACT_RELU6, st));
    {
      Gemm g = dense(dwo, m.mid, (int)pout, m.mid, m.pw_w, m.mid, m.cout, y, DE_DTYPE_BF16, m.cout);
      g.bias = m.pw_b;
      if (m.residual) {
        g.add1 = x;
        g.add1_dtype = DE_DTYPE_BF16;
        g.ld_add1 = m.cin;
      }
      run(g, c);
    }
    x = y;
    H = Ho;
    W = Wo;
    if (m.last) {
      const int lvl = m.stage == 1 ? 0 : m.stage == 2 ? 1 : m.stage == 4 ? 2 : m.stage == 6 ? 3 : -1;
      if (lvl >= 0) feats[lvl] = Feat{x, H, W, m.cout};
    }
    ++bi;
  }
  decoder(c, feats);
}

// Optional (DEB_SMALL_LANES=2..4, default 1 = off): cut the batch into equal lanes of >= 4 frames, each with its own
// buffers and stream, forked from / joined to `st` with events (parallel branches inside a captured graph), so that one
// lane's latency-bound kernels could fill the bubbles of another's.  MEASURED NEGATIVE on B200 (32 frames 720p:
// 1.64 ms with 1 lane, 1.73 with 2, 1.80 with 4): every GEMM CTA holds a whole SM (227 KB of shared memory), so kernels
// of different lanes do not co-reside, PDL already overlaps each kernel's prologue with its predecessor's tail, and the
// smaller per-lane grids lose more than the overlap gains.  Kept as a switch for other batch sizes / future kernels.
static int small_lanes(int B) {
  static int cap = -1;
  if (cap < 0) {
    const char* v = getenv("DEB_SMALL_LANES");
    cap = v ? atoi(v) : 1;
    if (cap < 1) cap = 1;
    if (cap > Engine::kMaxLanes) cap = Engine::kMaxLanes;
  }
  for (int n = cap; n > 1; --n)
    if (B % n == 0 && B / n >= 4) return n;
  return 1;
}

void forward_small(Engine& e, Plan& p, cudaStream_t st) {
  const int lanes = small_lanes(p.B);
  if (lanes == 1) {
    Ctx c(e, p, st);
    small_sequence(c, p.x_in);
    return;
  }
  const int Bl = p.B / lanes;
  const size_t in_stride = (size_t)Bl * 3 * p.net_h * p.net_w * 2;            // K1 layout 0: bf16 NCHW
  const size_t out_stride = (size_t)Bl * p.net_h * p.net_w;
  e.ensure_lanes(lanes);
  ENG_CUDA(cudaEventRecord(e.lane_fork, st));
  for (int l = 0; l < lanes; ++l) {
    cudaStream_t ls = l == 0 ? st : e.lane_stream[l];
    if (l > 0) ENG_CUDA(cudaStreamWaitEvent(ls, e.lane_fork, 0));
    Ctx c(e, p, ls);
    c.B = Bl;
    c.pred = p.pred + l * out_stride;
    c.lane = l;
    if (l > 0) {
      c.splitk_ws_bytes = 16 << 20;
      c.splitk_ws = c.buf("splitk", (size_t)c.splitk_ws_bytes);
    }
    small_sequence(c, reinterpret_cast<const uint8_t*>(p.x_in) + l * in_stride);
    if (l > 0) {
      ENG_CUDA(cudaEventRecord(e.lane_join[l], ls));
      ENG_CUDA(cudaStreamWaitEvent(st, e.lane_join[l], 0));
    }
  }
}

}  // namespace eng
}  // namespace deb
