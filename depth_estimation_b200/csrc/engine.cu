// Handle-level C ABI: de_create / de_set_weight / de_finalize_weights / de_plan / de_predict /
// de_forward / de_destroy.  Replaces, for a non-Python host, what `DepthEstimator.__init__` and `.predict` do in
// /root/reference/depth_midas.py:7-35: load the model's weights (here: take them in upstream state_dict naming, fold
// BatchNorm / weight standardisation and repack them once), then per frame transform -> network -> bicubic resize.
// This file: weight ingestion + packing, plans, CUDA-graph capture and the entry points.  engine_nets.cu: the launch
// sequences of the three hub models.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>

#include <mutex>

#include "engine.h"

namespace deb {
void add_launches(int64_t n);   // api.cu: replayed-graph launches join the library's launch counter
namespace eng {

void raise(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  throw EngineError{code, buf};
}

// ------------------------------------------------------------------------------------------------ packing
BfVec pack_plain(const float* w, int64_t n) {
  BfVec o((size_t)n);
  for (int64_t i = 0; i < n; ++i) o[i] = f2bf(w[i]);
  return o;
}
static int64_t ceil64(int64_t c) { return (c + 63) / 64 * 64; }

BfVec pack_conv3x3(const HostT& w) {
  const int64_t co = w.dim(0), ci = w.dim(1), cp = ceil64(ci);
  BfVec o((size_t)(co * 9 * cp), 0);
  for (int64_t n = 0; n < co; ++n)
    for (int64_t c = 0; c < ci; ++c)
      for (int t = 0; t < 9; ++t) o[(n * 9 + t) * cp + c] = f2bf(w.v[(n * ci + c) * 9 + t]);
  return o;
}
BfVec pack_conv3x3_flat(const HostT& w) {
  const int64_t co = w.dim(0), ci = w.dim(1);
  BfVec o((size_t)(co * 9 * ci));
  for (int64_t n = 0; n < co; ++n)
    for (int64_t c = 0; c < ci; ++c)
      for (int t = 0; t < 9; ++t) o[(n * 9 + t) * ci + c] = f2bf(w.v[(n * ci + c) * 9 + t]);
  return o;
}
BfVec pack_conv_transpose(const HostT& w) {
  const int64_t ci = w.dim(0), co = w.dim(1), s = w.dim(2);
  BfVec o((size_t)(s * s * co * ci));
  for (int64_t c = 0; c < ci; ++c)
    for (int64_t n = 0; n < co; ++n)
      for (int64_t i = 0; i < s; ++i)
        for (int64_t j = 0; j < s; ++j) o[((i * s + j) * co + n) * ci + c] = f2bf(w.v[((c * co + n) * s + i) * s + j]);
  return o;
}
FVec conv_transpose_rows(const HostT& w) {
  const int64_t ci = w.dim(0), co = w.dim(1), s = w.dim(2);
  FVec o((size_t)(s * s * co * ci));
  for (int64_t c = 0; c < ci; ++c)
    for (int64_t n = 0; n < co; ++n)
      for (int64_t i = 0; i < s; ++i)
        for (int64_t j = 0; j < s; ++j) o[((i * s + j) * co + n) * ci + c] = w.v[((c * co + n) * s + i) * s + j];
  return o;
}
BfVec split3_rows(const float* w, int64_t rows, int64_t cols) {
  BfVec o((size_t)(rows * 3 * cols));
  for (int64_t r = 0; r < rows; ++r)
    for (int64_t c = 0; c < cols; ++c) {
      const float v = w[r * cols + c];
      const uint16_t hi = f2bf(v), lo = f2bf(v - bf2f(hi));
      o[r * 3 * cols + c] = hi;
      o[r * 3 * cols + cols + c] = lo;
      o[r * 3 * cols + 2 * cols + c] = hi;
    }
  return o;
}
BfVec split3_conv3x3(const float* w, int64_t co, int64_t ci) {
  if ((3 * ci) % 64 != 0) raise(DE_ERR_UNSUPPORTED, "split3 conv3x3: 3*Cin=%lld is not a multiple of 64", (long long)(3 * ci));
  BfVec o((size_t)(co * 27 * ci));
  for (int64_t n = 0; n < co; ++n)
    for (int64_t c = 0; c < ci; ++c)
      for (int t = 0; t < 9; ++t) {
        const float v = w[(n * ci + c) * 9 + t];
        const uint16_t hi = f2bf(v), lo = f2bf(v - bf2f(hi));
        uint16_t* row = &o[(n * 9 + t) * 3 * ci];
        row[c] = hi;
        row[ci + c] = lo;
        row[2 * ci + c] = hi;
      }
  return o;
}
void fold_bn(FVec& w, int64_t cout, const FVec& g, const FVec& b, const FVec& mean, const FVec& var, float eps, FVec& bias_out) {
  const int64_t per = (int64_t)w.size() / cout;
  bias_out.resize((size_t)cout);
  for (int64_t n = 0; n < cout; ++n) {
    const float scale = g[n] / sqrtf(var[n] + eps);
    for (int64_t i = 0; i < per; ++i) w[n * per + i] *= scale;
    bias_out[n] = b[n] - mean[n] * scale;
  }
}
FVec standardize_weight(const HostT& w, double eps) {
  const int64_t co = w.dim(0), per = w.numel() / co;
  FVec o((size_t)w.numel());
  for (int64_t n = 0; n < co; ++n) {
    double m = 0.0;
    for (int64_t i = 0; i < per; ++i) m += (double)w.v[n * per + i];
    m /= (double)per;
    double var = 0.0;
    for (int64_t i = 0; i < per; ++i) {
      const double d = (double)w.v[n * per + i] - m;
      var += d * d;
    }
    var /= (double)per;                                       // biased, as torch.var(unbiased=False)
    const double inv = 1.0 / sqrt(var + eps);
    for (int64_t i = 0; i < per; ++i) o[n * per + i] = (float)(((double)w.v[n * per + i] - m) * inv);
  }
  return o;
}

// ------------------------------------------------------------------------------------------------ device memory
void* DevPool::alloc(size_t n) {
  void* p = nullptr;
  if (n == 0) n = 16;
  ENG_CUDA(cudaMalloc(&p, n));
  allocs.push_back(p);
  bytes += n;
  return p;
}
void* DevPool::upload(const void* host, size_t n) {
  void* p = alloc(n);
  ENG_CUDA(cudaMemcpy(p, host, n, cudaMemcpyHostToDevice));
  return p;
}
void DevPool::release() {
  for (void* p : allocs) cudaFree(p);
  allocs.clear();
  bytes = 0;
}
void* WorkSet::get(const std::string& key, size_t bytes, bool zero) {
  auto it = bufs.find(key);
  if (it != bufs.end()) {
    if (it->second.second != bytes) raise(DE_ERR_INVALID, "work buffer %s: size changed (%zu -> %zu)", key.c_str(), it->second.second, bytes);
    return it->second.first;
  }
  void* p = pool.alloc(bytes);
  if (zero) ENG_CUDA(cudaMemset(p, 0, bytes));
  bufs.emplace(key, std::make_pair(p, bytes));
  return p;
}
Plan::~Plan() {
  for (auto& kv : graphs)
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
}

void Engine::ensure_lanes(int n) {
  if (!lane_fork) ENG_CUDA(cudaEventCreateWithFlags(&lane_fork, cudaEventDisableTiming));
  for (int i = 1; i < n && i < kMaxLanes; ++i) {
    if (!lane_stream[i]) ENG_CUDA(cudaStreamCreateWithFlags(&lane_stream[i], cudaStreamNonBlocking));
    if (!lane_join[i]) ENG_CUDA(cudaEventCreateWithFlags(&lane_join[i], cudaEventDisableTiming));
  }
}

// ------------------------------------------------------------------------------------------------ weights
const HostT& Engine::W(const std::string& key, std::initializer_list<int64_t> shape) {
  auto it = raw.find(key);
  if (it == raw.end()) raise(DE_ERR_INVALID, "weight %s was never set (de_set_weight)", key.c_str());
  const HostT& t = it->second;
  if (t.shape != std::vector<int64_t>(shape)) {
    std::string got, want;
    for (auto d : t.shape) got += std::to_string(d) + ",";
    for (auto d : shape) want += std::to_string(d) + ",";
    raise(DE_ERR_INVALID, "weight %s has shape [%s], expected [%s]", key.c_str(), got.c_str(), want.c_str());
  }
  return t;
}
float* Engine::f32(const std::string& key, std::initializer_list<int64_t> shape) { return wpool.put(W(key, shape).v); }

void Engine::build_vit(const std::string& pre, int dim, int depth, int heads, bool x3v) {
  vit.dim = dim;
  vit.depth = depth;
  vit.heads = heads;
  vit.x3 = x3v;
  auto lin = [&](const HostT& w, int64_t n, int64_t k) -> void* {
    return x3v ? (void*)wpool.put(split3_rows(w.v.data(), n, k)) : (void*)wpool.put(pack_plain(w.v.data(), n * k));
  };
  vit.cls = f32(pre + "cls_token", {1, 1, dim});
  vit.pos_raw = W(pre + "pos_embed", {1, 577, dim});
  const int64_t D = dim;
  for (int i = 0; i < depth; ++i) {
    const std::string p = pre + "blocks." + std::to_string(i) + ".";
    VitBlockW b{};
    b.n1w = f32(p + "norm1.weight", {D});
    b.n1b = f32(p + "norm1.bias", {D});
    b.qkv_w = lin(W(p + "attn.qkv.weight", {3 * D, D}), 3 * D, D);
    b.qkv_b = f32(p + "attn.qkv.bias", {3 * D});
    b.proj_w = lin(W(p + "attn.proj.weight", {D, D}), D, D);
    b.proj_b = f32(p + "attn.proj.bias", {D});
    b.n2w = f32(p + "norm2.weight", {D});
    b.n2b = f32(p + "norm2.bias", {D});
    b.fc1_w = lin(W(p + "mlp.fc1.weight", {4 * D, D}), 4 * D, D);
    b.fc1_b = f32(p + "mlp.fc1.bias", {4 * D});
    b.fc2_w = lin(W(p + "mlp.fc2.weight", {D, 4 * D}), D, 4 * D);
    b.fc2_b = f32(p + "mlp.fc2.bias", {D});
    vit.blocks.push_back(b);
  }
}

// [UPSTREAM] vit._resize_pos_embed: grid part bilinearly resized (align_corners=False, ATen's fp32 coordinate rule),
// cls part kept.  One-time weight preparation per token grid, not part of the per-frame path.
float* Engine::pos_embed(int gh, int gw) {
  auto key = std::make_pair(gh, gw);
  auto it = vit.pos.find(key);
  if (it != vit.pos.end()) return it->second;
  const int D = vit.dim, g = 24;
  const float* src = vit.pos_raw.v.data();
  FVec out((size_t)(1 + gh * gw) * D);
  memcpy(out.data(), src, sizeof(float) * D);
  const float sh = (float)g / (float)gh, sw = (float)g / (float)gw;
  for (int y = 0; y < gh; ++y) {
    float fy = sh * ((float)y + 0.5f) - 0.5f;
    if (fy < 0.f) fy = 0.f;
    const int y0 = (int)fy < g - 1 ? (int)fy : g - 1, y1 = y0 + 1 < g ? y0 + 1 : g - 1;
    const float ly = fy - (float)y0, hy = 1.f - ly;
    for (int x = 0; x < gw; ++x) {
      float fx = sw * ((float)x + 0.5f) - 0.5f;
      if (fx < 0.f) fx = 0.f;
      const int x0 = (int)fx < g - 1 ? (int)fx : g - 1, x1 = x0 + 1 < g ? x0 + 1 : g - 1;
      const float lx = fx - (float)x0, hx = 1.f - lx;
      const float *p00 = src + (size_t)(1 + y0 * g + x0) * D, *p01 = src + (size_t)(1 + y0 * g + x1) * D;
      const float *p10 = src + (size_t)(1 + y1 * g + x0) * D, *p11 = src + (size_t)(1 + y1 * g + x1) * D;
      float* o = out.data() + (size_t)(1 + y * gw + x) * D;
      for (int c = 0; c < D; ++c) o[c] = hy * (hx * p00[c] + lx * p01[c]) + ly * (hx * p10[c] + lx * p11[c]);
    }
  }
  float* d = wpool.put(out);
  vit.pos[key] = d;
  return d;
}

void Engine::build_reasm(int idx, int dim, int cout, int kind, int s, bool x3r) {
  const std::string p = "pretrained.act_postprocess" + std::to_string(idx) + ".";
  const int64_t D = dim, c = cout;
  ReasmW r{};
  r.dim = dim;
  r.cout = cout;
  r.kind = kind;
  r.s = s;
  r.x3 = x3r;
  const HostT& wr = W(p + "0.project.0.weight", {D, 2 * D});                // [D, 2D] = [tokens | cls]
  r.b_read = f32(p + "0.project.0.bias", {D});
  const HostT& w1 = W(p + "3.weight", {c, D, 1, 1});
  r.b1 = f32(p + "3.bias", {c});
  if (!x3r) {
    r.w_read = wpool.put(pack_plain(wr.v.data(), 2 * D * D));
    r.w1 = wpool.put(pack_plain(w1.v.data(), c * D));
  } else {
    FVec wt((size_t)(D * D)), wc((size_t)(D * D));
    for (int64_t n = 0; n < D; ++n) {
      memcpy(&wt[n * D], &wr.v[n * 2 * D], sizeof(float) * D);
      memcpy(&wc[n * D], &wr.v[n * 2 * D + D], sizeof(float) * D);
    }
    r.w_tok3 = wpool.put(split3_rows(wt.data(), D, D));
    r.w_cls3 = wpool.put(split3_rows(wc.data(), D, D));
    r.w1 = wpool.put(split3_rows(w1.v.data(), c, D));
  }
  if (kind == 1) {
    const HostT& w4 = W(p + "4.weight", {c, c, s, s});
    if (x3r) {
      const FVec rows = conv_transpose_rows(w4);
      r.w4 = wpool.put(split3_rows(rows.data(), (int64_t)s * s * c, c));
    } else {
      r.w4 = wpool.put(pack_conv_transpose(w4));
    }
    r.b4 = f32(p + "4.bias", {c});
  } else if (kind == 2) {
    const HostT& w4 = W(p + "4.weight", {c, c, 3, 3});
    r.w4 = x3r ? (void*)wpool.put(split3_conv3x3(w4.v.data(), c, c)) : (void*)wpool.put(pack_conv3x3_flat(w4));
    r.b4 = f32(p + "4.bias", {c});
  }
  reasm.push_back(r);
}

void Engine::build_decoder(const int* cin, const int* feat, int head_mid, bool head_ac, bool split3, const bool* split3_in) {
  DecoderW& d = dec;
  d.head_ac = head_ac;
  d.split3 = split3;
  auto p3 = [&](const HostT& w, bool sp) -> void* {
    return sp ? wpool.put(split3_conv3x3(w.v.data(), w.dim(0), w.dim(1))) : wpool.put(pack_conv3x3(w));
  };
  for (int i = 0; i < 4; ++i) {
    d.cin[i] = cin[i];
    d.feat[i] = feat[i];
    d.split3_in[i] = split3 && split3_in && split3_in[i];
    d.rn_w[i] = p3(W("scratch.layer" + std::to_string(i + 1) + "_rn.weight", {feat[i], cin[i], 3, 3}), d.split3_in[i]);
  }
  for (int i = 0; i < 4; ++i) {
    const std::string p = "scratch.refinenet" + std::to_string(i + 1) + ".";
    const int64_t f = feat[i], fo = i > 0 ? feat[i - 1] : feat[0];
    RefineW& r = d.ref[i];
    const char* un[2] = {"resConfUnit1", "resConfUnit2"};
    const char* cn[2] = {"conv1", "conv2"};
    for (int u = 0; u < 2; ++u)
      for (int c = 0; c < 2; ++c) {
        const std::string k = p + un[u] + "." + cn[c];
        r.w[u][c] = p3(W(k + ".weight", {f, f, 3, 3}), split3);
        r.b[u][c] = f32(k + ".bias", {f});
      }
    const HostT& ow = W(p + "out_conv.weight", {fo, f, 1, 1});
    r.out_w = split3 ? wpool.put(split3_rows(ow.v.data(), fo, f)) : wpool.put(pack_plain(ow.v.data(), fo * f));
    r.out_b = f32(p + "out_conv.bias", {fo});
    r.cout = (int)fo;
  }
  const std::string oc = "scratch.output_conv.";
  if (!split3 && head_mid < 64) {
    // MiDaS_small (32 head channels): pad output_conv.0 to 64 output channels (zero filters, zero bias) so the head tail
    // runs on the halo-tiled fused kernel (de_head_up_conv needs Cin % 64 == 0) instead of upsample + a 32-wide GEMM
    // whose epilogue dominated it (0.23 of the 1.8 ms MiDaS_small step).  The padded channels are exact zeros.
    const HostT& w0 = W(oc + "0.weight", {head_mid, feat[0], 3, 3});
    HostT wp;
    wp.shape = {64, feat[0], 3, 3};
    wp.v.assign((size_t)64 * feat[0] * 9, 0.f);
    memcpy(wp.v.data(), w0.v.data(), w0.v.size() * sizeof(float));
    FVec bp(64, 0.f);
    const FVec& b0 = W(oc + "0.bias", {head_mid}).v;
    memcpy(bp.data(), b0.data(), b0.size() * sizeof(float));
    d.h0_c = 64;
    d.h0_w = wpool.put(pack_conv3x3(wp));
    d.h0_b = wpool.put(bp);
    d.h2_w = wpool.put(pack_conv3x3(W(oc + "2.weight", {32, head_mid, 3, 3})));   // Cin padded to 64 with zeros by the packer
  } else {
    d.h0_c = head_mid;
    d.h0_w = p3(W(oc + "0.weight", {head_mid, feat[0], 3, 3}), split3);
    d.h0_b = f32(oc + "0.bias", {head_mid});
    d.h2_w = p3(W(oc + "2.weight", {32, head_mid, 3, 3}), split3);
  }
  d.h2_b = f32(oc + "2.bias", {32});
  d.h4_w = f32(oc + "4.weight", {1, 32, 1, 1});
  d.h4_b = W(oc + "4.bias", {1}).v[0];
}

void Engine::build_hybrid() {
  const std::string bp = "pretrained.model.patch_embed.backbone.";
  {
    // StdConv 7x7/2 stem on the tensor cores: split3 im2col rows x [Wh | Wl | Wh], 147 taps padded to 152, (kh,kw,c) order
    const FVec ws = standardize_weight(W(bp + "stem.conv.weight", {64, 3, 7, 7}));
    FVec wp((size_t)64 * 152, 0.f);
    for (int n = 0; n < 64; ++n)
      for (int c = 0; c < 3; ++c)
        for (int t = 0; t < 49; ++t) wp[(size_t)n * 152 + t * 3 + c] = ws[((size_t)n * 3 + c) * 49 + t];
    hyb.stem_w3 = wpool.put(split3_rows(wp.data(), 64, 152));
    hyb.stem_g = f32(bp + "stem.norm.weight", {64});
    hyb.stem_b = f32(bp + "stem.norm.bias", {64});
  }
  const int couts[3] = {256, 512, 1024}, depths[3] = {3, 4, 9};
  int cin = 64;
  for (int si = 0; si < 3; ++si) {
    std::vector<BottleneckW> blocks;
    const int64_t cout = couts[si], mid = cout / 4;
    for (int b = 0; b < depths[si]; ++b) {
      const std::string p = bp + "stages." + std::to_string(si) + ".blocks." + std::to_string(b) + ".";
      const int64_t bi = b == 0 ? cin : cout;
      BottleneckW k{};
      k.cin = (int)bi;
      k.cout = (int)cout;
      k.mid = (int)mid;
      k.stride = b == 0 ? (si == 0 ? 1 : 2) : 1;
      k.has_proj = b == 0;
      k.w1 = wpool.put(split3_rows(standardize_weight(W(p + "conv1.weight", {mid, bi, 1, 1})).data(), mid, bi));
      k.w2 = wpool.put(split3_conv3x3(standardize_weight(W(p + "conv2.weight", {mid, mid, 3, 3})).data(), mid, mid));
      k.w3 = wpool.put(split3_rows(standardize_weight(W(p + "conv3.weight", {cout, mid, 1, 1})).data(), cout, mid));
      const int64_t nc[3] = {mid, mid, cout};
      for (int i = 0; i < 3; ++i) {
        k.ng[i] = f32(p + "norm" + std::to_string(i + 1) + ".weight", {nc[i]});
        k.nb[i] = f32(p + "norm" + std::to_string(i + 1) + ".bias", {nc[i]});
      }
      if (k.has_proj) {
        k.wd = wpool.put(split3_rows(standardize_weight(W(p + "downsample.conv.weight", {cout, bi, 1, 1})).data(), cout, bi));
        k.ndg = f32(p + "downsample.norm.weight", {cout});
        k.ndb = f32(p + "downsample.norm.bias", {cout});
      }
      blocks.push_back(k);
    }
    hyb.stages.push_back(blocks);
    cin = (int)cout;
  }
  hyb.pe_w = wpool.put(split3_rows(W("pretrained.model.patch_embed.proj.weight", {768, 1024, 1, 1}).v.data(), 768, 1024));
  hyb.pe_b = f32("pretrained.model.patch_embed.proj.bias", {768});
}

// [UPSTREAM] tf_efficientnet_lite3 stages after channel / depth scaling: (ir?, cout, repeats, kernel, stride)
static const int kLite3[7][5] = {{0, 24, 1, 3, 1}, {1, 32, 3, 3, 2}, {1, 48, 3, 5, 2}, {1, 96, 5, 3, 2},
                                 {1, 136, 5, 5, 1}, {1, 232, 6, 5, 2}, {1, 384, 1, 3, 1}};
// upstream Sequential positions of stages 0..6 (blocks._make_efficientnet_backbone)
static const char* const kLite3Layer[7] = {"layer1.3", "layer1.4", "layer2.0", "layer3.0", "layer3.1", "layer4.0", "layer4.1"};

void Engine::build_small() {
  const float eps = 1e-3f;
  auto bn = [&](const std::string& p, int64_t c, FVec& w, FVec& bias) {
    fold_bn(w, c, W(p + "weight", {c}).v, W(p + "bias", {c}).v, W(p + "running_mean", {c}).v, W(p + "running_var", {c}).v, eps, bias);
  };
  auto dw_pack = [&](const FVec& w, int64_t c, int k) {       // [C,1,k,k] -> tap-major bf16 [k*k][C]
    BfVec o((size_t)(c * k * k));
    for (int64_t ch = 0; ch < c; ++ch)
      for (int t = 0; t < k * k; ++t) o[(size_t)t * c + ch] = f2bf(w[ch * k * k + t]);
    return o;
  };
  {
    FVec w = W("pretrained.layer1.0.weight", {32, 3, 3, 3}).v, b;
    bn("pretrained.layer1.1.", 32, w, b);
    FVec wt((size_t)32 * 27);                                  // (kh,kw,c) order
    for (int n = 0; n < 32; ++n)
      for (int c = 0; c < 3; ++c)
        for (int t = 0; t < 9; ++t) wt[(size_t)n * 27 + t * 3 + c] = w[((size_t)n * 3 + c) * 9 + t];
    small.stem_w = wpool.put(wt);
    small.stem_b = wpool.put(b);
  }
  int cin = 32;
  for (int si = 0; si < 7; ++si) {
    const int ir = kLite3[si][0], cout = kLite3[si][1], reps = kLite3[si][2], k = kLite3[si][3], s = kLite3[si][4];
    for (int r = 0; r < reps; ++r) {
      const std::string p = std::string("pretrained.") + kLite3Layer[si] + "." + std::to_string(r) + ".";
      MbBlockW m{};
      m.ir = ir != 0;
      m.k = k;
      m.stride = r == 0 ? s : 1;
      m.cin = cin;
      m.cout = cout;
      m.stage = si;
      m.last = r == reps - 1;
      FVec w, b;
      if (!m.ir) {
        m.mid = cin;
        w = W(p + "conv_dw.weight", {cin, 1, k, k}).v;
        bn(p + "bn1.", cin, w, b);
        m.dw_w = wpool.put(dw_pack(w, cin, k));
        m.dw_b = wpool.put(b);
        w = W(p + "conv_pw.weight", {cout, cin, 1, 1}).v;
        bn(p + "bn2.", cout, w, b);
        m.pw_w = wpool.put(pack_plain(w.data(), (int64_t)w.size()));
        m.pw_b = wpool.put(b);
      } else {
        const int64_t mid = (int64_t)cin * 6;
        m.mid = (int)mid;
        w = W(p + "conv_pw.weight", {mid, cin, 1, 1}).v;
        bn(p + "bn1.", mid, w, b);
        m.exp_w = wpool.put(pack_plain(w.data(), (int64_t)w.size()));
        m.exp_b = wpool.put(b);
        w = W(p + "conv_dw.weight", {mid, 1, k, k}).v;
        bn(p + "bn2.", mid, w, b);
        m.dw_w = wpool.put(dw_pack(w, mid, k));
        m.dw_b = wpool.put(b);
        w = W(p + "conv_pwl.weight", {cout, mid, 1, 1}).v;
        bn(p + "bn3.", cout, w, b);
        m.pw_w = wpool.put(pack_plain(w.data(), (int64_t)w.size()));
        m.pw_b = wpool.put(b);
      }
      m.residual = m.stride == 1 && cin == cout;
      small.blocks.push_back(m);
      cin = cout;
    }
  }
}

void Engine::finalize() {
  if (finalized) raise(DE_ERR_INVALID, "de_finalize_weights: already finalized");
  ENG_CUDA(cudaSetDevice(device));
  const int f256[4] = {256, 256, 256, 256};
  x3 = precision == DE_PRECISION_BF16X3;
  const bool all3[4] = {true, true, true, true};
  if (kind == kLarge) {
    build_vit("pretrained.model.", 1024, 24, 16, x3);
    const HostT& pw = W("pretrained.model.patch_embed.proj.weight", {1024, 3, 16, 16});
    pe_w = x3 ? (void*)wpool.put(split3_rows(pw.v.data(), 1024, 768)) : (void*)wpool.put(pack_plain(pw.v.data(), 1024 * 768));
    pe_b = f32("pretrained.model.patch_embed.proj.bias", {1024});
    build_reasm(1, 1024, 256, 1, 4, x3);
    build_reasm(2, 1024, 512, 1, 2, x3);
    build_reasm(3, 1024, 1024, 0, 0, x3);
    build_reasm(4, 1024, 1024, 2, 0, x3);
    const int cin[4] = {256, 512, 1024, 1024};
    build_decoder(cin, f256, 128, true, x3, all3);
  } else if (kind == kHybrid) {
    build_hybrid();
    build_vit("pretrained.model.", 768, 12, 12, x3);
    build_reasm(3, 768, 768, 0, 0, x3);
    build_reasm(4, 768, 768, 2, 0, x3);
    const int cin[4] = {256, 512, 768, 768};
    const bool sp_in[4] = {true, true, false, false};
    hyb.decoder_split3 = precision != DE_PRECISION_BF16;
    build_decoder(cin, f256, 128, true, hyb.decoder_split3, x3 ? all3 : sp_in);
  } else {
    if (x3) raise(DE_ERR_UNSUPPORTED, "DE_PRECISION_BF16X3 is available for DPT_Large and DPT_Hybrid (MiDaS_small's 24 / 48 / 136-channel maps do not form 64-channel split3 operands)");
    build_small();
    const int cin[4] = {32, 48, 136, 384}, feat[4] = {64, 128, 256, 512};
    build_decoder(cin, feat, 32, false, false, nullptr);
  }
  weight_bytes = (int64_t)wpool.bytes;
  // keep only what plan-time preparation needs (the raw pos-embed was copied into vit.pos_raw)
  raw.clear();
  finalized = true;
}

// ------------------------------------------------------------------------------------------------ plans / predict
static Plan& get_plan(Engine& e, int B, int H, int W) {
  if (!e.finalized) raise(DE_ERR_INVALID, "weights are not finalized (de_finalize_weights)");
  if (B < 1 || H < 1 || W < 1) raise(DE_ERR_INVALID, "bad geometry B=%d H=%d W=%d", B, H, W);
  const std::vector<int64_t> key = {B, H, W};
  auto it = e.plans.find(key);
  if (it != e.plans.end()) return *it->second;
  std::unique_ptr<Plan> p(new Plan());
  p->B = B;
  p->H = H;
  p->W = W;
  ENG_CK(de_get_size(W, H, e.target, e.method, &p->net_w, &p->net_h));
  if (p->net_h < 32 || p->net_w < 32) raise(DE_ERR_UNSUPPORTED, "network input %dx%d too small", p->net_w, p->net_h);
  const size_t px = (size_t)B * p->net_h * p->net_w;
  size_t in_bytes;
  if (e.layout == 1) in_bytes = (size_t)B * (p->net_h / 16) * (p->net_w / 16) * 768 * 2;
  else in_bytes = px * 3 * (e.layout == 2 ? 4 : 2);
  p->x_in = p->ws.pool.alloc(in_bytes);
  ENG_CUDA(cudaMemset(p->x_in, 0, in_bytes));
  p->pred = reinterpret_cast<float*>(p->ws.pool.alloc(px * 4));
  p->splitk_ws_bytes = 64 << 20;                              // K-slice partial tiles of the GEMMs small enough to be split
  p->splitk_ws = p->ws.pool.alloc((size_t)p->splitk_ws_bytes);
  Plan& ref = *p;
  e.plans[key] = std::move(p);
  return ref;
}

static void run_net(Engine& e, Plan& p, cudaStream_t st) {
  if (e.kind == kLarge) forward_large(e, p, st);
  else if (e.kind == kHybrid) forward_hybrid(e, p, st);
  else forward_small(e, p, st);
}

// allocate every work buffer (and the library's tap tables for this geometry) by running the network once on the
// engine's own stream, so that nothing allocates during a later capture
static void warm_up(Engine& e, Plan& p) {
  if (p.warm) return;
  if (!e.side) ENG_CUDA(cudaStreamCreateWithFlags(&e.side, cudaStreamNonBlocking));
  ENG_CUDA(cudaDeviceSynchronize());                          // weights uploaded with synchronous copies; be explicit
  run_net(e, p, e.side);
  ENG_CUDA(cudaStreamSynchronize(e.side));
  p.warm = true;
}

static void run_io(Engine& e, Plan& p, const DePredictIO& io, cudaStream_t st) {
  const int64_t rs = io.row_stride ? io.row_stride : (int64_t)p.W * (io.uv ? 1 : 3);
  const int64_t fs = io.frame_stride ? io.frame_stride : rs * p.H;
  if (io.uv) {
    const int64_t urs = io.uv_row_stride ? io.uv_row_stride : (int64_t)p.W;
    const int64_t ufs = io.uv_frame_stride ? io.uv_frame_stride : urs * (p.H / 2);
    ENG_CK(de_preprocess_nv12(io.frames, io.uv, p.B, p.H, p.W, rs, fs, urs, ufs, p.net_h, p.net_w, e.mean, e.stdv, e.layout,
                              p.x_in, st));
  } else {
    ENG_CK(de_preprocess(io.frames, p.B, p.H, p.W, rs, fs, p.net_h, p.net_w, e.mean, e.stdv, e.layout, p.x_in, st));
  }
  run_net(e, p, st);
  ENG_CK(de_postprocess(p.pred, DE_DTYPE_F32, p.B, p.net_h, p.net_w, p.H, p.W, io.out_map, io.out_metric, io.metric_scale,
                        io.metric_shift, io.out_vis, io.sample_rc, io.n_samples, io.out_samples, st));
}

static void predict(Engine& e, int B, int H, int W, const DePredictIO& io, int use_graph, cudaStream_t st) {
  if (!io.frames) raise(DE_ERR_INVALID, "de_predict: io.frames is NULL");
  if (!io.out_map && !io.out_metric && !io.out_vis && !(io.sample_rc && io.out_samples)) raise(DE_ERR_INVALID, "de_predict: no output requested");
  if (io.uv && (H % 2 || W % 2)) raise(DE_ERR_INVALID, "de_predict: NV12 frames need even H and W");
  ENG_CUDA(cudaSetDevice(e.device));
  Plan& p = get_plan(e, B, H, W);
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  ENG_CUDA(cudaStreamIsCapturing(st, &cs));
  if (cs != cudaStreamCaptureStatusNone) {                    // the caller is recording its own graph: just record into it
    if (!p.warm) raise(DE_ERR_INVALID, "de_predict during stream capture needs a planned geometry (call de_plan first)");
    run_io(e, p, io, st);
    return;
  }
  warm_up(e, p);
  if (!use_graph) {
    run_io(e, p, io, st);
    return;
  }
  uint64_t fbits[2];
  memcpy(&fbits[0], &io.metric_scale, 4);
  memcpy(&fbits[1], &io.metric_shift, 4);
  const std::vector<uint64_t> key = {(uint64_t)io.frames, (uint64_t)io.uv, (uint64_t)io.row_stride, (uint64_t)io.frame_stride,
                                     (uint64_t)io.uv_row_stride, (uint64_t)io.uv_frame_stride, (uint64_t)io.out_map,
                                     (uint64_t)io.out_metric, fbits[0] & 0xffffffffu, fbits[1] & 0xffffffffu, (uint64_t)io.out_vis,
                                     (uint64_t)io.sample_rc, (uint64_t)io.n_samples, (uint64_t)io.out_samples};
  auto it = p.graphs.find(key);
  if (it == p.graphs.end()) {
    if (p.graphs.size() >= 64) {                              // callers cycling through fresh buffers: do not hoard graphs
      for (auto& kv : p.graphs) cudaGraphExecDestroy(kv.second.exec);
      p.graphs.clear();
    }
    // This call runs eagerly (which also builds K1 / K8's tap tables for the geometry); the same launch sequence is
    // then recorded on the engine's own stream - a capture executes nothing, and the instantiated graph can be
    // replayed on any stream, including the legacy default stream that cannot itself be captured.
    run_io(e, p, io, st);
    if (!e.side) ENG_CUDA(cudaStreamCreateWithFlags(&e.side, cudaStreamNonBlocking));
    GraphEntry ge;
    cudaGraph_t g = nullptr;
    const int64_t n0 = de_launch_count();
    ENG_CUDA(cudaStreamBeginCapture(e.side, cudaStreamCaptureModeThreadLocal));
    try {
      run_io(e, p, io, e.side);
    } catch (...) {
      cudaStreamEndCapture(e.side, &g);
      if (g) cudaGraphDestroy(g);
      throw;
    }
    ENG_CUDA(cudaStreamEndCapture(e.side, &g));
    ge.launches = de_launch_count() - n0;
    add_launches(-ge.launches);                               // recorded, not executed
    cudaError_t ie = cudaGraphInstantiate(&ge.exec, g, 0);
    cudaGraphDestroy(g);
    if (ie != cudaSuccess) raise(DE_ERR_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(ie));
    it = p.graphs.emplace(key, ge).first;
    return;                                                   // the eager pass above already produced this call's outputs
  }
  ENG_CUDA(cudaGraphLaunch(it->second.exec, st));
  add_launches(it->second.launches);
}

}  // namespace eng
}  // namespace deb

// ================================================================================================== C ABI
using namespace deb;
using namespace deb::eng;

namespace {
template <typename F>
int guarded(F&& fn) {
  try {
    fn();
    return DE_OK;
  } catch (const EngineError& e) {
    deb::set_error("%s", e.msg.c_str());
    return e.code;
  } catch (const std::bad_alloc&) {
    deb::set_error("out of host memory");
    return DE_ERR_INVALID;
  } catch (const std::exception& e) {
    deb::set_error("%s", e.what());
    return DE_ERR_INVALID;
  }
}
Engine* as_engine(de_handle h) {
  if (!h) raise(DE_ERR_INVALID, "null handle");
  return reinterpret_cast<Engine*>(h);
}
}  // namespace

extern "C" int de_create(const char* model_type, int device, int precision, de_handle* out) {
  return guarded([&] {
    if (!model_type || !out) raise(DE_ERR_INVALID, "de_create: null argument");
    if (precision != DE_PRECISION_DEFAULT && precision != DE_PRECISION_BF16 && precision != DE_PRECISION_BF16X3)
      raise(DE_ERR_INVALID, "de_create: precision %d", precision);
    int ndev = 0;
    ENG_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) raise(DE_ERR_INVALID, "de_create: device %d of %d", device, ndev);
    std::unique_ptr<Engine> e(new Engine());
    const std::string mt = model_type;
    const float half[3] = {0.5f, 0.5f, 0.5f}, im_mean[3] = {0.485f, 0.456f, 0.406f}, im_std[3] = {0.229f, 0.224f, 0.225f};
    if (mt == "DPT_Large" || mt == "DPT_Hybrid") {             // dpt_transform (depth_midas.py:13-14): 384, "minimal", 0.5 / 0.5
      e->kind = mt == "DPT_Large" ? kLarge : kHybrid;
      e->target = 384;
      e->method = 1;
      // K1 output: ViT patch matrix | fp32 NCHW (the hybrid's bf16x3 stem; DPT_Large at bf16x3, re-tiled by de_patchify_split3)
      e->layout = (mt == "DPT_Large" && precision != DE_PRECISION_BF16X3) ? 1 : 2;
      memcpy(e->mean, half, sizeof(half));
      memcpy(e->stdv, half, sizeof(half));
    } else if (mt == "MiDaS_small") {                         // small_transform (:15-16): 256, "upper_bound", ImageNet
      e->kind = kSmall;
      e->target = 256;
      e->method = 0;
      e->layout = 0;
      memcpy(e->mean, im_mean, sizeof(im_mean));
      memcpy(e->stdv, im_std, sizeof(im_std));
    } else {
      raise(DE_ERR_INVALID, "de_create: unknown model_type '%s' (DPT_Large | DPT_Hybrid | MiDaS_small)", model_type);
    }
    e->device = device;
    e->precision = precision;
    *out = reinterpret_cast<de_handle>(e.release());
  });
}

extern "C" int de_set_weight(de_handle h, const char* key, const void* host_ptr, const int64_t* shape, int ndim, int dtype) {
  return guarded([&] {
    Engine* e = as_engine(h);
    if (e->finalized) raise(DE_ERR_INVALID, "de_set_weight after de_finalize_weights");
    if (!key || !host_ptr || ndim < 0 || ndim > 8 || (ndim > 0 && !shape)) raise(DE_ERR_INVALID, "de_set_weight: bad arguments");
    if (dtype != DE_DTYPE_F32 && dtype != DE_DTYPE_BF16) raise(DE_ERR_INVALID, "de_set_weight: dtype %d", dtype);
    HostT t;
    int64_t n = 1;
    for (int i = 0; i < ndim; ++i) {
      if (shape[i] < 0) raise(DE_ERR_INVALID, "de_set_weight: negative extent");
      t.shape.push_back(shape[i]);
      n *= shape[i];
    }
    t.v.resize((size_t)n);
    if (dtype == DE_DTYPE_F32) memcpy(t.v.data(), host_ptr, (size_t)n * 4);
    else
      for (int64_t i = 0; i < n; ++i) t.v[i] = bf2f(reinterpret_cast<const uint16_t*>(host_ptr)[i]);
    e->raw[key] = std::move(t);
  });
}

extern "C" int de_finalize_weights(de_handle h) {
  return guarded([&] { as_engine(h)->finalize(); });
}

extern "C" int de_plan(de_handle h, int B, int H, int W, int* net_h, int* net_w) {
  return guarded([&] {
    Engine* e = as_engine(h);
    ENG_CUDA(cudaSetDevice(e->device));
    Plan& p = get_plan(*e, B, H, W);
    warm_up(*e, p);
    if (net_h) *net_h = p.net_h;
    if (net_w) *net_w = p.net_w;
  });
}

extern "C" int de_predict(de_handle h, int B, int H, int W, const DePredictIO* io, int use_graph, void* stream) {
  return guarded([&] {
    if (!io) raise(DE_ERR_INVALID, "de_predict: io is NULL");
    predict(*as_engine(h), B, H, W, *io, use_graph, reinterpret_cast<cudaStream_t>(stream));
  });
}

extern "C" int de_forward(de_handle h, int B, int H, int W, void** net_input, float** pred, int* net_h, int* net_w, int run, void* stream) {
  return guarded([&] {
    Engine* e = as_engine(h);
    ENG_CUDA(cudaSetDevice(e->device));
    Plan& p = get_plan(*e, B, H, W);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (run) {
      cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
      ENG_CUDA(cudaStreamIsCapturing(st, &cs));
      if (cs == cudaStreamCaptureStatusNone) warm_up(*e, p);
      else if (!p.warm) raise(DE_ERR_INVALID, "de_forward during stream capture needs a planned geometry (call de_plan first)");
      run_net(*e, p, st);
    }
    if (net_input) *net_input = p.x_in;
    if (pred) *pred = p.pred;
    if (net_h) *net_h = p.net_h;
    if (net_w) *net_w = p.net_w;
  });
}

extern "C" int de_engine_info(de_handle h, int* layout, int64_t* weight_bytes, int64_t* work_bytes) {
  return guarded([&] {
    Engine* e = as_engine(h);
    if (layout) *layout = e->layout;
    if (weight_bytes) *weight_bytes = e->weight_bytes;
    if (work_bytes) {
      int64_t n = 0;
      for (auto& kv : e->plans) n += (int64_t)kv.second->ws.pool.bytes;
      *work_bytes = n;
    }
  });
}

extern "C" void de_destroy(de_handle h) {
  if (!h) return;
  Engine* e = reinterpret_cast<Engine*>(h);
  cudaSetDevice(e->device);
  cudaDeviceSynchronize();
  e->plans.clear();
  if (e->side) cudaStreamDestroy(e->side);
  for (int i = 0; i < Engine::kMaxLanes; ++i) {
    if (e->lane_stream[i]) cudaStreamDestroy(e->lane_stream[i]);
    if (e->lane_join[i]) cudaEventDestroy(e->lane_join[i]);
  }
  if (e->lane_fork) cudaEventDestroy(e->lane_fork);
  delete e;
}
