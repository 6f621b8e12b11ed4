// Handle-level engine behind de_create / de_set_weight / de_finalize_weights / de_plan / de_predict / de_forward /
// de_destroy (include/depth_b200.h): everything between the frame and the depth map that is not a
// kernel - weight folding / repacking into the kernel-native layouts, per-geometry work buffers, the launch
// sequences of the three hub models and CUDA-graph capture - in C++, so that any host language that can call a
// C ABI runs the path of /root/reference/depth_midas.py:21-35.  Internal header (engine*.cu only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "depth_b200.h"
#include "host_util.h"

namespace deb {
namespace eng {

struct EngineError {
  int code;
  std::string msg;
};
[[noreturn]] void raise(int code, const char* fmt, ...);
// operator call that already set de_last_error(): propagate
#define ENG_CK(call)                                          \
  do {                                                        \
    int rc__ = (call);                                        \
    if (rc__ != DE_OK) throw EngineError{rc__, de_last_error()}; \
  } while (0)
#define ENG_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess) raise(DE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e__)); \
  } while (0)

// ---------------------------------------------------------------------------------------------- host tensors
struct HostT {
  std::vector<float> v;
  std::vector<int64_t> shape;
  int64_t numel() const { return (int64_t)v.size(); }
  int64_t dim(int i) const { return shape[i]; }
};

inline uint16_t f2bf(float f) {          // round to nearest even, like torch .to(bfloat16)
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);   // NaN stays NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
inline float bf2f(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

typedef std::vector<uint16_t> BfVec;
typedef std::vector<float> FVec;

// one-time repacking of upstream-format weights (SURVEY.md section 5 "Checkpoint / resume")
BfVec pack_plain(const float* w, int64_t n);                               // bf16 copy ([N,K] matrices, 1x1 convs)
BfVec pack_conv3x3(const HostT& w);                                        // [Cout,Cin,3,3] -> [Cout, 9*ceil64(Cin)] tap-major
BfVec pack_conv3x3_flat(const HostT& w);                                   // -> [Cout, 9*Cin] (im2col + GEMM)
BfVec pack_conv_transpose(const HostT& w);                                 // [Cin,Cout,s,s] -> [(i*s+j)*Cout+co, Cin]
FVec conv_transpose_rows(const HostT& w);                                  // the same matrix in fp32
BfVec split3_rows(const float* w, int64_t rows, int64_t cols);             // [rows, cols] -> [rows, 3*cols] = [Wh | Wl | Wh]
BfVec split3_conv3x3(const float* w, int64_t cout, int64_t cin);           // [Cout,Cin,3,3] -> [Cout, 9*3*Cin] per tap [Wh | Wl | Wh]
void fold_bn(FVec& w, int64_t cout, const FVec& g, const FVec& b, const FVec& mean, const FVec& var, float eps, FVec& bias_out);
FVec standardize_weight(const HostT& w, double eps = 1e-8);                // timm StdConv2dSame, folded at load

// ---------------------------------------------------------------------------------------------- device memory
struct DevPool {
  std::vector<void*> allocs;
  size_t bytes = 0;
  void* alloc(size_t n);
  void* upload(const void* host, size_t n);
  template <typename T>
  T* put(const std::vector<T>& v) { return reinterpret_cast<T*>(upload(v.data(), v.size() * sizeof(T))); }
  void release();
  ~DevPool() { release(); }
};

// named work buffers of one geometry: allocated during the eager warm-up pass, looked up during capture / eager runs
struct WorkSet {
  DevPool pool;
  std::unordered_map<std::string, std::pair<void*, size_t>> bufs;
  void* get(const std::string& key, size_t bytes, bool zero = false);
};

// ---------------------------------------------------------------------------------------------- model weights
struct VitBlockW {
  float *n1w, *n1b, *qkv_b, *proj_b, *n2w, *n2b, *fc1_b, *fc2_b;
  void *qkv_w, *proj_w, *fc1_w, *fc2_w;
};
struct VitW {
  int dim = 0, depth = 0, heads = 0;
  bool x3 = false;                                             // bf16x3: weights [Wh | Wl | Wh] along K, activations split3
  float* cls = nullptr;
  HostT pos_raw;                                              // [1, 1+g*g, D]
  std::vector<VitBlockW> blocks;
  std::map<std::pair<int, int>, float*> pos;                  // resized pos-embed per token grid, device fp32 [(1+gh*gw), D]
};
struct ReasmW {
  int dim = 0, cout = 0, kind = 0, s = 0;                      // kind 0 same, 1 up (ConvT k == s), 2 down2 (conv3x3 s2)
  bool x3 = false;
  void *w_read = nullptr, *w1 = nullptr, *w4 = nullptr;        // w_read bf16 [D, 2D] = [tokens | cls]
  void *w_tok3 = nullptr, *w_cls3 = nullptr;                   // bf16x3: the two halves packed separately, [D, 3D] each
  float *b_read = nullptr, *b1 = nullptr, *b4 = nullptr;
};
struct RefineW {
  void* w[2][2];                                               // [resConfUnit1|2][conv1|2]
  float* b[2][2];
  void* out_w;
  float* out_b;
  int cout;
};
struct DecoderW {
  int cin[4], feat[4];
  bool head_ac = true, split3 = false, split3_in[4] = {false, false, false, false};
  void* rn_w[4];
  RefineW ref[4];
  void *h0_w, *h2_w;
  float *h0_b, *h2_b, *h4_w;
  float h4_b = 0.f;
  int h0_c = 0;
};
struct BottleneckW {
  int cin, cout, mid, stride;
  bool has_proj;
  void *w1, *w2, *w3, *wd = nullptr;
  float *ng[3], *nb[3], *ndg = nullptr, *ndb = nullptr;
};
struct HybridW {
  void* stem_w3;                                               // split3 [64, 3*152]
  float *stem_g, *stem_b;
  std::vector<std::vector<BottleneckW>> stages;
  void* pe_w;
  float* pe_b;
  bool decoder_split3 = true;
};
struct MbBlockW {
  bool ir;
  int k, stride, cin, cout, mid, stage;
  bool last, residual;
  void *exp_w = nullptr, *dw_w, *pw_w;
  float *exp_b = nullptr, *dw_b, *pw_b;
};
struct SmallW {
  float *stem_w, *stem_b;                                      // fp32 [32, 27], [32]
  std::vector<MbBlockW> blocks;
};

enum ModelKind { kLarge = 0, kHybrid = 1, kSmall = 2 };

struct Plan;

struct Engine {
  ModelKind kind;
  int device = 0;
  int precision = 0;                                           // DE_PRECISION_*
  bool finalized = false;
  std::map<std::string, HostT> raw;                            // upstream key -> fp32 host copy (dropped by finalize)
  DevPool wpool;                                               // packed weights
  VitW vit;
  std::vector<ReasmW> reasm;
  DecoderW dec;
  HybridW hyb;
  SmallW small;
  bool x3 = false;                                             // DE_PRECISION_BF16X3
  void* pe_w = nullptr;                                        // DPT_Large patch-embed [D, 768] (bf16x3: [D, 3*768])
  float* pe_b = nullptr;
  float mean[3], stdv[3];
  int target = 384, method = 1, layout = 0;                    // transform: get_size target / method, K1 output layout
  std::map<std::vector<int64_t>, std::unique_ptr<Plan>> plans; // key (B, H, W)
  cudaStream_t side = nullptr;                                 // warm-up stream
  static const int kMaxLanes = 4;                              // MiDaS_small: concurrent sub-batches (engine_nets.cu)
  cudaStream_t lane_stream[kMaxLanes] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t lane_fork = nullptr, lane_join[kMaxLanes] = {nullptr, nullptr, nullptr, nullptr};
  void ensure_lanes(int n);
  int64_t weight_bytes = 0;

  // weight access during finalize
  const HostT& W(const std::string& key, std::initializer_list<int64_t> shape);
  float* f32(const std::string& key, std::initializer_list<int64_t> shape);
  void finalize();
  void build_vit(const std::string& prefix, int dim, int depth, int heads, bool x3);
  void build_reasm(int idx, int dim, int cout, int kind, int s, bool x3);
  void build_decoder(const int* cin, const int* feat, int head_mid, bool head_ac, bool split3, const bool* split3_in);
  void build_hybrid();
  void build_small();
  float* pos_embed(int gh, int gw);
};

struct GraphEntry {
  cudaGraphExec_t exec = nullptr;
  int64_t launches = 0;
};

struct Plan {
  int B, H, W, net_h, net_w;
  WorkSet ws;
  void* x_in = nullptr;                                        // K1 output (network input)
  float* pred = nullptr;                                       // network output [B, net_h, net_w]
  void* splitk_ws = nullptr;                                   // fp32 scratch for the K-slices of split-K GEMMs
  int64_t splitk_ws_bytes = 0;
  bool warm = false;
  std::map<std::vector<uint64_t>, GraphEntry> graphs;          // keyed on the IO pointers / scalars baked into the capture
  ~Plan();
};

// launch sequences (engine_nets.cu); x_in / pred live in the plan
void forward_large(Engine& e, Plan& p, cudaStream_t st);
void forward_hybrid(Engine& e, Plan& p, cudaStream_t st);
void forward_small(Engine& e, Plan& p, cudaStream_t st);

}  // namespace eng
}  // namespace deb
