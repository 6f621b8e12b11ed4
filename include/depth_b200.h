/* depth_b200.h - C ABI of libdepthb200.so: the B200 (sm_100a) kernels behind the per-frame MiDaS
 * depth path of juhoward/depth_estimation.
 *
 * The reference has no FFI of its own (it is pure Python): the boundary it offers is
 * the Python class DepthEstimator (/root/reference/depth_midas.py:6-35 == midas.py:6-35).  Every entry
 * point below replaces one slice of DepthEstimator.predict or of the un-vendored upstream modules it
 * calls, and is bound with ctypes by depth_estimation_b200/_lib.py.
 *
 * Conventions: plain pointers + sizes, no torch types.  All data pointers are DEVICE pointers owned by
 * the caller unless a parameter says "host".  Work is enqueued on `stream` (a cudaStream_t passed as
 * void*), never synchronised.  Return value: 0 on success, negative DE_ERR_* otherwise;
 * de_last_error() returns a thread-local message for the last failure.
 */
#ifndef DEPTH_B200_H_
#define DEPTH_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DE_OK 0
#define DE_ERR_INVALID (-1)
#define DE_ERR_CUDA (-2)
#define DE_ERR_UNSUPPORTED (-3)

#define DE_DTYPE_BF16 0
#define DE_DTYPE_F32 1

#define DE_ACT_NONE 0
#define DE_ACT_RELU 1
#define DE_ACT_GELU 2  /* exact erf GELU (timm nn.GELU) */
#define DE_ACT_RELU6 3

const char* de_last_error(void);
int de_version(void);
/* Number of kernels this library has launched in the calling process (bench.py "gpu_launches"). */
int64_t de_launch_count(void);

/* Per-operator timing for measurement runs (bench.py's roofline): between de_profile_start and de_profile_stop every
 * operator entry point of this library brackets its launches with CUDA events on the stream it is given (eager launches
 * only; calls recorded into a stream capture are skipped).  de_profile_stop synchronises the device and returns one
 * record per call in issue order.  dims: operator-specific extents (de_gemm: M, N, K, a_mode*100 + act*10 + accumulate). */
typedef struct DeProfileRecord {
  char name[24];
  double ms;
  double flops; /* algorithmic 2*MACs of the call; 0 for bandwidth operators */
  int64_t dims[4];
} DeProfileRecord;
int de_profile_start(void);
int de_profile_stop(DeProfileRecord* out, int max_records, int* n_records);

/* ---- get_size: [UPSTREAM] midas/transforms.py Resize.get_size, reached from depth_midas.py:23.
 * method 0 = "upper_bound" (small_transform, T=256), 1 = "minimal" (dpt_transform, T=384). */
int de_get_size(int width, int height, int target, int method, int* net_w, int* net_h);

/* ---- cubic tap tables (host): cv2.resize INTER_CUBIC coordinate rule (mode 0; used by the
 * transform, depth_midas.py:23) or torch F.interpolate bicubic align_corners=False rule (mode 1;
 * depth_midas.py:28-33).  idx: int32 [dst*4] border-clamped source indices, w: float [dst*4]. */
int de_cubic_taps(int src, int dst, int mode, int32_t* idx_host, float* w_host);

/* ---- K1: replaces depth_midas.py:22-23 (cvtColor BGR2RGB + upstream transform chain
 * /255 -> Resize(INTER_CUBIC, multiple of 32) -> NormalizeImage -> PrepareForNet).
 * bgr: uint8 [B,H,W,3] (row_stride bytes per row, frame_stride bytes per frame).
 * Tap tables (de_cubic_taps mode 0) are built and cached inside the library per (H,W,net_h,net_w);
 * the first call for a new shape allocates - warm up before CUDA-graph capture.
 * mean/std: host float[3] in RGB order.
 * layout 0: out = bf16 NCHW [B,3,net_h,net_w].
 * layout 1: out = bf16 patch matrix [B*(net_h/16)*(net_w/16), 768], column = c*256+py*16+px
 *           (the im2col of the ViT patch-embed conv, feeding de_gemm directly).
 * layout 2: out = float32 NCHW (parity/debug). */
int de_preprocess(const uint8_t* bgr, int B, int H, int W, int64_t row_stride, int64_t frame_stride,
                  int net_h, int net_w, const float* mean_host, const float* std_host, int layout, void* out,
                  void* stream);

/* K1 on an NV12 frame as a hardware decoder delivers it (replaces the host BGR hop in front of
 * depth_midas.py:22): y_plane uint8 [B,H,W] (strides in bytes), uv_plane uint8 [B,H/2,W/2,2] interleaved U,V.  The
 * colour conversion is fused into the resampling kernel and is bit-exact with
 * cv2.cvtColor(nv12, COLOR_YUV2BGR_NV12) followed by de_preprocess.  H and W even. */
int de_preprocess_nv12(const uint8_t* y_plane, const uint8_t* uv_plane, int B, int H, int W, int64_t y_row_stride,
                       int64_t y_frame_stride, int64_t uv_row_stride, int64_t uv_frame_stride, int net_h, int net_w,
                       const float* mean_host, const float* std_host, int layout, void* out, void* stream);

/* ---- K8: replaces depth_midas.py:28-33 (F.interpolate bicubic, align_corners=False, to the source
 * resolution) fused with the metric tail of face.py:106-108,137 (ri_depth sampling, /2.54).
 * pred: [B,h,w] (in_dtype bf16 or f32).  Tap tables (mode 1) cached inside the library.
 * out_map: [B,H,W] f32 (may be NULL).  out_metric: [B,H,W] bf16 = map*metric_scale + metric_shift (may be NULL):
 * (1/2.54, 0) is face.py:137 rel2abs; (m/30.48, b/30.48) the calibrated rel2abs_2 of face.py:118-130 in feet.
 * out_vis: [B,H,W] uint8 (may be NULL) = the reference's to_video_frame grey level (midas_mediapipe.py:141-150,
 * depth_midas.py:83): uint8(map) * 5 + 10 with numpy's wrapping uint8 arithmetic.
 * sample_rc: int32 [B,n_samples,2] (row, col) device indices into the OUTPUT map (may be NULL);
 * out_samples: float [B,n_samples] receives the interpolated value at those points. */
int de_postprocess(const void* pred, int in_dtype, int B, int h, int w, int H, int W, float* out_map,
                   void* out_metric, float metric_scale, float metric_shift, uint8_t* out_vis, const int32_t* sample_rc,
                   int n_samples, float* out_samples, void* stream);

/* ---- tcgen05 GEMM / implicit-GEMM convolution.  D = epilogue(A * Wt^T):
 *   a_mode 0: A is a bf16 row-major matrix [M,K] (lda elements); K % 8 == 0.
 *   a_mode 1: A is a bf16 NHWC activation [B,H,W,Cin]; 3x3, stride 1, pad 1 convolution; the weight
 *             is [N, 9*cin_pad] with k = tap*cin_pad + ci (tap = kh*3+kw, cin_pad = ceil64(Cin)).
 *   w: bf16 [N,K] row-major (ldw elements) - the torch Linear / flattened-conv layout.
 * Epilogue, per output element (row m, column n), fp32:
 *   v = acc + bias[n % bias_mod]; if (add1 && add1_pre) v += add1; v = act(v);
 *   if (add1 && !add1_pre) v += add1; if (add2) v += add2; store(out); if (out_relu) store(relu(v)).
 * Row maps (m -> storage row): out_rowmode 0: m; 1: m + (m/grp)*pad + off;
 *   2: pixel shuffle (ConvTranspose k==s): rows are (b,y,x) on a gh x gw grid, columns are
 *      (i,j,co); out is NHWC [B,gh*s,gw*s,cout].
 * add1_rowmode 0: same storage row as out; 1: m / grp (per-group broadcast); 2: (m % grp) + off.
 * head_w != NULL (requires N == 32): out is float [rows]; out[m] = relu(sum_n v[n]*head_w[n] + head_b). */
typedef struct DeGemm {
  int a_mode;
  const void* a;
  int64_t lda;
  int M, N, K;
  int B, H, W, Cin; /* a_mode 1 */
  const void* w;
  int64_t ldw;
  const float* bias;
  int bias_mod; /* 0 -> N */
  int act;
  const void* add1;
  int add1_dtype;
  int64_t ld_add1;
  int add1_rowmode;
  int add1_pre;
  const void* add2; /* bf16 or f32 (add2_dtype), storage row of out, pitch ld_add2 */
  void* out;
  int out_dtype;
  int64_t ldo;
  void* out_relu; /* bf16, same rows as out, pitch ld_relu */
  int out_rowmode;
  int grp, pad, off;
  int ps_gh, ps_gw, ps_s, ps_cout; /* out_rowmode 2 */
  const float* head_w;
  float head_b;
  int block_n; /* 0 = auto; else 32/64/128/256 */
  int epilogue;  /* 0 = auto (TMA-store epilogue whenever the call is bias + activation -> dense output); 1 = always
                  * the general register epilogue (cross-check of the two code paths) */
  int cta_group; /* 0 = auto; 1 = one CTA per 128-row tile; 2 = CTA pair per 256-row tile (block_n 256) */
  /* bf16x3 ("split3") plumbing, general epilogue only.  A split3 operand row is [hi | hi | lo] (hi = bf16(v),
   * lo = bf16(v - hi)), each part N (pixel shuffle: ps_cout) wide; against weights packed [Wh | Wl | Wh] along K the
   * ordinary bf16 MMA accumulates Ah*Wh + Ah*Wl + Al*Wh in fp32 (~16 mantissa bits). */
  int out_split;     /* 1: `out` (bf16, ldo >= 3N) is written as a split3 row */
  int relu_split;    /* 1: `out_relu` is written as a split3 row */
  int64_t ld_relu;   /* row pitch of out_relu in elements; 0 = ldo */
  int add2_dtype;    /* DE_DTYPE_BF16 (default) or DE_DTYPE_F32 (requires add1_pre == 0) */
  int64_t ld_add2;   /* row pitch of add2 in elements; 0 = ldo (pixel shuffle: ps_cout) */
  /* Split-K (small grids: the batch-1 latency path).  k_splits 0 = automatic, 1 = off, > 1 = forced.  Each K-slice stores
   * its fp32 partial tiles into its own region of splitk_ws (device scratch, 16-byte aligned, no initial contents
   * required; a slice is ceil128(M)*N*4 bytes, a_mode 1: B*H*W*N*4; the slice count is capped by the size given) and
   * one finalize launch adds the slices in a fixed order - bitwise reproducible - and applies the epilogue (accumulate
   * included).  Without a workspace a call never splits. */
  int k_splits;
  void* splitk_ws;
  int64_t splitk_ws_bytes;
  int accumulate; /* 1: out += epilogue(A W^T) instead of out = ... (fp32 dense output, a_mode 0, bias + activation
                   * only): the tile leaves through a TMA reduce-add, so the ViT residual stream x += proj(..) /
                   * x += fc2(..) needs neither a bf16 branch buffer nor an add pass.  Each element receives exactly one
                   * add per call - deterministic. */
} DeGemm;
int de_gemm(const DeGemm* g, void* stream);

/* ---- (residual add +) LayerNorm over the last dim (timm ViT norm1/norm2, eps 1e-6).
 * if y != NULL: x += y in place (fp32 residual stream; y = the branch output a GEMM wrote, y_dtype f32 or bf16);
 * if out_bf16 != NULL: out = LN(x) * gamma + beta in bf16.  At least one of y / out_bf16 is required. */
int de_layernorm(float* x, int64_t ldx, const void* y, int y_dtype, int64_t ldy, const float* gamma, const float* beta,
                 int rows, int dim, float eps, void* out_bf16, int64_t ldo, void* stream);

/* LayerNorm whose output row is the bf16x3 ("split3") operand [hi | hi | lo] (each `dim` wide, ldo >= 3*dim) - see
 * DeGemm.out_split; used by precision DE_PRECISION_BF16X3. */
int de_layernorm_split3(float* x, int64_t ldx, const float* gamma, const float* beta, int rows, int dim, float eps,
                        void* out_split3, int64_t ldo, void* stream);

/* ---- multi-head self-attention on a packed qkv buffer (timm Attention.forward):
 * qkv: bf16 [B*T, 3*heads*64] (q | k | v, each heads x 64), out: bf16 [B*T, heads*64].
 * softmax(q k^T * scale) v, non-causal, fp32 softmax. head_dim must be 64.
 * tcgen05/TMEM flash kernel: S and O tiles in tensor memory, Q/K/V by TMA, V consumed MN-major. */
int de_attention(const void* qkv, void* out, int B, int T, int heads, float scale, void* stream);

/* The same attention with the output row written as a split3 operand: out bf16 [B*T, 3*heads*64] = [hi | hi | lo] of the
 * fp32 result (the qkv input and the probabilities stay bf16). */
int de_attention_split3(const void* qkv, void* out_split3, int B, int T, int heads, float scale, void* stream);

/* ---- small bandwidth kernels used between the GEMMs ---- */
/* fp32 [rows, dim] (ldx) -> bf16 [rows', dim]: drops the first `skip` rows of every group of `grp`
 * rows (cls token removal for the reassemble readout).  grp=0: plain cast. */
int de_cast_tokens(const float* x, int64_t ldx, int rows, int dim, int grp, int skip, void* out_bf16,
                   void* stream);
/* de_cast_tokens with split3 output rows: bf16 [rows', 3*dim] = [hi | hi | lo]. */
int de_cast_tokens_split3(const float* x, int64_t ldx, int rows, int dim, int grp, int skip, void* out_split3, void* stream);
/* fp32 NCHW network input [B,3,h,w] (de_preprocess layout 2) -> split3 ViT patch matrix bf16 [B*(h/16)*(w/16), 3*768],
 * column = c*256+py*16+px inside each part: the bf16x3 form of de_preprocess layout 1. */
int de_patchify_split3(const float* x_nchw, void* out_split3, int B, int h, int w, void* stream);
/* x[b*T + 0, :] = cls[:] + pos[0, :] (fp32 token stream) */
int de_fill_cls(float* x, int64_t ldx, int B, int T, int dim, const float* cls, const float* pos,
                void* stream);
/* bilinear x2 upsample NHWC bf16 [B,H,W,C] -> [B,2H,2W,C]; align_corners 1/0. */
int de_upsample2x(const void* x, void* out, int B, int H, int W, int C, int align_corners, void* stream);
/* The same resampling for the bf16x3 decoder: x NHWC bf16 or f32 (x_dtype); out_mode 0: bf16 [B,2H,2W,C]; 1: split3 bf16
 * [B,2H,2W,3C] = [hi | hi | lo] (see DeGemm.out_split); 2: f32 [B,2H,2W,C]. */
int de_upsample2x_ex(const void* x, int x_dtype, void* out, int out_mode, int B, int H, int W, int C, int align_corners,
                     void* stream);
/* im2col for a 3x3 window, stride s, explicit leading pad: NHWC bf16 -> [B*Ho*Wo, 9*C], k = tap*C + c.
 * torch Conv2d(k3,s2,p1) (act_postprocess4): pad 1, Ho = (H-1)/2+1.  TF-same s2 (ResNetV2): pad
 * ((ceil(H/2)-1)*2+3-H)/2, Ho = ceil(H/2). */
int de_im2col3x3(const void* x, void* out, int B, int H, int W, int C, int stride, int pad_y, int pad_x, int Ho,
                 int Wo, void* stream);
/* out[b,y,x,:] = x[b,y*s,x*s,:] (input gather of a 1x1 stride-s conv), NHWC bf16. */
int de_subsample(const void* x, void* out, int B, int H, int W, int C, int stride, void* stream);
/* GroupNorm over (HW, C/groups) per sample, NHWC: y = act((x-mean)*rstd*gamma+beta [+ add]).
 * x: bf16 or f32 (x_dtype).  add: bf16 or f32 (add_dtype), same shape, may be NULL.  Outputs (each may be NULL, at
 * least one required): out - bf16, out_mode 0: [.., C]; out_mode 1: the "split3" operand layout [.., 3C] =
 * [hi | hi | lo] with hi = bf16(y), lo = bf16(y - hi), which against weights packed [Wh | Wl | Wh] along K gives a
 * bf16x3 (~16 mantissa bit) product on the ordinary bf16 GEMM (DPT_Hybrid's ResNetV2 stages); out_f32 - fp32 copy
 * (the ResNetV2 residual stream); out_plain - bf16 [.., C] copy (decoder input).  Deterministic (fixed-order
 * reductions).  ws: device scratch of at least de_groupnorm_ws_bytes(B, HW, C, groups) bytes.  Its first 16 KB are
 * per-sample arrival counters: they must be ZERO before the first call (cudaMemset the buffer once) and every call
 * leaves them zero; calls sharing a workspace must be ordered on one stream.  B <= 4096.
 * (timm GroupNormAct, eps 1e-5, 32 groups.) */
int64_t de_groupnorm_ws_bytes(int B, int HW, int C, int groups); /* 0 for unsupported arguments */
int de_groupnorm(const void* x, int x_dtype, const float* gamma, const float* beta, const void* add, int add_dtype,
                 void* out, int out_mode, float* out_f32, void* out_plain, int B, int HW, int C, int groups, float eps,
                 int act, void* ws, int64_t ws_bytes, void* stream);
/* de_groupnorm without a caller workspace, for callers that may not allocate device memory (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arithmetic, arguments and outputs as de_groupnorm; the scratch is a file-scope
 * `static __device__` array of DE_GROUPNORM_STATIC_MAX_BYTES inside the library (every GPU that loads the library holds
 * it until exit).  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).  Shapes whose
 * de_groupnorm_ws_bytes(B, HW, C, groups) exceeds the limit return DE_ERR_UNSUPPORTED with the needed size in
 * de_last_error() - use de_groupnorm then.  All calls on one device share the one array, which keeps the last call's
 * contents (the arrival counters are re-zeroed on `stream` by every call): calls must be ordered on one stream; two
 * calls that overlap on the GPU race on it. */
#define DE_GROUPNORM_STATIC_MAX_BYTES (4 << 20)
int64_t de_groupnorm_static_max_bytes(void);
int de_groupnorm_static(const void* x, int x_dtype, const float* gamma, const float* beta, const void* add, int add_dtype,
                        void* out, int out_mode, float* out_f32, void* out_plain, int B, int HW, int C, int groups,
                        float eps, int act, void* stream);

/* Fused DPT head tail ([UPSTREAM] midas/dpt_depth.py head, output_conv[2..5]):
 * out[b,y,x] = relu(head_b + sum_n head_w[n] * relu(bias[n] + conv3x3(x, w)[b,y,x,n])), n < 32.
 * x: NHWC bf16 [B,H,W,Cin], Cin % 64 == 0 (filter bank must fit shared memory: Cin <= 192); w: bf16 [32, ldw >= 9*Cin],
 * k = (kh*3+kw)*Cin + ci (weights.pack_conv3x3); bias / head_w: float[32]; out: float [B,H,W].  Halo-tiled tcgen05
 * kernel with the filter bank resident in shared memory (csrc/head_conv.cu). */
int de_head_conv(const void* x, const void* w, int64_t ldw, const float* bias, const float* head_w, float head_b,
                 float* out, int B, int H, int W, int Cin, void* stream);

/* The same head tail with the bilinear x2 upsample in front of it fused in (output_conv[1..5]): x is the LOW-resolution
 * NHWC bf16 activation [B,Hs,Ws,Cin]; out: float [B,2*Hs,2*Ws].  The upsampled activation is interpolated into shared
 * memory (rounded to bf16 exactly as de_upsample2x does) and never written to HBM. */
int de_head_up_conv(const void* x, const void* w, int64_t ldw, const float* bias, const float* head_w, float head_b,
                    float* out, int B, int Hs, int Ws, int Cin, int align_corners, void* stream);

/* ---- direct (CUDA-core) convolutions for non-GEMM shapes; TF-"same" padding (geffnet Conv2dSame /
 * timm StdConv2dSame, MaxPool2dSame): pad_total = max((ceil(i/s)-1)*s+k-i, 0), extra pixel bottom/right. */
/* 3-channel stem: x NCHW bf16 or f32 [B,3,H,W] (K1 layout 0 / 2), w float [Cout][k*k*3] ((kh,kw,c) order, BN
 * folded), bias float[Cout] or NULL -> NHWC bf16 or f32 [B,ceil(H/s),ceil(W/s),Cout]. Cout % 16 == 0. */
int de_stem_conv(const void* x_nchw, int x_dtype, const float* w, const float* bias, void* out_nhwc, int out_dtype,
                 int B, int H, int W, int Cout, int k, int stride, int act, void* stream);
/* The same stem as a tensor-core GEMM operand: x NCHW f32 [B,3,H,W] -> bf16 [B*ceil(H/s)*ceil(W/s), 3*kp], row =
 * [hi(kp) | hi(kp) | lo(kp)] of the TF-"same"-padded k x k x 3 window in (kh,kw,c) order (zero for columns >= 3*k*k;
 * kp % 8 == 0), hi = bf16(v), lo = bf16(v - hi).  de_gemm against the filter bank packed [Wh | Wl | Wh] (each kp wide)
 * gives the stem convolution at bf16x3 precision (DPT_Hybrid's StdConv 7x7/2; replaces timm
 * ResNetV2.stem.conv inside the hub model of /root/reference/depth_midas.py:10). */
int de_stem_im2col_split3(const float* x_nchw, void* out_bf16, int B, int H, int W, int k, int stride, int kp, void* stream);
/* depthwise k in {3,5}: x NHWC bf16, w bf16 [k*k][C] (tap-major, BN folded), bias float[C]. */
int de_depthwise(const void* x, const void* w_tapmajor, const float* bias, void* out, int B, int H, int W, int C,
                 int k, int stride, int act, void* stream);
/* Fused depthwise k3 / k5 (stride 1, TF-same padding, + bias, ReLU6) -> pointwise-linear 1x1 (+ bias [+ residual]) of an
 * MBConv block ([UPSTREAM] geffnet InvertedResidual conv_dw/bn2/act/conv_pwl/bn3, BatchNorm folded): what
 * de_depthwise followed by de_gemm computes, without the depthwise output ever reaching global memory.
 * x: NHWC bf16 [B,H,W,C] (the expanded activation); dw_w: bf16 [k*k][C] tap-major; pw_w: bf16 [cout][C]; residual / out:
 * NHWC bf16 [B,H,W,cout].  C, cout multiples of 8, cout <= 384.  ws: device scratch of de_dwproject_ws_bytes(B,H,W,cout)
 * bytes whose first 16 KB (arrival counters) must be zero before the first call; every call leaves them zero; calls
 * sharing a workspace must be ordered on one stream.  Deterministic (fixed-order sums). */
int64_t de_dwproject_ws_bytes(int B, int H, int W, int cout);
int de_dwproject(const void* x, const void* dw_w, const float* dw_b, const void* pw_w, const float* pw_b,
                 const void* residual, void* out, int B, int H, int W, int C, int cout, int k, void* ws,
                 int64_t ws_bytes, void* stream);
/* MaxPool 3x3 stride 2 TF-same, NHWC bf16 or f32 in; out bf16 [.., C] (out_mode 0) or split3 [.., 3C] (1). */
int de_maxpool3x3s2(const void* x, int x_dtype, void* out, int out_mode, int B, int H, int W, int C, void* stream);

/* ================================================================================================================
 * Handle-level API: the whole of DepthEstimator (/root/reference/depth_midas.py:6-35) for a host
 * that is not Python.  de_create == `__init__` up to the hub load (:7-19), de_set_weight / de_finalize_weights take the
 * place of `torch.hub.load(...)`'s checkpoint (upstream state_dict naming; BatchNorm / weight standardisation folded
 * and everything repacked into the kernel-native layouts once), de_predict == `predict` (:21-35) for B frames that are
 * already in device memory.  One handle per GPU; calls on one handle must be serialised by the caller; distinct handles
 * may be used from distinct threads.  Nothing throws across the ABI.
 * ================================================================================================================ */
typedef struct de_engine* de_handle;

#define DE_PRECISION_DEFAULT 0 /* bf16 operands / fp32 accumulate; DPT_Hybrid: ResNetV2 stages + decoder as bf16x3 */
#define DE_PRECISION_BF16 1    /* DPT_Hybrid: decoder on plain bf16 operands (faster; 1.3e-3 instead of < 1e-3) */
#define DE_PRECISION_BF16X3 2  /* DPT_Large / DPT_Hybrid: every Linear / convolution as bf16x3 (split3 operands, ~16 mantissa
                                * bits, 3x the tensor work); attention's QK^T / PV stay bf16.  For maps whose dynamic range
                                * is wide against their maximum, where bf16 operands alone cost > 1e-3 */

/* model_type: "DPT_Large" | "DPT_Hybrid" | "MiDaS_small" (the strings of depth_midas.py:91-93). */
int de_create(const char* model_type, int device, int precision, de_handle* out);
/* One tensor of the model's state_dict under its upstream key (e.g. "pretrained.model.blocks.0.attn.qkv.weight",
 * "scratch.refinenet1.out_conv.bias"; SURVEY.md 8a footer), host memory, dtype DE_DTYPE_F32 or DE_DTYPE_BF16.  The
 * data is copied.  Keys the forward pass never reads are accepted and ignored. */
int de_set_weight(de_handle h, const char* upstream_key, const void* host_ptr, const int64_t* shape, int ndim, int dtype);
/* Validates that every tensor the architecture needs is present with the right shape, folds and repacks, uploads.
 * Fails with the offending key in de_last_error(). */
int de_finalize_weights(de_handle h);
/* Prepares the geometry (B frames of H x W): Resize.get_size, work buffers, one eager warm-up pass (synchronous).
 * Optional - de_predict plans on first use - but required before a de_predict that is issued inside a stream capture. */
int de_plan(de_handle h, int B, int H, int W, int* net_h, int* net_w);

typedef struct DePredictIO {
  const uint8_t* frames;  /* device: uint8 BGR [B,H,W,3] as cv2 delivers it - or, with uv != NULL, the NV12 Y plane [B,H,W] */
  int64_t row_stride;     /* bytes; 0 = dense */
  int64_t frame_stride;   /* bytes; 0 = dense */
  const uint8_t* uv;      /* device: NV12 interleaved UV plane [B,H/2,W/2,2], or NULL for BGR input */
  int64_t uv_row_stride, uv_frame_stride; /* bytes; 0 = dense */
  float* out_map;         /* device float [B,H,W]: relative inverse depth at source resolution (depth_midas.py:28-34), or NULL */
  void* out_metric;       /* device bf16 [B,H,W] = map * metric_scale + metric_shift (face.py:137: 1/2.54, 0), or NULL */
  float metric_scale, metric_shift;
  uint8_t* out_vis;       /* device uint8 [B,H,W]: to_video_frame grey level (midas_mediapipe.py:141-150), or NULL */
  const int32_t* sample_rc; /* device int32 [B,n_samples,2] (row, col) into the output map, or NULL */
  int n_samples;
  float* out_samples;     /* device float [B,n_samples] */
} DePredictIO;
/* K1 -> network -> K8 for B frames, enqueued on `stream`, no synchronisation once the geometry is planned.
 * use_graph != 0: the launch sequence for this (geometry, IO pointers) is captured into a CUDA graph on first use
 * and replayed afterwards (keep the buffers stable).  When `stream` is itself being captured, the kernels are simply
 * recorded into the caller's graph. */
int de_predict(de_handle h, int B, int H, int W, const DePredictIO* io, int use_graph, void* stream);
/* The network alone (depth_midas.py:25-26): returns the plan's input buffer (what de_preprocess must write for this
 * model: de_engine_info layout) and its output buffer float [B,net_h,net_w]; run != 0 enqueues the forward pass. */
int de_forward(de_handle h, int B, int H, int W, void** net_input, float** pred, int* net_h, int* net_w, int run, void* stream);
/* layout: the de_preprocess layout this model consumes; bytes of packed weights / of all planned work buffers. */
int de_engine_info(de_handle h, int* layout, int64_t* weight_bytes, int64_t* work_bytes);
void de_destroy(de_handle h);

#ifdef __cplusplus
}
#endif
#endif /* DEPTH_B200_H_ */
