// Host-side plumbing of the C ABI: error strings, launch counter, get_size and the cubic tap tables.
#include <stdlib.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>

#include <atomic>
#include <mutex>
#include <vector>

#include "depth_b200.h"
#include "host_util.h"

namespace deb {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
int cuda_fail(cudaError_t e, const char* what) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
  return DE_ERR_CUDA;
}
int cuda_check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, what);
  return DE_OK;
}
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
void add_launches(int64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---- per-operator profiler
struct ProfRec {
  char name[24];
  double flops;
  int64_t dims[4];
  cudaEvent_t e0, e1;
};
static std::atomic<bool> g_prof_on{false};
static std::mutex g_prof_mu;
static std::vector<ProfRec> g_prof;

ProfScope::ProfScope(const char* name, double flops, void* st, int64_t d0, int64_t d1, int64_t d2, int64_t d3) {
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(reinterpret_cast<cudaStream_t>(st), &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return;
  ProfRec r{};
  snprintf(r.name, sizeof(r.name), "%s", name);
  r.flops = flops;
  r.dims[0] = d0; r.dims[1] = d1; r.dims[2] = d2; r.dims[3] = d3;
  if (cudaEventCreate(&r.e0) != cudaSuccess || cudaEventCreate(&r.e1) != cudaSuccess) return;
  cudaEventRecord(r.e0, reinterpret_cast<cudaStream_t>(st));
  std::lock_guard<std::mutex> lk(g_prof_mu);
  idx = (int)g_prof.size();
  stream = st;
  g_prof.push_back(r);
}
ProfScope::~ProfScope() {
  if (idx < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (idx < (int)g_prof.size()) cudaEventRecord(g_prof[idx].e1, reinterpret_cast<cudaStream_t>(stream));
}

int num_sms() {
  static int cache[32] = {0};                  // per device
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 32 && cache[dev] > 0) return cache[dev];
  int n = 0;
  cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  if (n <= 0) n = 148;
  if (dev >= 0 && dev < 32) cache[dev] = n;
  return n;
}

bool pdl_enabled(int cls) {
  static int mask = -1;
  if (mask < 0) {
    const char* e = getenv("DEB_PDL");
    mask = e ? atoi(e) : 11;  // LayerNorm off: early-scheduled GEMM CTAs squat on the SMs and cost it occupancy (measured)
  }
  return (mask & cls) != 0;
}

TensorMapEncodeTiledFn get_tensor_map_encoder() {
  static TensorMapEncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
      set_error("cudaGetDriverEntryPoint(cuTensorMapEncodeTiled) failed: %s", cudaGetErrorString(e));
      return nullptr;
    }
    fn = reinterpret_cast<TensorMapEncodeTiledFn>(p);
  }
  return fn;
}

}  // namespace deb

extern "C" const char* de_last_error(void) { return deb::g_err; }
extern "C" int de_version(void) { return 1; }
extern "C" int64_t de_launch_count(void) { return deb::g_launches.load(); }

extern "C" int de_profile_start(void) {
  std::lock_guard<std::mutex> lk(deb::g_prof_mu);
  for (auto& r : deb::g_prof) {
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  deb::g_prof.clear();
  deb::g_prof_on.store(true);
  return DE_OK;
}
extern "C" int de_profile_stop(DeProfileRecord* out, int max_records, int* n_records) {
  deb::g_prof_on.store(false);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return deb::cuda_fail(e, "de_profile_stop: cudaDeviceSynchronize");
  std::lock_guard<std::mutex> lk(deb::g_prof_mu);
  int n = 0;
  for (auto& r : deb::g_prof) {
    float ms = 0.f;
    if (out && n < max_records && cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
      snprintf(out[n].name, sizeof(out[n].name), "%s", r.name);
      out[n].ms = ms;
      out[n].flops = r.flops;
      for (int k = 0; k < 4; ++k) out[n].dims[k] = r.dims[k];
      ++n;
    }
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  const int total = (int)deb::g_prof.size();
  deb::g_prof.clear();
  if (n_records) *n_records = out ? n : total;
  return DE_OK;
}

// [UPSTREAM] midas/transforms.py Resize.constrain_to_multiple_of / get_size (keep_aspect_ratio=True,
// ensure_multiple_of=32).  nearbyint() under the default rounding mode is round-half-to-even == np.round.
static int constrain32(double x, int max_val) {
  int y = (int)(nearbyint(x / 32.0) * 32.0);
  if (max_val > 0 && y > max_val) y = (int)(floor(x / 32.0) * 32.0);
  if (y < 0) y = (int)(ceil(x / 32.0) * 32.0);
  return y;
}
extern "C" int de_get_size(int width, int height, int target, int method, int* net_w, int* net_h) {
  if (width <= 0 || height <= 0 || target <= 0 || !net_w || !net_h) return deb::fail(DE_ERR_INVALID, "de_get_size: bad args");
  double sh = (double)target / height, sw = (double)target / width;
  if (method == 0) {  // upper_bound
    if (sw < sh) sh = sw; else sw = sh;
    *net_h = constrain32(sh * height, target);
    *net_w = constrain32(sw * width, target);
  } else if (method == 1) {  // minimal
    if (fabs(1 - sw) < fabs(1 - sh)) sh = sw; else sw = sh;
    *net_h = constrain32(sh * height, 0);
    *net_w = constrain32(sw * width, 0);
  } else {
    return deb::fail(DE_ERR_INVALID, "de_get_size: method %d", method);
  }
  return DE_OK;
}

// Keys cubic, A=-0.75, evaluated in float exactly like OpenCV interpolateCubic and ATen
// get_cubic_upsample_coefficients (same polynomial forms per library, see comments).
extern "C" int de_cubic_taps(int src, int dst, int mode, int32_t* idx, float* w) {
  if (src <= 0 || dst <= 0 || !idx || !w) return deb::fail(DE_ERR_INVALID, "de_cubic_taps: bad args");
  const float A = -0.75f;
  for (int d = 0; d < dst; ++d) {
    int s;
    float x;
    float c[4];
    if (mode == 0) {
      // OpenCV resize.cpp: scale = 1/(dst/src) in double; fx = (float)((d+0.5)*scale-0.5)
      const double inv_scale = (double)dst / (double)src;
      const double scale = 1.0 / inv_scale;
      const float f = (float)((d + 0.5) * scale - 0.5);
      s = (int)floorf(f);
      x = f - (float)s;
      c[0] = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A;
      c[1] = ((A + 2) * x - (A + 3)) * x * x + 1;
      c[2] = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1;
      c[3] = 1.f - c[0] - c[1] - c[2];
    } else {
      // ATen UpSample.h: scale = (float)src/dst; real = fma(scale, d+0.5f, -0.5f) (fp32, no clamp for cubic);
      // cubic_convolution2(x+1), cubic_convolution1(x), cubic_convolution1(1-x), cubic_convolution2(2-x)
      const float scale = (float)src / (float)dst;
      const float f = fmaf(scale, (float)d + 0.5f, -0.5f);  // ATen's kernels are built with FMA contraction
      s = (int)floorf(f);
      x = f - (float)s;
      const float x0 = x + 1.0f, x1 = x, x2 = 1.0f - x, x3 = 2.0f - x;
      c[0] = ((A * x0 - 5 * A) * x0 + 8 * A) * x0 - 4 * A;
      c[1] = ((A + 2) * x1 - (A + 3)) * x1 * x1 + 1;
      c[2] = ((A + 2) * x2 - (A + 3)) * x2 * x2 + 1;
      c[3] = ((A * x3 - 5 * A) * x3 + 8 * A) * x3 - 4 * A;
    }
    for (int k = 0; k < 4; ++k) {
      int i = s - 1 + k;
      i = i < 0 ? 0 : (i > src - 1 ? src - 1 : i);
      idx[d * 4 + k] = i;
      w[d * 4 + k] = c[k];
    }
  }
  return DE_OK;
}
