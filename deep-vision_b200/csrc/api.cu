// api.cu — the extern "C" surface of libdv_b200.so for the dense convolution, plus library-level entry points.
// (bn.cu, dwconv.cu and pool_misc.cu export their own entry points next to their kernels.)
#include "common.cuh"
#include <atomic>
#include <string.h>

namespace dv {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

int num_sms() {
  static int cached[16] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

static int check_conv(const dv_conv_params* p, const char* who) {
  DV_REQUIRE(p != nullptr, DV_ERR_ARG, "%s: params is null", who);
  DV_REQUIRE(p->x && p->w && p->y, DV_ERR_ARG, "%s: null tensor pointer", who);
  DV_REQUIRE(p->N > 0 && p->H > 0 && p->W > 0 && p->C > 0 && p->K > 0 && p->R > 0 && p->S > 0, DV_ERR_ARG, "%s: non-positive dimension", who);
  DV_REQUIRE(p->stride > 0 && p->pad >= 0, DV_ERR_ARG, "%s: bad stride/pad", who);
  DV_REQUIRE(p->H + 2 * p->pad >= p->R && p->W + 2 * p->pad >= p->S, DV_ERR_ARG, "%s: filter larger than padded input", who);
  DV_REQUIRE(p->dtype == DV_F32 || p->dtype == DV_BF16, DV_ERR_ARG, "%s: bad dtype %d", who, p->dtype);
  DV_REQUIRE(p->algo >= DV_ALGO_AUTO && p->algo <= DV_ALGO_TC, DV_ERR_ARG, "%s: bad algo %d", who, p->algo);
  return DV_OK;
}

static int resolve_algo(const dv_conv_params* p, int pass) {
  if (p->algo == DV_ALGO_SIMT) return DV_ALGO_SIMT;
  const char* why;
  const bool ok = conv_tc_supported(p, pass, &why);
  if (p->algo == DV_ALGO_TC) return DV_ALGO_TC;  // conv_tc reports the reason if unsupported
  return ok ? DV_ALGO_TC : DV_ALGO_SIMT;
}

static int run_conv(const dv_conv_params* p, int pass, void* stream, const char* who) {
  int rc = check_conv(p, who);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int algo = resolve_algo(p, pass);
  const bool folded = p->ep_scale || p->ep_shift || p->ep_relu || (p->ep_residual && pass != DV_CONV_DGRAD);
  DV_REQUIRE(!folded || (pass == DV_CONV_FPROP && algo == DV_ALGO_TC), DV_ERR_UNSUPPORTED,
             "%s: the folded-BatchNorm epilogue (ep_*) exists on the tensor-core fprop path only", who);
  DV_REQUIRE(p->ep_residual_mask == nullptr || (p->ep_residual != nullptr && pass == DV_CONV_DGRAD), DV_ERR_ARG,
             "%s: ep_residual_mask goes with ep_residual in dgrad", who);
  DV_REQUIRE(p->ep_residual == nullptr || pass != DV_CONV_DGRAD || algo == DV_ALGO_TC, DV_ERR_UNSUPPORTED,
             "%s: the shortcut-gradient epilogue (ep_residual in dgrad) exists on the tensor-core path only", who);
  DV_REQUIRE(p->bnb_x == nullptr || (pass == DV_CONV_DGRAD && algo == DV_ALGO_TC), DV_ERR_UNSUPPORTED,
             "%s: the BatchNorm-backward sums (bnb_*) exist on the tensor-core dgrad path only", who);
  return algo == DV_ALGO_TC ? conv_tc(p, pass, st) : conv_simt(p, pass, st);
}

}  // namespace dv

using namespace dv;

extern "C" int dv_conv_fprop(const dv_conv_params* p, void* stream) { return run_conv(p, DV_CONV_FPROP, stream, "dv_conv_fprop"); }
extern "C" int dv_conv_dgrad(const dv_conv_params* p, void* stream) { return run_conv(p, DV_CONV_DGRAD, stream, "dv_conv_dgrad"); }
extern "C" int dv_conv_wgrad(const dv_conv_params* p, void* stream) { return run_conv(p, DV_CONV_WGRAD, stream, "dv_conv_wgrad"); }

extern "C" int dv_conv_algo(const dv_conv_params* p, int pass) {
  if (!p) return DV_ALGO_SIMT;
  if (p->algo == DV_ALGO_SIMT) return DV_ALGO_SIMT;
  const char* why;
  return conv_tc_supported(p, pass, &why) ? DV_ALGO_TC : DV_ALGO_SIMT;
}

extern "C" size_t dv_conv_workspace_bytes(const dv_conv_params* p, int pass) {
  if (!p || p->N <= 0 || p->C <= 0 || p->K <= 0 || p->R <= 0 || p->S <= 0 || p->stride <= 0) return 0;
  if (p->algo == DV_ALGO_SIMT) return 0;
  const char* why;
  if (!conv_tc_supported(p, pass, &why)) return 0;
  return conv_tc_workspace(p, pass);
}

extern "C" int dv_version(void) { return DV_VERSION; }
extern "C" const char* dv_last_error(void) { return g_err; }
extern "C" uint64_t dv_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int dv_device_check(void) {
  int dev = 0;
  DV_CUDA_OK(cudaGetDevice(&dev));
  int major = 0, minor = 0;
  DV_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  DV_CUDA_OK(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  DV_REQUIRE(major == 10, DV_ERR_ARCH, "device %d is compute capability %d.%d; this library is built for sm_100a only", dev, major, minor);
  return DV_OK;
}
