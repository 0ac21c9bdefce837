// dwconv.cu — depth-wise 3x3 convolution (pad 1, stride 1|2, groups == C) for MobileNetV1's DepthwiseConv
// (MobileNet/pytorch/models/mobilenet_v1.py:109-133).  Arithmetic intensity is ~3.5 FLOP/B, so the design target is
// the HBM roof: NHWC, one 16-byte channel vector per thread, consecutive threads on consecutive channels (coalesced
// 128 B lines), the 9 taps of a channel slab cached in shared memory, neighbouring pixels' halo served by L1/L2.
//   fprop : y[n,p,q,c]  = sum_{r,s} x[n, p*st+r-1, q*st+s-1, c] * w[c,r,s]
//   dgrad : dx[n,h,w,c] = sum_{r,s} dy[n, (h+1-r)/st, (w+1-s)/st, c] * w[c,r,s]      (divisible & in range)
//   wgrad : dw[c,r,s]   = sum_{n,p,q} dy[n,p,q,c] * x[n, p*st+r-1, q*st+s-1, c]      (block partials -> fixed-order reduce)
#include "common.cuh"

namespace dv {
namespace {

constexpr int kThreads = 256;

struct DwDims {
  int N, H, W, C, stride, P, Q;
};

struct DwPlan {
  int cg, cgb, lanes, cchunks;
};
static DwPlan dw_plan(int C, int V) {
  DwPlan p;
  p.cg = C / V;
  p.cgb = p.cg < 128 ? p.cg : 128;  // <= 1024 channels of taps in smem
  p.lanes = kThreads / p.cgb;
  p.cchunks = ceil_div(p.cg, p.cgb);
  return p;
}

// DGRAD == false: fprop over output pixels (P,Q).  DGRAD == true: dgrad over input pixels (H,W).
template <typename T, bool DGRAD>
__global__ void __launch_bounds__(kThreads)
dwconv_kernel(const T* __restrict__ in, const float* __restrict__ w, T* __restrict__ out, DwDims d, int cgb, int lanes,
              int64_t pix_per_blk) {
  constexpr int V = Vec16<T>::N;
  __shared__ float ws[9][128 * V];
  const int cg0 = blockIdx.y * cgb;
  const int nch = min(cgb * V, d.C - cg0 * V);
  for (int i = threadIdx.x; i < 9 * nch; i += kThreads) {
    const int ch = i / 9, tap = i % 9;
    ws[tap][ch] = w[(size_t)(cg0 * V + ch) * 9 + tap];
  }
  __syncthreads();
  const int tcg = threadIdx.x % cgb, lane = threadIdx.x / cgb;
  if (lane >= lanes || (cg0 + tcg) * V >= d.C) return;
  const int c = (cg0 + tcg) * V;
  const int OH = DGRAD ? d.H : d.P, OW = DGRAD ? d.W : d.Q;  // extent this kernel writes
  const int IH = DGRAD ? d.P : d.H, IW = DGRAD ? d.Q : d.W;  // extent it reads
  const int64_t npix = (int64_t)d.N * OH * OW;
  const int64_t p0 = (int64_t)blockIdx.x * pix_per_blk;
  const int64_t p1 = (p0 + pix_per_blk < npix) ? p0 + pix_per_blk : npix;
  for (int64_t pix = p0 + lane; pix < p1; pix += lanes) {
    const int n = (int)(pix / (OH * OW)), rem = (int)(pix % (OH * OW));
    const int oh = rem / OW, ow = rem % OW;
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      int ih;
      if (DGRAD) {
        const int t = oh + 1 - r;
        if (t < 0 || t % d.stride) continue;
        ih = t / d.stride;
      } else {
        ih = oh * d.stride + r - 1;
      }
      if ((unsigned)ih >= (unsigned)IH) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        int iw;
        if (DGRAD) {
          const int t = ow + 1 - s;
          if (t < 0 || t % d.stride) continue;
          iw = t / d.stride;
        } else {
          iw = ow * d.stride + s - 1;
        }
        if ((unsigned)iw >= (unsigned)IW) continue;
        float xv[V];
        Vec16<T> v; v.load(in + (((size_t)n * IH + ih) * IW + iw) * d.C + c); v.unpack(xv);
        const float* wt = &ws[r * 3 + s][tcg * V];
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] = fmaf(xv[j], wt[j], acc[j]);
      }
    }
    Vec16<T> o; o.pack(acc); o.store(out + (size_t)pix * d.C + c);
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
dwconv_wgrad_kernel(const T* __restrict__ x, const T* __restrict__ dy, float* __restrict__ partial, DwDims d, int cgb, int lanes,
                    int64_t pix_per_blk) {
  constexpr int V = Vec16<T>::N;
  __shared__ float red[kThreads * V];
  const int cg0 = blockIdx.y * cgb;
  const int tcg = threadIdx.x % cgb, lane = threadIdx.x / cgb;
  const bool active = lane < lanes && (cg0 + tcg) * V < d.C;
  const int c = (cg0 + tcg) * V;
  float acc[9][V];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < V; ++j) acc[t][j] = 0.f;
  const int64_t npix = (int64_t)d.N * d.P * d.Q;
  const int64_t p0 = (int64_t)blockIdx.x * pix_per_blk;
  const int64_t p1 = (p0 + pix_per_blk < npix) ? p0 + pix_per_blk : npix;
  if (active) {
    for (int64_t pix = p0 + lane; pix < p1; pix += lanes) {
      const int n = (int)(pix / (d.P * d.Q)), rem = (int)(pix % (d.P * d.Q));
      const int p = rem / d.Q, q = rem % d.Q;
      float g[V];
      Vec16<T> vg; vg.load(dy + (size_t)pix * d.C + c); vg.unpack(g);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int h = p * d.stride + r - 1;
        if ((unsigned)h >= (unsigned)d.H) continue;
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          const int w = q * d.stride + s - 1;
          if ((unsigned)w >= (unsigned)d.W) continue;
          float xv[V];
          Vec16<T> vx; vx.load(x + (((size_t)n * d.H + h) * d.W + w) * d.C + c); vx.unpack(xv);
#pragma unroll
          for (int j = 0; j < V; ++j) acc[r * 3 + s][j] = fmaf(xv[j], g[j], acc[r * 3 + s][j]);
        }
      }
    }
  }
  // reduce the pixel lanes of this block, one tap at a time
#pragma unroll
  for (int t = 0; t < 9; ++t) {
#pragma unroll
    for (int j = 0; j < V; ++j) red[threadIdx.x * V + j] = acc[t][j];
    __syncthreads();
    if (lane == 0 && active) {
      float s[V];
#pragma unroll
      for (int j = 0; j < V; ++j) s[j] = 0.f;
      for (int l = 0; l < lanes; ++l)
#pragma unroll
        for (int j = 0; j < V; ++j) s[j] += red[(l * cgb + tcg) * V + j];
      float* dst = partial + ((size_t)blockIdx.x * 9 + t) * d.C + c;
#pragma unroll
      for (int j = 0; j < V; ++j) dst[j] = s[j];
    }
    __syncthreads();
  }
}

// dw[c][tap] = sum_blk partial[blk][tap][c]
__global__ void dwconv_wgrad_finalize_kernel(const float* __restrict__ partial, float* __restrict__ dw, int nblk, int C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // over 9*C, c fastest
  if (i >= 9 * C) return;
  const int t = i / C, c = i % C;
  double s = 0.0;
  for (int b = 0; b < nblk; ++b) s += (double)partial[((size_t)b * 9 + t) * C + c];
  dw[(size_t)c * 9 + t] = (float)s;
}

static int wgrad_blocks(int64_t npix, int lanes) {
  int64_t nb = ceil_div64(npix, (int64_t)lanes * 16);
  const int64_t cap = (int64_t)num_sms() * 2;
  if (nb > cap) nb = cap;
  return (int)(nb < 1 ? 1 : nb);
}

static DwDims dims_of(const dv_dwconv_params* p) {
  DwDims d;
  d.N = p->N; d.H = p->H; d.W = p->W; d.C = p->C; d.stride = p->stride;
  d.P = (p->H + 2 - 3) / p->stride + 1;
  d.Q = (p->W + 2 - 3) / p->stride + 1;
  return d;
}

template <typename T, bool DGRAD>
static int run_fd(const dv_dwconv_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  DV_REQUIRE(p->C % V == 0, DV_ERR_UNSUPPORTED, "dwconv: C=%d must be a multiple of %d", p->C, V);
  DwDims d = dims_of(p);
  DwPlan pl = dw_plan(p->C, V);
  const int64_t npix = (int64_t)d.N * (DGRAD ? d.H * d.W : d.P * d.Q);
  int64_t nb = ceil_div64(npix, (int64_t)pl.lanes * 8);
  const int64_t cap = (int64_t)num_sms() * 16;
  if (nb > cap) nb = cap;
  const int64_t ppb = ceil_div64(npix, nb);
  nb = ceil_div64(npix, ppb);
  dim3 grid((unsigned)nb, (unsigned)pl.cchunks);
  const T* in = (const T*)(DGRAD ? p->y : p->x);
  T* out = (T*)(DGRAD ? p->x : p->y);
  dwconv_kernel<T, DGRAD><<<grid, kThreads, 0, st>>>(in, (const float*)p->w, out, d, pl.cgb, pl.lanes, ppb);
  DV_LAUNCH_OK("dwconv_kernel");
  return DV_OK;
}

template <typename T>
static int run_wgrad(const dv_dwconv_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  DV_REQUIRE(p->C % V == 0, DV_ERR_UNSUPPORTED, "dwconv: C=%d must be a multiple of %d", p->C, V);
  DwDims d = dims_of(p);
  DwPlan pl = dw_plan(p->C, V);
  const int64_t npix = (int64_t)d.N * d.P * d.Q;
  const int nb0 = wgrad_blocks(npix, pl.lanes);
  const int64_t ppb = ceil_div64(npix, nb0);
  const int nb = (int)ceil_div64(npix, ppb);
  const size_t need = (size_t)nb * 9 * p->C * sizeof(float);
  DV_REQUIRE(p->workspace && p->workspace_bytes >= need, DV_ERR_WORKSPACE, "dwconv_wgrad: workspace %zu bytes required, %zu given", need,
             p->workspace_bytes);
  dim3 grid((unsigned)nb, (unsigned)pl.cchunks);
  dwconv_wgrad_kernel<T><<<grid, kThreads, 0, st>>>((const T*)p->x, (const T*)p->y, (float*)p->workspace, d, pl.cgb, pl.lanes, ppb);
  DV_LAUNCH_OK("dwconv_wgrad_kernel");
  dwconv_wgrad_finalize_kernel<<<ceil_div(9 * p->C, 256), 256, 0, st>>>((const float*)p->workspace, (float*)p->w, nb, p->C);
  DV_LAUNCH_OK("dwconv_wgrad_finalize_kernel");
  return DV_OK;
}

static int check(const dv_dwconv_params* p, const char* who) {
  DV_REQUIRE(p && p->x && p->w && p->y, DV_ERR_ARG, "%s: null pointer", who);
  DV_REQUIRE(p->N > 0 && p->H > 0 && p->W > 0 && p->C > 0, DV_ERR_ARG, "%s: bad dims", who);
  DV_REQUIRE(p->stride == 1 || p->stride == 2, DV_ERR_UNSUPPORTED, "%s: stride must be 1 or 2", who);
  DV_REQUIRE(p->dtype == DV_F32 || p->dtype == DV_BF16, DV_ERR_ARG, "%s: bad dtype", who);
  return DV_OK;
}

}  // namespace
}  // namespace dv

using namespace dv;

extern "C" size_t dv_dwconv_workspace_bytes(const dv_dwconv_params* p) {
  if (!p || p->C <= 0) return 0;
  DwDims d = dims_of(p);
  const int64_t npix = (int64_t)d.N * d.P * d.Q;
  // the fp32 plan has at least as many pixel lanes per channel group as bf16, so size for the larger block count
  size_t best = 0;
  for (int V : {4, 8}) {
    if (p->C % V) continue;
    DwPlan pl = dw_plan(p->C, V);
    const int nb0 = wgrad_blocks(npix, pl.lanes);
    const int64_t ppb = ceil_div64(npix, nb0);
    const size_t nb = (size_t)ceil_div64(npix, ppb);
    const size_t need = nb * 9 * p->C * sizeof(float);
    if (need > best) best = need;
  }
  return best;
}
extern "C" int dv_dwconv_fprop(const dv_dwconv_params* p, void* stream) {
  int rc = check(p, "dv_dwconv_fprop");
  if (rc) return rc;
  return p->dtype == DV_BF16 ? run_fd<bf16, false>(p, (cudaStream_t)stream) : run_fd<float, false>(p, (cudaStream_t)stream);
}
extern "C" int dv_dwconv_dgrad(const dv_dwconv_params* p, void* stream) {
  int rc = check(p, "dv_dwconv_dgrad");
  if (rc) return rc;
  return p->dtype == DV_BF16 ? run_fd<bf16, true>(p, (cudaStream_t)stream) : run_fd<float, true>(p, (cudaStream_t)stream);
}
extern "C" int dv_dwconv_wgrad(const dv_dwconv_params* p, void* stream) {
  int rc = check(p, "dv_dwconv_wgrad");
  if (rc) return rc;
  return p->dtype == DV_BF16 ? run_wgrad<bf16>(p, (cudaStream_t)stream) : run_wgrad<float>(p, (cudaStream_t)stream);
}
