// pool_misc.cu — the small memory-bound ops around the conv-BN-ReLU stacks: stem max-pool, tail average-pool,
// layout change at the model entrance, casts, FC bias, filter preparation, optimisers.  All NHWC, 16-byte vectors
// over channels where the channel count allows it.
#include "common.cuh"
#include <math_constants.h>

namespace dv {
namespace {

constexpr int kThreads = 256;

static int pool_cap() {
  static const int v = getenv("DV_POOL_CAP") ? atoi(getenv("DV_POOL_CAP")) : 8;
  return v > 0 ? v : 8;
}

static unsigned grid_for(int64_t n, int per_sm = 8) {
  int64_t b = ceil_div64(n, kThreads);
  const int64_t cap = (int64_t)num_sms() * per_sm;
  return (unsigned)(b < cap ? (b < 1 ? 1 : b) : cap);
}

// ---- max-pool 3x3 stride 2 pad 1 (resnet50.py:34).  Padding behaves as -inf; ties keep the first position in
//      row-major window order, as torch's kernel does (strict '>' update). ---------------------------------------
// Forward: one thread = one 16-byte channel vector of TWO horizontally adjacent outputs.  Their windows share a column,
// so the 3x5 input patch is loaded once (15 independent 16-byte loads issued before any compare) instead of 2x9.
template <typename T>
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ idx, int N, int H, int W, int C, int P, int Q) {
  constexpr int V = Vec16<T>::N;
  const int cg = C / V;
  const int Q2 = (Q + 1) >> 1;
  const int total = N * P * Q2 * cg;  // < 2^31, checked by the caller
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = (i % cg) * V;
    int pix = i / cg;
    const int q0 = (pix % Q2) * 2; pix /= Q2;
    const int p = pix % P;
    const int n = pix / P;
    Vec16<T> in[3][5];
    bool ok[3][5];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = p * 2 + r - 1;
#pragma unroll
      for (int s = 0; s < 5; ++s) {
        const int w = q0 * 2 + s - 1;
        ok[r][s] = (unsigned)h < (unsigned)H && (unsigned)w < (unsigned)W;
        if (ok[r][s]) in[r][s].load(x + (((size_t)n * H + h) * W + w) * C + c);
      }
    }
#pragma unroll
    for (int o = 0; o < 2; ++o) {
      const int q = q0 + o;
      if (q >= Q) break;
      float best[V];
      __align__(8) uint8_t bi[V];
#pragma unroll
      for (int j = 0; j < V; ++j) { best[j] = -CUDART_INF_F; bi[j] = 0; }
      bool first = true;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          if (!ok[r][2 * o + s]) continue;
          float v[V];
          in[r][2 * o + s].unpack(v);
#pragma unroll
          for (int j = 0; j < V; ++j) {
            if (first || v[j] > best[j]) { best[j] = v[j]; bi[j] = (uint8_t)(r * 3 + s); }
          }
          first = false;
        }
      }
      const size_t off = (((size_t)n * P + p) * Q + q) * C + c;
      Vec16<T> out; out.pack(best); out.store(y + off);
      if (V == 8) *reinterpret_cast<uint2*>(idx + off) = *reinterpret_cast<const uint2*>(bi);   // off is a multiple of V
      else *reinterpret_cast<uint32_t*>(idx + off) = *reinterpret_cast<const uint32_t*>(bi);
    }
  }
}

// Backward: one thread = one channel vector of a 2x2 input cell {2p, 2p+1} x {2q, 2q+1}.  The cell is covered by the four
// windows (p..p+1, q..q+1) only, so 4 gradient + 4 index loads (all independent) produce 4 outputs; a thread per input
// element would load 2.25 windows per output through data-dependent loops.
template <typename T>
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ idx, T* __restrict__ dx, int N, int H, int W, int C, int P, int Q) {
  constexpr int V = Vec16<T>::N;
  const int cg = C / V;
  const int H2 = (H + 1) >> 1, W2 = (W + 1) >> 1;
  const int total = N * H2 * W2 * cg;  // < 2^31, checked by the caller
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = (i % cg) * V;
    int pix = i / cg;
    const int q = pix % W2; pix /= W2;
    const int p = pix % H2;
    const int n = pix / H2;
    float g[2][2][V];
    __align__(8) uint8_t ti[2][2][V];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const bool ok = (p + a) < P && (q + b) < Q;
        if (ok) {
          const size_t o = (((size_t)n * P + p + a) * Q + q + b) * C + c;
          Vec16<T> v; v.load(dy + o); v.unpack(g[a][b]);
          if (V == 8) *reinterpret_cast<uint2*>(ti[a][b]) = *reinterpret_cast<const uint2*>(idx + o);
          else *reinterpret_cast<uint32_t*>(ti[a][b]) = *reinterpret_cast<const uint32_t*>(idx + o);
        } else {
#pragma unroll
          for (int j = 0; j < V; ++j) { g[a][b][j] = 0.f; ti[a][b][j] = 255; }
        }
      }
    // input (2p+dh, 2q+dw) sits at tap (r,s) = (dh+1-2a, dw+1-2b) of window (p+a, q+b): dh=0 -> a=0 only (r=1);
    // dh=1 -> a=0 (r=2) and a=1 (r=0); same along w.
#pragma unroll
    for (int dh = 0; dh < 2; ++dh) {
      const int h = 2 * p + dh;
      if (h >= H) break;
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        const int w = 2 * q + dw;
        if (w >= W) break;
        float acc[V];
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] = 0.f;
#pragma unroll
        for (int a = 0; a <= dh; ++a)
#pragma unroll
          for (int b = 0; b <= dw; ++b) {
            const uint8_t tap = (uint8_t)((dh + 1 - 2 * a) * 3 + (dw + 1 - 2 * b));
#pragma unroll
            for (int j = 0; j < V; ++j) acc[j] += (ti[a][b][j] == tap) ? g[a][b][j] : 0.f;
          }
        Vec16<T> out; out.pack(acc); out.store(dx + (((size_t)n * H + h) * W + w) * C + c);
      }
    }
  }
}

// ---- global average pool (resnet50.py:43): y[n,c] = mean_hw x[n,hw,c] -----------------------------------------
template <typename T>
__global__ void avgpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int HW, int C) {
  constexpr int V = Vec16<T>::N;
  const int cg = C / V;
  const int64_t total = (int64_t)N * cg;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % cg) * V;
    const int n = (int)(i / cg);
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
    for (int p = 0; p < HW; ++p) {
      float v[V];
      Vec16<T> a; a.load(x + ((size_t)n * HW + p) * C + c); a.unpack(v);
#pragma unroll
      for (int j = 0; j < V; ++j) acc[j] += v[j];
    }
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] *= inv;
    Vec16<T> out; out.pack(acc); out.store(y + (size_t)n * C + c);
  }
}
template <typename T>
__global__ void avgpool_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int N, int HW, int C) {
  constexpr int V = Vec16<T>::N;
  const int cg = C / V;
  const int64_t total = (int64_t)N * HW * cg;
  const float inv = 1.f / (float)HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % cg) * V;
    const int n = (int)(i / ((int64_t)HW * cg));
    float g[V];
    Vec16<T> a; a.load(dy + (size_t)n * C + c); a.unpack(g);
#pragma unroll
    for (int j = 0; j < V; ++j) g[j] *= inv;
    a.pack(g); a.store(dx + (size_t)(i / cg) * C + c);
  }
}

// ---- NCHW fp32 -> NHWC T ------------------------------------------------------------------------------------------
template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, T* __restrict__ y, int N, int C, int H, int W) {
  const int64_t total = (int64_t)N * H * W;
  const int64_t hw = (int64_t)H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / hw, p = i % hw;
    for (int c = 0; c < C; ++c) y[i * C + c] = from_f32<T>(x[(n * C + c) * hw + p]);
  }
}

template <typename TS, typename TD>
__global__ void cast_kernel(const TS* __restrict__ s, TD* __restrict__ d, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    d[i] = from_f32<TD>(to_f32(s[i]));
}

__global__ void bias_add_kernel(float* __restrict__ y, const float* __restrict__ b, int64_t total, int C) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) y[i] += b[i % C];
}
__global__ void bias_grad_kernel(const float* __restrict__ dy, float* __restrict__ db, int64_t M, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float s = 0.f;
  for (int64_t m = 0; m < M; ++m) s += dy[m * C + c];
  db[c] = s;
}

// ---- filters: fp32 KRSC master -> compute dtype KRSC, and [C][R][S][K] with flipped taps for dgrad -----------------
template <typename T>
__global__ void filter_prep_kernel(const float* __restrict__ w, T* __restrict__ wo, T* __restrict__ wt, int K, int R, int S, int C) {
  const int64_t total = (int64_t)K * R * S * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = w[i];
    if (wo != nullptr) wo[i] = from_f32<T>(v);
    if (wt != nullptr) {
      int64_t t = i;
      const int c = (int)(t % C); t /= C;
      const int s = (int)(t % S); t /= S;
      const int r = (int)(t % R);
      const int k = (int)(t / R);
      wt[(((size_t)c * R + (R - 1 - r)) * S + (S - 1 - s)) * K + k] = from_f32<T>(v);
    }
  }
}

// ---- explicit patch matrix for the C=3 stems ---------------------------------------------------------------------------
// One CTA per output row (n,p): the R input rows it touches are staged in shared memory WITH their zero margins, so every
// filter row of a patch is S*C CONTIGUOUS staged elements and needs no bounds test.  ncu on the first version (7x7 stem, batch
// 256): issue slots 82 % busy, math-pipe throttled, 0.52 ms for a 1.2 GB write — bound by instruction count (a division per
// staged element, a table look-up per gathered element).  Now: rows are staged with 16-byte loads (the slot is laid out so the
// image row starts on a 16-byte boundary), a lane keeps one 16-byte group of patch columns for the whole row so its eight source
// offsets live in registers, and a warp takes 4 pixels x 8 groups per step (conflict-free gathers, whole-line stores).
template <typename T>
__global__ void __launch_bounds__(256)
im2col_kernel(const T* __restrict__ x, T* __restrict__ out, int N, int H, int W, int C, int R, int S, int stride, int pad, int P,
              int Q, int Kp, int lead, int Lp) {
  extern __shared__ __align__(16) unsigned char im2col_smem[];
  constexpr int V = 16 / (int)sizeof(T);
  T* rows = reinterpret_cast<T*>(im2col_smem);        // [R][Lp]: `lead` elements of slack, pad*C zeros, the row, zeros
  const int n = blockIdx.x / P, p = blockIdx.x % P;
  const int rsc = R * S * C, sc = S * C, wc = W * C;
  const int data0 = lead + pad * C;                    // element index of the image row inside a slot (multiple of V)
  for (int r = 0; r < R; ++r) {
    const int h = p * stride + r - pad;
    T* slot = rows + (size_t)r * Lp;
    const bool live = (unsigned)h < (unsigned)H;
    const T* src = x + ((size_t)n * H + (live ? h : 0)) * wc;
    if (live && (wc % V) == 0) {
      for (int i = threadIdx.x; i < wc / V; i += blockDim.x)
        *reinterpret_cast<uint4*>(slot + data0 + i * V) = *reinterpret_cast<const uint4*>(src + i * V);
      for (int e = threadIdx.x; e < data0; e += blockDim.x) slot[e] = from_f32<T>(0.f);
      for (int e = data0 + wc + threadIdx.x; e < Lp; e += blockDim.x) slot[e] = from_f32<T>(0.f);
    } else {
      for (int e = threadIdx.x; e < Lp; e += blockDim.x) {
        const int we = e - data0;
        slot[e] = (live && (unsigned)we < (unsigned)wc) ? src[we] : from_f32<T>(0.f);
      }
    }
  }
  // columns past R*S*C read from a zero region behind the slots (long enough for any pixel's offset), so the gathers are
  // unconditional; shared addresses are formed once as 32-bit shared-window offsets (the generic-address form cost an
  // S2R + LEA per predicated load)
  const int zero0 = R * Lp;
  for (int e = threadIdx.x; e < Q * stride * C + 8; e += blockDim.x) rows[zero0 + e] = from_f32<T>(0.f);
  __syncthreads();
  const int groups = Kp / 8, gtiles = (groups + 7) / 8, qtiles = (Q + 3) / 4;
  T* orow = out + ((size_t)n * P + p) * Q * Kp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(rows);
  for (int gt = 0; gt < gtiles; ++gt) {
    const int g = gt * 8 + (lane & 7);
    uint32_t off[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = g * 8 + j;
      off[j] = sbase + (uint32_t)sizeof(T) * (uint32_t)((g < groups && k < rsc) ? (k / sc) * Lp + lead + (k % sc) : zero0);
    }
    if (g >= groups) continue;
    for (int qt = warp; qt < qtiles; qt += nwarps) {
      const int q = qt * 4 + (lane >> 3);
      if (q >= Q) continue;
      const uint32_t base = (uint32_t)sizeof(T) * (uint32_t)(q * stride * C);
      T vals[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (sizeof(T) == 2) {
          unsigned short u;
          asm volatile("ld.shared.u16 %0, [%1];" : "=h"(u) : "r"(off[j] + base));
          reinterpret_cast<unsigned short*>(vals)[j] = u;
        } else {
          uint32_t u;
          asm volatile("ld.shared.u32 %0, [%1];" : "=r"(u) : "r"(off[j] + base));
          reinterpret_cast<uint32_t*>(vals)[j] = u;
        }
      }
      T* dst = orow + (size_t)q * Kp + g * 8;
      if (sizeof(T) == 2) {
        *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(vals);
      } else {
        *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(vals);
        *reinterpret_cast<uint4*>(dst + 4) = *reinterpret_cast<const uint4*>(vals + 4);
      }
    }
  }
}

template <typename T>
__global__ void pad_rows_kernel(const float* __restrict__ src, T* __restrict__ dst, int rows, int cols, int pitch) {
  const int64_t total = (int64_t)rows * pitch;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % pitch);
    const int64_t r = i / pitch;
    dst[i] = from_f32<T>(c < cols ? src[r * cols + c] : 0.f);
  }
}
__global__ void unpad_rows_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int cols, int pitch) {
  const int64_t total = (int64_t)rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % cols);
    const int64_t r = i / cols;
    dst[i] = src[r * pitch + c];
  }
}

// ---- cross entropy, one CTA per row (train.py:358,452) -------------------------------------------------------------
__global__ void __launch_bounds__(256)
cross_entropy_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, float* __restrict__ loss,
                     float* __restrict__ dlogits, int B, int C) {
  __shared__ float red[8];
  const int b = blockIdx.x;
  const float* row = logits + (size_t)b * C;
  float m = -CUDART_INF_F;
  for (int c = threadIdx.x; c < C; c += blockDim.x) m = fmaxf(m, row[c]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) m = fmaxf(m, red[i]);
  __syncthreads();
  float s = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s += expf(row[c] - m);
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += red[i];
  const int label = (int)labels[b];
  const float inv_b = 1.f / (float)B, inv_s = 1.f / s;
  for (int c = threadIdx.x; c < C; c += blockDim.x)
    dlogits[(size_t)b * C + c] = (expf(row[c] - m) * inv_s - (c == label ? 1.f : 0.f)) * inv_b;
  if (threadIdx.x == 0) atomicAdd(loss, (logf(s) + m - row[label]) * inv_b);
}
__global__ void scale_by_kernel(float* __restrict__ x, const float* __restrict__ s, int64_t n) {
  const float f = s[0];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) x[i] *= f;
}

// ---- optimisers (train.py:166-205) -----------------------------------------------------------------------------------
__global__ void sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ buf, bf16* __restrict__ p16, int64_t n,
                           float lr, float mom, float wd, float gscale, int first) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float pv = p[i];
    const float gv = fmaf(wd, pv, g[i] * gscale);
    const float b = first ? gv : fmaf(mom, buf[i], gv);
    buf[i] = b;
    const float pn = pv - lr * b;
    p[i] = pn;
    if (p16 != nullptr) p16[i] = __float2bfloat16_rn(pn);
  }
}
__global__ void rmsprop_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ sq, bf16* __restrict__ p16, int64_t n,
                               float lr, float alpha, float eps, float wd, float gscale) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float pv = p[i];
    const float gv = fmaf(wd, pv, g[i] * gscale);
    const float s = alpha * sq[i] + (1.f - alpha) * gv * gv;
    sq[i] = s;
    const float pn = pv - lr * gv / (sqrtf(s) + eps);
    p[i] = pn;
    if (p16 != nullptr) p16[i] = __float2bfloat16_rn(pn);
  }
}

}  // namespace
}  // namespace dv

using namespace dv;

#define DV_DISPATCH_T(dtype, CALL_BF, CALL_F32) \
  do {                                           \
    if ((dtype) == DV_BF16) { CALL_BF; } else { CALL_F32; } \
  } while (0)

static int vec_ok(int C, int dtype, const char* who) {
  const int V = dtype == DV_BF16 ? 8 : 4;
  DV_REQUIRE(C % V == 0, DV_ERR_UNSUPPORTED, "%s: C=%d must be a multiple of %d", who, C, V);
  return DV_OK;
}

extern "C" int dv_maxpool_fwd(const void* x, void* y, uint8_t* idx, int N, int H, int W, int C, int dtype, void* stream) {
  DV_REQUIRE(x && y && idx && N > 0 && H > 0 && W > 0 && C > 0, DV_ERR_ARG, "dv_maxpool_fwd: bad arguments");
  int rc = vec_ok(C, dtype, "dv_maxpool_fwd");
  if (rc) return rc;
  const int P = (H + 2 - 3) / 2 + 1, Q = (W + 2 - 3) / 2 + 1;
  const int64_t total = (int64_t)N * P * ((Q + 1) / 2) * (C / (dtype == DV_BF16 ? 8 : 4));  // two outputs per thread
  DV_REQUIRE((int64_t)N * H * W * C < (1LL << 31), DV_ERR_UNSUPPORTED, "dv_maxpool_fwd: tensor exceeds 2^31 elements");
  cudaStream_t st = (cudaStream_t)stream;
  DV_DISPATCH_T(dtype, (maxpool_fwd_kernel<bf16><<<grid_for(total, pool_cap()), kThreads, 0, st>>>((const bf16*)x, (bf16*)y, idx, N, H, W, C, P, Q)),
                (maxpool_fwd_kernel<float><<<grid_for(total), kThreads, 0, st>>>((const float*)x, (float*)y, idx, N, H, W, C, P, Q)));
  DV_LAUNCH_OK("maxpool_fwd_kernel");
  return DV_OK;
}
extern "C" int dv_maxpool_bwd(const void* dy, const uint8_t* idx, void* dx, int N, int H, int W, int C, int dtype, void* stream) {
  DV_REQUIRE(dy && dx && idx && N > 0 && H > 0 && W > 0 && C > 0, DV_ERR_ARG, "dv_maxpool_bwd: bad arguments");
  int rc = vec_ok(C, dtype, "dv_maxpool_bwd");
  if (rc) return rc;
  const int P = (H + 2 - 3) / 2 + 1, Q = (W + 2 - 3) / 2 + 1;
  const int64_t total = (int64_t)N * ((H + 1) / 2) * ((W + 1) / 2) * (C / (dtype == DV_BF16 ? 8 : 4));  // a 2x2 cell per thread
  DV_REQUIRE((int64_t)N * H * W * C < (1LL << 31), DV_ERR_UNSUPPORTED, "dv_maxpool_bwd: tensor exceeds 2^31 elements");
  cudaStream_t st = (cudaStream_t)stream;
  DV_DISPATCH_T(dtype, (maxpool_bwd_kernel<bf16><<<grid_for(total, pool_cap()), kThreads, 0, st>>>((const bf16*)dy, idx, (bf16*)dx, N, H, W, C, P, Q)),
                (maxpool_bwd_kernel<float><<<grid_for(total), kThreads, 0, st>>>((const float*)dy, idx, (float*)dx, N, H, W, C, P, Q)));
  DV_LAUNCH_OK("maxpool_bwd_kernel");
  return DV_OK;
}
extern "C" int dv_avgpool_fwd(const void* x, void* y, int N, int HW, int C, int dtype, void* stream) {
  DV_REQUIRE(x && y && N > 0 && HW > 0 && C > 0, DV_ERR_ARG, "dv_avgpool_fwd: bad arguments");
  int rc = vec_ok(C, dtype, "dv_avgpool_fwd");
  if (rc) return rc;
  const int64_t total = (int64_t)N * (C / (dtype == DV_BF16 ? 8 : 4));
  cudaStream_t st = (cudaStream_t)stream;
  DV_DISPATCH_T(dtype, (avgpool_fwd_kernel<bf16><<<grid_for(total), kThreads, 0, st>>>((const bf16*)x, (bf16*)y, N, HW, C)),
                (avgpool_fwd_kernel<float><<<grid_for(total), kThreads, 0, st>>>((const float*)x, (float*)y, N, HW, C)));
  DV_LAUNCH_OK("avgpool_fwd_kernel");
  return DV_OK;
}
extern "C" int dv_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, int dtype, void* stream) {
  DV_REQUIRE(dy && dx && N > 0 && HW > 0 && C > 0, DV_ERR_ARG, "dv_avgpool_bwd: bad arguments");
  int rc = vec_ok(C, dtype, "dv_avgpool_bwd");
  if (rc) return rc;
  const int64_t total = (int64_t)N * HW * (C / (dtype == DV_BF16 ? 8 : 4));
  cudaStream_t st = (cudaStream_t)stream;
  DV_DISPATCH_T(dtype, (avgpool_bwd_kernel<bf16><<<grid_for(total), kThreads, 0, st>>>((const bf16*)dy, (bf16*)dx, N, HW, C)),
                (avgpool_bwd_kernel<float><<<grid_for(total), kThreads, 0, st>>>((const float*)dy, (float*)dx, N, HW, C)));
  DV_LAUNCH_OK("avgpool_bwd_kernel");
  return DV_OK;
}
extern "C" int dv_nchw_to_nhwc(const float* x, void* y, int N, int C, int H, int W, int dtype, void* stream) {
  DV_REQUIRE(x && y && N > 0 && H > 0 && W > 0 && C > 0, DV_ERR_ARG, "dv_nchw_to_nhwc: bad arguments");
  const int64_t total = (int64_t)N * H * W;
  cudaStream_t st = (cudaStream_t)stream;
  DV_DISPATCH_T(dtype, (nchw_to_nhwc_kernel<bf16><<<grid_for(total), kThreads, 0, st>>>(x, (bf16*)y, N, C, H, W)),
                (nchw_to_nhwc_kernel<float><<<grid_for(total), kThreads, 0, st>>>(x, (float*)y, N, C, H, W)));
  DV_LAUNCH_OK("nchw_to_nhwc_kernel");
  return DV_OK;
}
extern "C" int dv_cast(const void* src, int sd, void* dst, int dd, int64_t n, void* stream) {
  DV_REQUIRE(src && dst && n > 0, DV_ERR_ARG, "dv_cast: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned g = grid_for(n);
  if (sd == DV_F32 && dd == DV_BF16) cast_kernel<float, bf16><<<g, kThreads, 0, st>>>((const float*)src, (bf16*)dst, n);
  else if (sd == DV_BF16 && dd == DV_F32) cast_kernel<bf16, float><<<g, kThreads, 0, st>>>((const bf16*)src, (float*)dst, n);
  else if (sd == DV_F32 && dd == DV_F32) cast_kernel<float, float><<<g, kThreads, 0, st>>>((const float*)src, (float*)dst, n);
  else if (sd == DV_BF16 && dd == DV_BF16) cast_kernel<bf16, bf16><<<g, kThreads, 0, st>>>((const bf16*)src, (bf16*)dst, n);
  else { set_error("dv_cast: bad dtypes %d -> %d", sd, dd); return DV_ERR_ARG; }
  DV_LAUNCH_OK("cast_kernel");
  return DV_OK;
}
extern "C" int dv_bias_add(float* y, const float* bias, int64_t M, int C, void* stream) {
  DV_REQUIRE(y && bias && M > 0 && C > 0, DV_ERR_ARG, "dv_bias_add: bad arguments");
  bias_add_kernel<<<grid_for(M * C), kThreads, 0, (cudaStream_t)stream>>>(y, bias, M * C, C);
  DV_LAUNCH_OK("bias_add_kernel");
  return DV_OK;
}
extern "C" int dv_bias_grad(const float* dy, float* dbias, int64_t M, int C, void* stream) {
  DV_REQUIRE(dy && dbias && M > 0 && C > 0, DV_ERR_ARG, "dv_bias_grad: bad arguments");
  bias_grad_kernel<<<ceil_div(C, 128), 128, 0, (cudaStream_t)stream>>>(dy, dbias, M, C);
  DV_LAUNCH_OK("bias_grad_kernel");
  return DV_OK;
}
extern "C" int dv_filter_prep(const float* w, void* wo, void* wt, int K, int R, int S, int C, int dtype, void* stream) {
  DV_REQUIRE(w && (wo || wt) && K > 0 && R > 0 && S > 0 && C > 0, DV_ERR_ARG, "dv_filter_prep: bad arguments");
  const int64_t total = (int64_t)K * R * S * C;
  cudaStream_t st = (cudaStream_t)stream;
  DV_DISPATCH_T(dtype, (filter_prep_kernel<bf16><<<grid_for(total), kThreads, 0, st>>>(w, (bf16*)wo, (bf16*)wt, K, R, S, C)),
                (filter_prep_kernel<float><<<grid_for(total), kThreads, 0, st>>>(w, (float*)wo, (float*)wt, K, R, S, C)));
  DV_LAUNCH_OK("filter_prep_kernel");
  return DV_OK;
}
extern "C" int dv_im2col(const void* x, void* patches, int N, int H, int W, int C, int R, int S, int stride, int pad, int Kp, int dtype,
                         void* stream) {
  DV_REQUIRE(x && patches && N > 0 && H > 0 && W > 0 && C > 0 && R > 0 && S > 0 && stride > 0 && pad >= 0, DV_ERR_ARG, "dv_im2col: bad arguments");
  DV_REQUIRE(Kp % 8 == 0 && Kp >= R * S * C, DV_ERR_ARG, "dv_im2col: Kp=%d must be a multiple of 8 and >= R*S*C=%d", Kp, R * S * C);
  const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t esz = dtype == DV_BF16 ? 2 : 4;
  const int vec = (int)(16 / esz);
  const int lead = (vec - (pad * C) % vec) % vec;                       // slack so that the image row starts on a 16-byte boundary
  const int Lp = ceil_div(lead + (W + 2 * pad) * C, vec) * vec;         // slot length, a whole number of 16-byte vectors
  const size_t smem = ((size_t)R * Lp + (size_t)Q * stride * C + 8) * esz;   // R staged rows + the zero region
  DV_REQUIRE(smem <= 48 * 1024, DV_ERR_UNSUPPORTED, "dv_im2col: %zu bytes of staging exceed 48 KB (R*(W+2*pad)*C too large)", smem);
  DV_REQUIRE((int64_t)N * P <= 0x7fffffff, DV_ERR_UNSUPPORTED, "dv_im2col: too many output rows");
  DV_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(patches) & 15) == 0, DV_ERR_ARG,
             "dv_im2col: pointers must be 16-byte aligned");
  DV_DISPATCH_T(dtype, (im2col_kernel<bf16><<<(unsigned)(N * P), 256, smem, st>>>((const bf16*)x, (bf16*)patches, N, H, W, C, R, S, stride, pad, P, Q, Kp, lead, Lp)),
                (im2col_kernel<float><<<(unsigned)(N * P), 256, smem, st>>>((const float*)x, (float*)patches, N, H, W, C, R, S, stride, pad, P, Q, Kp, lead, Lp)));
  DV_LAUNCH_OK("im2col_kernel");
  return DV_OK;
}
extern "C" int dv_pad_rows(const float* src, void* dst, int rows, int cols, int dst_pitch, int dtype, void* stream) {
  DV_REQUIRE(src && dst && rows > 0 && cols > 0 && dst_pitch >= cols, DV_ERR_ARG, "dv_pad_rows: bad arguments");
  const int64_t total = (int64_t)rows * dst_pitch;
  cudaStream_t st = (cudaStream_t)stream;
  DV_DISPATCH_T(dtype, (pad_rows_kernel<bf16><<<grid_for(total), kThreads, 0, st>>>(src, (bf16*)dst, rows, cols, dst_pitch)),
                (pad_rows_kernel<float><<<grid_for(total), kThreads, 0, st>>>(src, (float*)dst, rows, cols, dst_pitch)));
  DV_LAUNCH_OK("pad_rows_kernel");
  return DV_OK;
}
extern "C" int dv_unpad_rows(const float* src, float* dst, int rows, int cols, int src_pitch, void* stream) {
  DV_REQUIRE(src && dst && rows > 0 && cols > 0 && src_pitch >= cols, DV_ERR_ARG, "dv_unpad_rows: bad arguments");
  unpad_rows_kernel<<<grid_for((int64_t)rows * cols), kThreads, 0, (cudaStream_t)stream>>>(src, dst, rows, cols, src_pitch);
  DV_LAUNCH_OK("unpad_rows_kernel");
  return DV_OK;
}
extern "C" int dv_cross_entropy(const float* logits, const int64_t* labels, float* loss, float* dlogits, int B, int C, void* stream) {
  DV_REQUIRE(logits && labels && loss && dlogits && B > 0 && C > 0, DV_ERR_ARG, "dv_cross_entropy: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  DV_CUDA_OK(cudaMemsetAsync(loss, 0, sizeof(float), st));
  cross_entropy_kernel<<<B, 256, 0, st>>>(logits, labels, loss, dlogits, B, C);
  DV_LAUNCH_OK("cross_entropy_kernel");
  return DV_OK;
}
extern "C" int dv_scale_by(float* x, const float* s, int64_t n, void* stream) {
  DV_REQUIRE(x && s && n > 0, DV_ERR_ARG, "dv_scale_by: bad arguments");
  scale_by_kernel<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(x, s, n);
  DV_LAUNCH_OK("scale_by_kernel");
  return DV_OK;
}
extern "C" int dv_sgd_step(float* p, const float* g, float* buf, int64_t n, float lr, float momentum, float wd, float gscale, int first,
                           void* stream) {
  DV_REQUIRE(p && g && buf && n > 0, DV_ERR_ARG, "dv_sgd_step: bad arguments");
  sgd_kernel<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(p, g, buf, nullptr, n, lr, momentum, wd, gscale, first);
  DV_LAUNCH_OK("sgd_kernel");
  return DV_OK;
}
extern "C" int dv_sgd_step_bf16(float* p, const float* g, float* buf, void* p16, int64_t n, float lr, float momentum, float wd, float gscale,
                                int first, void* stream) {
  DV_REQUIRE(p && g && buf && p16 && n > 0, DV_ERR_ARG, "dv_sgd_step_bf16: bad arguments");
  sgd_kernel<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(p, g, buf, (bf16*)p16, n, lr, momentum, wd, gscale, first);
  DV_LAUNCH_OK("sgd_kernel");
  return DV_OK;
}
extern "C" int dv_rmsprop_step(float* p, const float* g, float* sq, int64_t n, float lr, float alpha, float eps, float wd, float gscale,
                               void* stream) {
  DV_REQUIRE(p && g && sq && n > 0, DV_ERR_ARG, "dv_rmsprop_step: bad arguments");
  rmsprop_kernel<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(p, g, sq, nullptr, n, lr, alpha, eps, wd, gscale);
  DV_LAUNCH_OK("rmsprop_kernel");
  return DV_OK;
}
extern "C" int dv_rmsprop_step_bf16(float* p, const float* g, float* sq, void* p16, int64_t n, float lr, float alpha, float eps, float wd,
                                    float gscale, void* stream) {
  DV_REQUIRE(p && g && sq && p16 && n > 0, DV_ERR_ARG, "dv_rmsprop_step_bf16: bad arguments");
  rmsprop_kernel<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(p, g, sq, (bf16*)p16, n, lr, alpha, eps, wd, gscale);
  DV_LAUNCH_OK("rmsprop_kernel");
  return DV_OK;
}
