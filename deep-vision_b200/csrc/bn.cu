// bn.cu — batch normalisation over [M rows][C channels] (NHWC activations flattened), the HBM-bound half of the
// conv-BN-ReLU block.  All passes stream 16-byte vectors (8 bf16 / 4 fp32 channels per thread, consecutive threads on
// consecutive channels => fully coalesced 128-byte lines) and keep per-channel statistics in fp32.
//
//   stats      : per-channel shifted sums  S1 = sum(x-k), S2 = sum((x-k)^2), k = x[0][c]  -> block partials in scratch
//                finalize: mean, biased var -> invstd, running stats (unbiased var)          [2 reads of nothing extra]
//   apply      : y = relu((x-mean)*invstd*gamma + beta [+ residual])
//   bwd_reduce : dbeta = sum(dz), dgamma = sum(dz*xhat), dz = dy * (y > 0)
//   bwd_apply  : dx = gamma*invstd*(dz - dbeta/M - xhat*dgamma/M);  dres = dz
// Partials are reduced in a fixed order, so every result is bit-reproducible run to run.  The ReLU mask of the backward passes is
// recomputed from x (no identity), read from the 1-bit mask the apply pass wrote (identity added), or taken from a stored y.
// Grids are one wave of resident CTAs (occupancy query per kernel): the backward kernels hold 114-128 registers.
#include "common.cuh"
#ifndef DV_BN_RED_U
#define DV_BN_RED_U 6
#endif
#include <map>
#include <mutex>
#include <mutex>

namespace dv {
namespace {

static int env_or(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && atoi(v) > 0 ? atoi(v) : dflt;
}

constexpr int kThreads = 256;
constexpr int kMaxBlocks = 1184;

// CTAs of `kernel` one SM holds at kThreads threads (register-limited: 2 for the backward kernels, 3-4 for the forward ones).
// Grids are sized to ONE wave of resident CTAs: with the 8-per-SM grids these kernels started with, three quarters of the
// CTAs of bn_bwd_apply began only after an earlier CTA had finished (measured in the captured step: 18.87 -> 18.40 ms).
template <typename K>
static int resident_ctas(K kernel) {
  static std::mutex mu;   // keyed on the pointer: K is a function-pointer TYPE shared by instantiations of equal signature
  static std::map<const void*, int> table;
  std::lock_guard<std::mutex> lock(mu);
  auto it = table.find((const void*)kernel);
  if (it != table.end()) return it->second;
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, 0) != cudaSuccess || n < 1) n = 2;
  if (n > 8) n = 8;
  table[(const void*)kernel] = n;
  return n;
}  // 8 x 148 SMs: every SM fully occupied (2048 threads)

struct Plan {
  int cg;        // channel groups (C / V)
  int cgb;       // channel groups handled concurrently by a block (<= 256)
  int rows_par;  // rows handled concurrently by a block
  int nblk;
  int64_t rows_per_blk;
};

static Plan make_plan(int64_t M, int C, int V) {
  Plan p;
  p.cg = C / V;
  p.cgb = p.cg < kThreads ? p.cg : kThreads;
  p.rows_par = kThreads / p.cgb;
  int64_t nb = ceil_div64(M, (int64_t)p.rows_par * 32);
  // wide layers: a block already spans all C channels, so fewer, longer blocks keep the bytes in flight and shrink the
  // partial array the finalize pass has to read (nblk x 2 x C floats)
  const int max_blocks = C >= 512 ? kMaxBlocks / 4 : (C >= 256 ? kMaxBlocks / 2 : kMaxBlocks);
  if (nb > max_blocks) nb = max_blocks;
  if (nb < 1) nb = 1;
  p.rows_per_blk = ceil_div64(M, nb);
  p.nblk = (int)ceil_div64(M, p.rows_per_blk);
  return p;
}

// fewer, longer row ranges so that the grid is one wave of resident CTAs (workspace sized for the uncapped plan still fits)
static void cap_plan(Plan& p, int64_t M, int max_blocks) {
  static const int forced = env_or("DV_BN_RED_BLOCKS", 0);
  if (forced) max_blocks = forced;
  if (p.nblk > max_blocks && max_blocks > 0) {
    p.rows_per_blk = ceil_div64(M, max_blocks);
    p.nblk = (int)ceil_div64(M, p.rows_per_blk);
  }
}

// MODE 0: forward statistics of x.  MODE 1: backward sums (dz, dz*xhat).  RELU (compile time, so that the registers of the
// variants not taken are not reserved): see below.
template <typename T, int MODE, int RELU>
__global__ void __launch_bounds__(kThreads)
bn_reduce_kernel(const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ dy, const float* __restrict__ mean,
                 const float* __restrict__ invstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                 float* __restrict__ partial, int64_t M, int C, int cgb, int rows_par, int64_t rows_per_blk,
                 float* __restrict__ atomic_s1 = nullptr, float* __restrict__ atomic_s2 = nullptr,
                 const uint8_t* __restrict__ mask = nullptr) {
  // relu: 0 = no activation, 1 = mask from the stored output y, 2 = mask recomputed from x exactly as bn_apply computed it,
  //       3 = mask from the bit mask bn_apply wrote (one byte per 16-byte bf16 vector)
  constexpr int V = Vec16<T>::N;
  constexpr int relu = RELU;
  __shared__ float red[2][kThreads * V];
  const int tcg = threadIdx.x % cgb, trow = threadIdx.x / cgb;
  const bool active = trow < rows_par;
  const int cg = C / V;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_blk;
  const int64_t r1 = (r0 + rows_per_blk < M) ? r0 + rows_per_blk : M;

  for (int g0 = 0; g0 < cg; g0 += cgb) {
    const int g = g0 + tcg;
    const int c = g * V;
    float a1[V], a2[V], k0[V], k1[V], msc[V], msh[V];
#pragma unroll
    for (int i = 0; i < V; ++i) { a1[i] = 0.f; a2[i] = 0.f; k0[i] = 0.f; k1[i] = 1.f; msc[i] = 0.f; msh[i] = 0.f; }
    if (active && g < cg) {
      if (MODE == 0) {
        Vec16<T> v0; v0.load(x + c);  // row 0 as the shift
        v0.unpack(k0);
      } else {
#pragma unroll
        for (int i = 0; i < V; ++i) { k0[i] = mean[c + i]; k1[i] = invstd[c + i]; }
        if (relu == 2) {
#pragma unroll
          for (int i = 0; i < V; ++i) { msc[i] = k1[i] * gamma[c + i]; msh[i] = beta[c + i] - k0[i] * msc[i]; }
        }
      }
      // U independent 16-byte loads per stream in flight per thread (memory-level parallelism), then the arithmetic
      constexpr int U = (MODE == 1 && RELU != 1) ? DV_BN_RED_U : 4;
      for (int64_t r = r0 + trow; r < r1; r += (int64_t)rows_par * U) {
        Vec16<T> vx[U], vg[U], vy[U];
        uint32_t mb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t rr = r + (int64_t)u * rows_par;
          if (rr < r1) {
            const size_t off = (size_t)rr * C + c;
            vx[u].load(x + off);
            if (MODE == 1) {
              vg[u].load(dy + off);
              if (relu == 1) vy[u].load(y + off);
              if (relu == 3) mb[u] = mask[(size_t)rr * cg + g];
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t rr = r + (int64_t)u * rows_par;
          if (rr >= r1) break;
          float xv[V];
          vx[u].unpack(xv);
          if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < V; ++i) { const float d = xv[i] - k0[i]; a1[i] += d; a2[i] = fmaf(d, d, a2[i]); }
          } else {
            float gv[V];
            vg[u].unpack(gv);
            if (relu == 1) {
              float yv[V];
              vy[u].unpack(yv);
#pragma unroll
              for (int i = 0; i < V; ++i) gv[i] = yv[i] > 0.f ? gv[i] : 0.f;
            } else if (relu == 2) {
#pragma unroll
              for (int i = 0; i < V; ++i) gv[i] = fmaf(xv[i], msc[i], msh[i]) > 0.f ? gv[i] : 0.f;
            } else if (relu == 3) {
#pragma unroll
              for (int i = 0; i < V; ++i) gv[i] = ((mb[u] >> i) & 1u) ? gv[i] : 0.f;
            }
#pragma unroll
            for (int i = 0; i < V; ++i) { a1[i] += gv[i]; a2[i] = fmaf(gv[i], (xv[i] - k0[i]) * k1[i], a2[i]); }
          }
        }
      }
    }
    // reduce over the rows_par row lanes of this block
#pragma unroll
    for (int i = 0; i < V; ++i) { red[0][threadIdx.x * V + i] = a1[i]; red[1][threadIdx.x * V + i] = a2[i]; }
    __syncthreads();
    if (trow == 0 && g < cg) {
      float s1[V], s2[V];
#pragma unroll
      for (int i = 0; i < V; ++i) { s1[i] = 0.f; s2[i] = 0.f; }
      for (int rr = 0; rr < rows_par; ++rr) {
        const int t = rr * cgb + tcg;
#pragma unroll
        for (int i = 0; i < V; ++i) { s1[i] += red[0][t * V + i]; s2[i] += red[1][t * V + i]; }
      }
      if (atomic_s1 != nullptr) {
#pragma unroll
        for (int i = 0; i < V; ++i) { atomicAdd(atomic_s1 + c + i, s1[i]); atomicAdd(atomic_s2 + c + i, s2[i]); }
      } else {
        float* p1 = partial + ((size_t)blockIdx.x * 2) * C + c;
        float* p2 = p1 + C;
#pragma unroll
        for (int i = 0; i < V; ++i) { p1[i] = s1[i]; p2[i] = s2[i]; }
      }
    }
    __syncthreads();
  }
}

// Finalize: 32 channels x 16 partial-lanes per CTA.  Each thread sums every 16th block partial (fp64), the 16 lanes are
// combined through shared memory, lane 0 writes the per-channel results.
constexpr int kFinLanes = 16;

__device__ __forceinline__ void finalize_sums(const float* __restrict__ partial, int nblk, int C, int c, double& s1, double& s2) {
  __shared__ double sh[2][kFinLanes][32];
  const int cx = threadIdx.x & 31, ly = threadIdx.x >> 5;
  double a1 = 0.0, a2 = 0.0;
  if (c < C) {
    for (int b = ly; b < nblk; b += kFinLanes) {
      a1 += (double)partial[((size_t)b * 2) * C + c];
      a2 += (double)partial[((size_t)b * 2 + 1) * C + c];
    }
  }
  sh[0][ly][cx] = a1;
  sh[1][ly][cx] = a2;
  __syncthreads();
  s1 = 0.0; s2 = 0.0;
  if (ly == 0) {
#pragma unroll
    for (int l = 0; l < kFinLanes; ++l) { s1 += sh[0][l][cx]; s2 += sh[1][l][cx]; }
  }
}

template <typename T>
__global__ void __launch_bounds__(32 * kFinLanes)
bn_stats_finalize_kernel(const T* __restrict__ x, const float* __restrict__ partial, int nblk, int C, int64_t M,
                         float eps, float momentum, float* __restrict__ save_mean, float* __restrict__ save_invstd,
                         float* __restrict__ running_mean, float* __restrict__ running_var) {
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  double s1, s2;
  finalize_sums(partial, nblk, C, c, s1, s2);
  if ((threadIdx.x >> 5) != 0 || c >= C) return;
  const double k = (double)to_f32(x[c]);
  const double n = (double)M;
  const double m1 = s1 / n;
  double var = s2 / n - m1 * m1;
  if (var < 0.0) var = 0.0;
  const double mean = k + m1;
  save_mean[c] = (float)mean;
  save_invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
  if (running_mean != nullptr) {
    const double unbiased = M > 1 ? var * n / (n - 1.0) : var;
    running_mean[c] = (float)((1.0 - momentum) * (double)running_mean[c] + momentum * mean);
    running_var[c] = (float)((1.0 - momentum) * (double)running_var[c] + momentum * unbiased);
  }
}

// statistics from sums accumulated by the convolution epilogue; clears the sums for the next step
__global__ void bn_from_sums_kernel(float* __restrict__ sums, int C, int64_t M, float eps, float momentum, float* __restrict__ save_mean,
                                    float* __restrict__ save_invstd, float* __restrict__ running_mean, float* __restrict__ running_var) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double s1 = (double)sums[c], s2 = (double)sums[C + c];
  sums[c] = 0.f;
  sums[C + c] = 0.f;
  const double n = (double)M;
  const double mean = s1 / n;
  double var = s2 / n - mean * mean;
  if (var < 0.0) var = 0.0;
  save_mean[c] = (float)mean;
  save_invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
  if (running_mean != nullptr) {
    const double unbiased = M > 1 ? var * n / (n - 1.0) : var;
    running_mean[c] = (float)((1.0 - momentum) * (double)running_mean[c] + momentum * mean);
    running_var[c] = (float)((1.0 - momentum) * (double)running_var[c] + momentum * unbiased);
  }
}

__global__ void __launch_bounds__(32 * kFinLanes)
bn_bwd_finalize_kernel(const float* __restrict__ partial, int nblk, int C, float* __restrict__ dgamma, float* __restrict__ dbeta) {
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  double s1, s2;
  finalize_sums(partial, nblk, C, c, s1, s2);
  if ((threadIdx.x >> 5) != 0 || c >= C) return;
  dbeta[c] = (float)s1;
  dgamma[c] = (float)s2;
}

// The grid-stride is a multiple of the number of channel groups whenever C/V divides the CTA size (every channel count
// on the path: C/V is a power of two <= 256), so a thread meets the SAME channels on every iteration: the per-channel
// constants are folded into registers once (FIXED) instead of being re-read per 16-byte vector, and the loop is
// unrolled four deep so each thread keeps 4-12 independent 16-byte loads in flight.
constexpr int kUnroll = 4;

template <typename T, bool FIXED>
__global__ void __launch_bounds__(kThreads)
bn_apply_kernel(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y, const float* __restrict__ gamma,
                const float* __restrict__ beta, const float* __restrict__ mean, const float* __restrict__ invstd_or_var,
                int64_t nvec, int C, int relu, int eval_mode, float eps, uint8_t* __restrict__ mask, int reverse) {
  // reverse: walk the tensor from its END.  The producer of x (a convolution's epilogue, walking tiles upwards) leaves the
  // LAST bytes it wrote in the L2; a consumer that starts there meets them before they are evicted, and leaves the START of
  // y in the L2 for the next convolution, which walks upwards again.
  constexpr int V = Vec16<T>::N;
  const int cg = C / V;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t last = nvec - 1;
  float sc[V], sh[V];
  auto load_consts = [&](int c) {
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const float is = eval_mode ? rsqrtf(invstd_or_var[c + j] + eps) : invstd_or_var[c + j];
      sc[j] = is * gamma[c + j];
      sh[j] = beta[c + j] - mean[c + j] * sc[j];
    }
  };
  if (FIXED) load_consts((int)((reverse ? last - i0 : i0) % cg) * V);   // nvec is a multiple of cg: the group stays fixed either way
  for (int64_t i = i0; i < nvec; i += stride * kUnroll) {
    Vec16<T> vx[kUnroll], vr[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t kk = i + u * stride;
      if (kk < nvec) {
        const int64_t k = reverse ? last - kk : kk;
        vx[u].load(x + k * V);
        if (res != nullptr) vr[u].load(res + k * V);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t kk = i + u * stride;
      if (kk >= nvec) break;
      const int64_t k = reverse ? last - kk : kk;
      if (!FIXED) load_consts((int)(k % cg) * V);
      float xv[V], rv[V], out[V];
      vx[u].unpack(xv);
      if (res != nullptr) vr[u].unpack(rv);
      uint32_t bits = 0;
#pragma unroll
      for (int j = 0; j < V; ++j) {
        float v = fmaf(xv[j], sc[j], sh[j]);
        if (res != nullptr) v += rv[j];
        if (relu) v = fmaxf(v, 0.f);
        out[j] = v;
      }
      Vec16<T> vo; vo.pack(out); vo.store(y + k * V);
      if (mask != nullptr) {  // the sign of the ROUNDED output, i.e. exactly what `y > 0` reads back
        float ov[V];
        vo.unpack(ov);
#pragma unroll
        for (int j = 0; j < V; ++j) bits |= (ov[j] > 0.f ? 1u : 0u) << j;
        mask[k] = (uint8_t)bits;
      }
    }
  }
}

// U = vectors per thread per iteration.  The residual variant has 3 loads per vector in flight and streams at 0.97 of the HBM
// roof with U = 2; the variant without identity (2 loads, 1 store) sat at 0.77 with U = 2 (ncu: 16 warps/SM at 114 registers,
// 8 warps per issue slot waiting on loads) and gets U = 4.
template <typename T, bool FIXED, int U>
__global__ void __launch_bounds__(kThreads, 2)   // two CTAs per SM: 128 registers at most (the U = 4 variant sits right at that line)
bn_bwd_apply_kernel(const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx,
                    T* __restrict__ dres, const float* __restrict__ gamma, const float* __restrict__ beta,
                    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ dgamma,
                    const float* __restrict__ dbeta, int64_t nvec, int C, int relu, float inv_m,
                    const uint8_t* __restrict__ mask, int reverse) {
  constexpr int V = Vec16<T>::N;
  const int cg = C / V;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t last = nvec - 1;   // reverse: walk from the end (see bn_apply_kernel)
  // dx = k1*dz - k2*(x - mean) - k3   with k1 = gamma*invstd, k2 = k1*invstd*dgamma/M, k3 = k1*dbeta/M
  float k1[V], k2[V], k3[V], mu[V], msc[V], msh[V];
  auto load_consts = [&](int c) {
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const float is = invstd[c + j];
      k1[j] = gamma[c + j] * is;
      k2[j] = k1[j] * is * dgamma[c + j] * inv_m;
      k3[j] = k1[j] * dbeta[c + j] * inv_m;
      mu[j] = mean[c + j];
      if (relu == 2) {  // same expressions as bn_apply_kernel, so the recomputed sign is the forward's
        msc[j] = is * gamma[c + j];
        msh[j] = beta[c + j] - mu[j] * msc[j];
      }
    }
  };
  if (FIXED) load_consts((int)((reverse ? last - i0 : i0) % cg) * V);
  for (int64_t i = i0; i < nvec; i += stride * U) {
    Vec16<T> vx[U], vg[U], vy[U];
    uint32_t mb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t kk = i + u * stride;
      if (kk < nvec) {
        const int64_t k = reverse ? last - kk : kk;
        vx[u].load(x + k * V);
        vg[u].load(dy + k * V);
        if (relu == 1) vy[u].load(y + k * V);
        if (relu == 3) mb[u] = mask[k];
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t kk = i + u * stride;
      if (kk >= nvec) break;
      const int64_t k = reverse ? last - kk : kk;
      if (!FIXED) load_consts((int)(k % cg) * V);
      float xv[V], gv[V], out[V];
      vx[u].unpack(xv);
      vg[u].unpack(gv);
      if (relu == 1) {
        float yv[V];
        vy[u].unpack(yv);
#pragma unroll
        for (int j = 0; j < V; ++j) gv[j] = yv[j] > 0.f ? gv[j] : 0.f;
      } else if (relu == 2) {
#pragma unroll
        for (int j = 0; j < V; ++j) gv[j] = fmaf(xv[j], msc[j], msh[j]) > 0.f ? gv[j] : 0.f;
      } else if (relu == 3) {
#pragma unroll
        for (int j = 0; j < V; ++j) gv[j] = ((mb[u] >> j) & 1u) ? gv[j] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < V; ++j) out[j] = fmaf(k1[j], gv[j], -fmaf(k2[j], xv[j] - mu[j], k3[j]));
      Vec16<T> vo; vo.pack(out); vo.store(dx + k * V);
      if (dres != nullptr) { Vec16<T> vr; vr.pack(gv); vr.store(dres + k * V); }
    }
  }
}

template <typename T>
__global__ void relu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t nvec) {
  constexpr int V = Vec16<T>::N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    float v[V];
    Vec16<T> a; a.load(x + i * V); a.unpack(v);
#pragma unroll
    for (int j = 0; j < V; ++j) v[j] = fmaxf(v[j], 0.f);
    a.pack(v); a.store(y + i * V);
  }
}
template <typename T>
__global__ void relu_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, int64_t nvec) {
  constexpr int V = Vec16<T>::N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    float v[V], g[V];
    Vec16<T> a; a.load(y + i * V); a.unpack(v);
    Vec16<T> b; b.load(dy + i * V); b.unpack(g);
#pragma unroll
    for (int j = 0; j < V; ++j) g[j] = v[j] > 0.f ? g[j] : 0.f;
    b.pack(g); b.store(dx + i * V);
  }
}

// grid of a streaming kernel: `per_sm` CTAs per SM (what the kernel's register count lets be resident: one wave, no tail of
// late-starting CTAs), fewer when the tensor is small
static unsigned stream_grid(int64_t nvec, int per_sm) {
  int64_t b = ceil_div64(nvec, kThreads);
  const int64_t cap = (int64_t)num_sms() * per_sm;
  return (unsigned)(b < cap ? (b < 1 ? 1 : b) : cap);
}

template <typename T>
static int stats_t(const dv_bn_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  DV_REQUIRE(p->C % V == 0, DV_ERR_UNSUPPORTED, "bn: C=%d must be a multiple of %d", p->C, V);
  Plan pl = make_plan(p->M, p->C, V);
  cap_plan(pl, p->M, resident_ctas(bn_reduce_kernel<T, 0, 0>) * num_sms());
  const size_t need = (size_t)pl.nblk * 2 * p->C * sizeof(float);
  DV_REQUIRE(p->workspace && p->workspace_bytes >= need, DV_ERR_WORKSPACE, "bn_stats: workspace %zu bytes required, %zu given", need,
             p->workspace_bytes);
  bn_reduce_kernel<T, 0, 0><<<pl.nblk, kThreads, 0, st>>>((const T*)p->x, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                                                         (float*)p->workspace, p->M, p->C, pl.cgb, pl.rows_par, pl.rows_per_blk);
  DV_LAUNCH_OK("bn_reduce_kernel<stats>");
  bn_stats_finalize_kernel<T><<<ceil_div(p->C, 32), 32 * kFinLanes, 0, st>>>((const T*)p->x, (const float*)p->workspace, pl.nblk, p->C, p->M, p->eps,
                                                                  p->momentum, p->save_mean, p->save_invstd, p->running_mean,
                                                                  p->running_var);
  DV_LAUNCH_OK("bn_stats_finalize_kernel");
  return DV_OK;
}

template <typename T>
static int apply_t(const dv_bn_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  DV_REQUIRE(p->C % V == 0, DV_ERR_UNSUPPORTED, "bn: C=%d must be a multiple of %d", p->C, V);
  const int64_t nvec = p->M * p->C / V;
  const float* mean = p->eval_mode ? p->running_mean : p->save_mean;
  const float* is = p->eval_mode ? p->running_var : p->save_invstd;
  DV_REQUIRE(mean && is, DV_ERR_ARG, "bn_apply: statistics pointers are null");
  static const int apply_forced = env_or("DV_BN_APPLY_CTAS", 0);
  DV_REQUIRE(p->relu_mask == nullptr || (V == 8 && p->relu), DV_ERR_UNSUPPORTED, "bn_apply: relu_mask needs bf16 and relu == 1");
  uint8_t* mask = (uint8_t*)p->relu_mask;
  static const int rev_env = env_or("DV_BN_APPLY_REV", 0);   // experiment knob: 1 = always walk downwards, 2 = never
  const int reverse = rev_env == 1 ? 1 : (rev_env == 2 ? 0 : (p->reverse ? 1 : 0));
  if (kThreads % (p->C / V) == 0) {
    const unsigned grid = stream_grid(ceil_div64(nvec, kUnroll), apply_forced ? apply_forced : resident_ctas(bn_apply_kernel<T, true>));
    bn_apply_kernel<T, true><<<grid, kThreads, 0, st>>>((const T*)p->x, (const T*)p->residual, (T*)p->y, p->gamma, p->beta, mean, is, nvec,
                                                       p->C, p->relu, p->eval_mode, p->eps, mask, reverse);
  } else {
    const unsigned grid = stream_grid(ceil_div64(nvec, kUnroll), apply_forced ? apply_forced : resident_ctas(bn_apply_kernel<T, false>));
    bn_apply_kernel<T, false><<<grid, kThreads, 0, st>>>((const T*)p->x, (const T*)p->residual, (T*)p->y, p->gamma, p->beta, mean, is, nvec,
                                                        p->C, p->relu, p->eval_mode, p->eps, mask, reverse);
  }
  DV_LAUNCH_OK("bn_apply_kernel");
  return DV_OK;
}

template <typename T>
static int bwd_reduce_t(const dv_bn_bwd_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  DV_REQUIRE(p->C % V == 0, DV_ERR_UNSUPPORTED, "bn: C=%d must be a multiple of %d", p->C, V);
  DV_REQUIRE(!p->relu || p->y || p->beta || p->relu_mask, DV_ERR_ARG, "bn_bwd_reduce: relu needs relu_mask, y (stored mask) or beta (recomputed mask)");
  DV_REQUIRE(p->relu_mask == nullptr || V == 8, DV_ERR_UNSUPPORTED, "bn_bwd_reduce: relu_mask needs bf16");
  const int relu_mode = !p->relu ? 0 : (p->relu_mask ? 3 : (p->y ? 1 : 2));
  Plan pl = make_plan(p->M, p->C, V);
  int resident = 2;
  switch (relu_mode) {
    case 0: resident = resident_ctas(bn_reduce_kernel<T, 1, 0>); break;
    case 1: resident = resident_ctas(bn_reduce_kernel<T, 1, 1>); break;
    case 2: resident = resident_ctas(bn_reduce_kernel<T, 1, 2>); break;
    default: resident = resident_ctas(bn_reduce_kernel<T, 1, 3>); break;
  }
  cap_plan(pl, p->M, resident * num_sms());
  const size_t need = p->atomic_sums ? 0 : (size_t)pl.nblk * 2 * p->C * sizeof(float);
  DV_REQUIRE(need == 0 || (p->workspace && p->workspace_bytes >= need), DV_ERR_WORKSPACE,
             "bn_bwd_reduce: workspace %zu bytes required, %zu given", need, p->workspace_bytes);
#define DV_BWD_REDUCE(R_)                                                                                                          \
  bn_reduce_kernel<T, 1, R_><<<pl.nblk, kThreads, 0, st>>>((const T*)p->x, (const T*)p->y, (const T*)p->dy, p->save_mean, p->save_invstd, \
                                                          p->gamma, p->beta, (float*)p->workspace, p->M, p->C, pl.cgb, pl.rows_par,     \
                                                          pl.rows_per_blk, p->atomic_sums ? p->dbeta : nullptr,                          \
                                                          p->atomic_sums ? p->dgamma : nullptr, (const uint8_t*)p->relu_mask)
  switch (relu_mode) {
    case 0: DV_BWD_REDUCE(0); break;
    case 1: DV_BWD_REDUCE(1); break;
    case 2: DV_BWD_REDUCE(2); break;
    default: DV_BWD_REDUCE(3); break;
  }
#undef DV_BWD_REDUCE
  DV_LAUNCH_OK("bn_reduce_kernel<bwd>");
  if (p->atomic_sums) return DV_OK;  // dbeta += sum(dz), dgamma += sum(dz*xhat) were accumulated in place
  bn_bwd_finalize_kernel<<<ceil_div(p->C, 32), 32 * kFinLanes, 0, st>>>((const float*)p->workspace, pl.nblk, p->C, p->dgamma, p->dbeta);
  DV_LAUNCH_OK("bn_bwd_finalize_kernel");
  return DV_OK;
}

template <typename T>
static int bwd_apply_t(const dv_bn_bwd_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  DV_REQUIRE(p->C % V == 0, DV_ERR_UNSUPPORTED, "bn: C=%d must be a multiple of %d", p->C, V);
  DV_REQUIRE(!p->relu || p->y || p->beta || p->relu_mask, DV_ERR_ARG, "bn_bwd_apply: relu needs relu_mask, y (stored mask) or beta (recomputed mask)");
  DV_REQUIRE(p->relu_mask == nullptr || V == 8, DV_ERR_UNSUPPORTED, "bn_bwd_apply: relu_mask needs bf16");
  const int relu_mode = !p->relu ? 0 : (p->relu_mask ? 3 : (p->y ? 1 : 2));
  const int64_t nvec = p->M * p->C / V;
  const bool wide = (relu_mode != 1 && p->dres == nullptr);   // two loads per vector only: keep four vectors in flight
  static const int bwd_forced = env_or("DV_BN_BWD_CTAS", 0);
  static const int rev_env = env_or("DV_BN_BWD_REV", 0);   // experiment knob: 1 = always walk downwards, 2 = never
  const int reverse = rev_env == 1 ? 1 : (rev_env == 2 ? 0 : (p->reverse ? 1 : 0));
#define DV_BWD_APPLY(FIXED_, U_)                                                                                                     \
  do {                                                                                                                               \
    const unsigned grid = stream_grid(ceil_div64(nvec, U_), bwd_forced ? bwd_forced : resident_ctas(bn_bwd_apply_kernel<T, FIXED_, U_>)); \
    bn_bwd_apply_kernel<T, FIXED_, U_><<<grid, kThreads, 0, st>>>((const T*)p->x, (const T*)p->y, (const T*)p->dy, (T*)p->dx, (T*)p->dres, \
                                                                 p->gamma, p->beta, p->save_mean, p->save_invstd, p->dgamma, p->dbeta, nvec, \
                                                                 p->C, relu_mode, 1.0f / (float)p->M, (const uint8_t*)p->relu_mask, reverse);  \
  } while (0)
  if (kThreads % (p->C / V) == 0) {
    if (wide) DV_BWD_APPLY(true, 4); else DV_BWD_APPLY(true, 2);
  } else {
    if (wide) DV_BWD_APPLY(false, 4); else DV_BWD_APPLY(false, 2);
  }
#undef DV_BWD_APPLY
  DV_LAUNCH_OK("bn_bwd_apply_kernel");
  return DV_OK;
}

// ---- BatchNorm + ReLU + MaxPool(3, 2, 1) in one pass each way (the ResNet stem: resnet50.py:30-34,49-52) -------------------------
// conv1's output (411 MB at batch 256) is the largest activation of the network.  Unfused it is normalised into a second
// 411 MB tensor that only the max-pool reads, and on the way back the pool's 411 MB gradient is written, then read by both
// BatchNorm backward passes.  Fused: the forward reads x once and writes the pooled activation + the window index (154 MB);
// the two backward passes gather the pooled gradient through the index on the fly (103 + 51 MB instead of 411 MB).  Same
// arithmetic, same bits: a = bf16(relu(fma(x, sc, sh))) is formed per input element before the comparison, ties keep the
// first window position, the ReLU mask is recomputed from x as everywhere else.
template <typename T>
__global__ void __launch_bounds__(256)
bn_relu_maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ idx, const float* __restrict__ gamma,
                           const float* __restrict__ beta, const float* __restrict__ mean, const float* __restrict__ invstd, int N, int H,
                           int W, int C, int P, int Q) {
  constexpr int V = Vec16<T>::N;
  const int cg = C / V;
  const int Q2 = (Q + 1) >> 1;
  const int total = N * P * Q2 * cg;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = (i % cg) * V;
    int pix = i / cg;
    const int q0 = (pix % Q2) * 2; pix /= Q2;
    const int p = pix % P;
    const int n = pix / P;
    float sc[V], sh[V];
#pragma unroll
    for (int j = 0; j < V; ++j) {   // the expressions of bn_apply_kernel
      sc[j] = invstd[c + j] * gamma[c + j];
      sh[j] = beta[c + j] - mean[c + j] * sc[j];
    }
    Vec16<T> in[3][5];
    bool ok[3][5];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = p * 2 + r - 1;
#pragma unroll
      for (int s2 = 0; s2 < 5; ++s2) {
        const int w = q0 * 2 + s2 - 1;
        ok[r][s2] = (unsigned)h < (unsigned)H && (unsigned)w < (unsigned)W;
        if (ok[r][s2]) in[r][s2].load(x + (((size_t)n * H + h) * W + w) * C + c);
      }
    }
    // normalise + ReLU + round each loaded vector once (what dv_bn_apply would have written)
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int s2 = 0; s2 < 5; ++s2) {
        if (!ok[r][s2]) continue;
        float v[V];
        in[r][s2].unpack(v);
#pragma unroll
        for (int j = 0; j < V; ++j) v[j] = fmaxf(fmaf(v[j], sc[j], sh[j]), 0.f);
        in[r][s2].pack(v);
      }
#pragma unroll
    for (int o = 0; o < 2; ++o) {
      const int q = q0 + o;
      if (q >= Q) break;
      float best[V];
      __align__(8) uint8_t bi[V];
#pragma unroll
      for (int j = 0; j < V; ++j) { best[j] = 0.f; bi[j] = 0; }
      bool first = true;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int s2 = 0; s2 < 3; ++s2) {
          if (!ok[r][2 * o + s2]) continue;
          float v[V];
          in[r][2 * o + s2].unpack(v);
#pragma unroll
          for (int j = 0; j < V; ++j) {
            if (first || v[j] > best[j]) { best[j] = v[j]; bi[j] = (uint8_t)(r * 3 + s2); }
          }
          first = false;
        }
      }
      const size_t off = (((size_t)n * P + p) * Q + q) * C + c;
      Vec16<T> out; out.pack(best); out.store(y + off);
      if (V == 8) *reinterpret_cast<uint2*>(idx + off) = *reinterpret_cast<const uint2*>(bi);
      else *reinterpret_cast<uint32_t*>(idx + off) = *reinterpret_cast<const uint32_t*>(bi);
    }
  }
}

// MODE 0: block partials of sum(dz), sum(dz * xhat);  MODE 1: dx = k1*dz - k2*(x - mean) - k3.
// One thread = one 16-byte channel vector of a 2x2 block of x; dz = [gradient gathered from the <= 4 pooling windows whose
// maximum sits on this element] * [bn(x) > 0].  The grid stride is a multiple of C/V (the host checks 256 % (C/V) == 0), so a
// thread keeps its channels.
template <typename T, int MODE>
__global__ void __launch_bounds__(256)
bn_relu_maxpool_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, const uint8_t* __restrict__ idx, T* __restrict__ dx,
                           const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                           const float* __restrict__ invstd, const float* __restrict__ dgamma, const float* __restrict__ dbeta,
                           float* __restrict__ partial, float inv_m, int N, int H, int W, int C, int P, int Q) {
  constexpr int V = Vec16<T>::N;
  __shared__ float red[2][256 * V];
  const int cg = C / V;
  const int H2 = (H + 1) >> 1, W2 = (W + 1) >> 1;
  const int total = N * H2 * W2 * cg;
  const int i0 = blockIdx.x * blockDim.x + threadIdx.x;
  const int c = (i0 % cg) * V;
  float sc[V], sh[V], mu[V], is[V], k1[V], k2[V], k3[V], a1[V], a2[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    is[j] = invstd[c + j];
    mu[j] = mean[c + j];
    sc[j] = is[j] * gamma[c + j];
    sh[j] = beta[c + j] - mu[j] * sc[j];
    a1[j] = 0.f; a2[j] = 0.f;
    if (MODE == 1) {
      k1[j] = gamma[c + j] * is[j];
      k2[j] = k1[j] * is[j] * dgamma[c + j] * inv_m;
      k3[j] = k1[j] * dbeta[c + j] * inv_m;
    }
  }
  for (int i = i0; i < total; i += gridDim.x * blockDim.x) {
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
        const bool okw = (p + a) < P && (q + b) < Q;
        if (okw) {
          const size_t o = (((size_t)n * P + p + a) * Q + q + b) * C + c;
          Vec16<T> v; v.load(dy + o); v.unpack(g[a][b]);
          if (V == 8) *reinterpret_cast<uint2*>(ti[a][b]) = *reinterpret_cast<const uint2*>(idx + o);
          else *reinterpret_cast<uint32_t*>(ti[a][b]) = *reinterpret_cast<const uint32_t*>(idx + o);
        } else {
#pragma unroll
          for (int j = 0; j < V; ++j) { g[a][b][j] = 0.f; ti[a][b][j] = 255; }
        }
      }
    Vec16<T> xin[2][2];
#pragma unroll
    for (int dh = 0; dh < 2; ++dh)
#pragma unroll
      for (int dw = 0; dw < 2; ++dw)
        if (2 * p + dh < H && 2 * q + dw < W) xin[dh][dw].load(x + (((size_t)n * H + 2 * p + dh) * W + 2 * q + dw) * C + c);
#pragma unroll
    for (int dh = 0; dh < 2; ++dh) {
      const int h = 2 * p + dh;
      if (h >= H) break;
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        const int w = 2 * q + dw;
        if (w >= W) break;
        float acc[V], xv[V];
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
        // the pool's gradient is a bf16 tensor in the unfused pipeline: round it the same way before BatchNorm's arithmetic
        { Vec16<T> t; t.pack(acc); t.unpack(acc); }
        xin[dh][dw].unpack(xv);
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] = fmaf(xv[j], sc[j], sh[j]) > 0.f ? acc[j] : 0.f;
        if (MODE == 0) {
#pragma unroll
          for (int j = 0; j < V; ++j) { a1[j] += acc[j]; a2[j] = fmaf(acc[j], (xv[j] - mu[j]) * is[j], a2[j]); }
        } else {
          float out[V];
#pragma unroll
          for (int j = 0; j < V; ++j) out[j] = fmaf(k1[j], acc[j], -fmaf(k2[j], xv[j] - mu[j], k3[j]));
          Vec16<T> o; o.pack(out); o.store(dx + (((size_t)n * H + h) * W + w) * C + c);
        }
      }
    }
  }
  if (MODE == 0) {
    const int tcg = threadIdx.x % cg, lanes = 256 / cg;
#pragma unroll
    for (int j = 0; j < V; ++j) { red[0][threadIdx.x * V + j] = a1[j]; red[1][threadIdx.x * V + j] = a2[j]; }
    __syncthreads();
    if (threadIdx.x < cg) {
      float s1[V], s2[V];
#pragma unroll
      for (int j = 0; j < V; ++j) { s1[j] = 0.f; s2[j] = 0.f; }
      for (int l = 0; l < lanes; ++l)
#pragma unroll
        for (int j = 0; j < V; ++j) { s1[j] += red[0][(l * cg + tcg) * V + j]; s2[j] += red[1][(l * cg + tcg) * V + j]; }
      float* p1 = partial + ((size_t)blockIdx.x * 2) * C + tcg * V;
#pragma unroll
      for (int j = 0; j < V; ++j) { p1[j] = s1[j]; p1[C + j] = s2[j]; }
    }
  }
}

// ---- bf16 versions, written against the instruction-issue limit ---------------------------------------------------------------
// The generic kernels above spend ~80 instructions per pooled element and run at a third of the HBM roof.  Here everything after
// the fp32 fma stays in packed bf16x2 registers: ReLU, the window maximum and the comparison are 16-bit SIMD integer / half ops on
// the (non-negative, canonical) bit patterns, and the window position is tracked with bitwise selects on 16-bit masks.
// Same values bit for bit: bf16(max(v, 0)) == max(bf16(v), +0) (rounding is monotone, and as int16 a negative bf16 or -0 is < 0).
__device__ __forceinline__ float2 bfx2_to_f2(uint32_t u) { return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u)); }
__device__ __forceinline__ uint32_t f2_to_bfx2(float2 v) {
  __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint32_t bfx2_eq_mask(uint32_t a, uint32_t b) {   // 0xFFFF per equal half
  return __heq2_mask(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
}
__device__ __forceinline__ uint32_t bfx2_gt_mask(uint32_t a, uint32_t b) {   // 0xFFFF per half where a > b
  return __hgt2_mask(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
}
constexpr uint32_t kNegInf2 = 0xFF80FF80u;   // bf16x2 (-inf, -inf): a window position outside the image

// Division by a run-time constant as multiply-high + shift (Granlund-Montgomery), exact for dividends below 2^31.
struct FastDiv {
  uint32_t mul, shift, d;
  __host__ FastDiv() : mul(1), shift(0), d(1) {}
  __host__ explicit FastDiv(uint32_t div) : d(div) {
    shift = 0;
    while ((1u << shift) < div) ++shift;
    mul = (uint32_t)((((uint64_t)1 << 32) * ((1ull << shift) - div)) / div + 1);
  }
  __device__ __forceinline__ uint32_t div(uint32_t n) const { return (__umulhi(mul, n) + n) >> shift; }
};

// One thread = a 2x2 block of pooled outputs x 8 channels: 5x5 input vectors, each normalised once, consumed row by row.
// ReLU is applied to the window maximum, not to the 25 inputs: max(relu(y_t)) == relu(max(y_t)), and the position of the
// maximum only differs when nothing in the window is positive — then every relu(y_t) is 0 and the first position inside the
// image wins, which is known without looking at the data.  (As int16 lanes negative bf16 values order backwards; that only
// reorders values that lose to any positive one, and an all-negative window takes the data-free branch.)
template <bool EDGE>
__device__ __forceinline__ void bn_pool_fwd_tile(const bf16* __restrict__ x, bf16* __restrict__ y, uint8_t* __restrict__ idx,
                                                 const float2 (&sc)[4], const float2 (&sh)[4], int n, int p0, int q0, int c, int H,
                                                 int W, int C, int P, int Q) {
  const int h0 = 2 * p0 - 1, w0 = 2 * q0 - 1;
  bool okw[5];
#pragma unroll
  for (int s = 0; s < 5; ++s) okw[s] = !EDGE || (unsigned)(w0 + s) < (unsigned)W;
  uint32_t best[2][2][4], bidx[2][2][4];   // running maximum and its window position (16 bits per channel)
  uint4 raw[2][5];
  const bf16* base = x + (((size_t)n * H + h0) * W + w0) * C + c;
  const size_t row_pitch = (size_t)W * C;
  auto load_row = [&](int r, uint4 (&dst)[5]) {
    const bool okh = !EDGE || (unsigned)(h0 + r) < (unsigned)H;
#pragma unroll
    for (int s = 0; s < 5; ++s)
      if (okh && okw[s]) dst[s] = *reinterpret_cast<const uint4*>(base + r * row_pitch + (size_t)s * C);
  };
  load_row(0, raw[0]);
#pragma unroll
  for (int r = 0; r < 5; ++r) {
    if (r < 4) load_row(r + 1, raw[(r + 1) & 1]);
    const bool okh = !EDGE || (unsigned)(h0 + r) < (unsigned)H;
    uint32_t yv[5][4];
#pragma unroll
    for (int s = 0; s < 5; ++s) {
      const uint32_t w4[4] = {raw[r & 1][s].x, raw[r & 1][s].y, raw[r & 1][s].z, raw[r & 1][s].w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t b = f2_to_bfx2(__ffma2_rn(bfx2_to_f2(w4[k]), sc[k], sh[k]));
        yv[s][k] = (!EDGE || (okh && okw[s])) ? b : kNegInf2;
      }
    }
    // this input row is window row `tr` of pooled row `orow`: r = 2*orow + tr
#pragma unroll
    for (int orow = 0; orow < 2; ++orow) {
      const int tr = r - 2 * orow;
      if (tr < 0 || tr > 2) continue;
#pragma unroll
      for (int oc = 0; oc < 2; ++oc) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint32_t a0 = yv[2 * oc][k], a1 = yv[2 * oc + 1][k], a2 = yv[2 * oc + 2][k];
          const uint32_t m = __vimax3_s16x2(a0, a1, a2);
          // first position of the row's maximum (scan from the right so the leftmost wins)
          uint32_t t = (uint32_t)(tr * 3 + 2) * 0x00010001u;
          const uint32_t e1 = bfx2_eq_mask(a1, m), e0 = bfx2_eq_mask(a0, m);
          t = (t & ~e1) | (((uint32_t)(tr * 3 + 1) * 0x00010001u) & e1);
          t = (t & ~e0) | (((uint32_t)(tr * 3 + 0) * 0x00010001u) & e0);
          if (tr == 0) {
            best[orow][oc][k] = m;
            bidx[orow][oc][k] = t;
          } else {   // a later row wins only when strictly greater
            const uint32_t g = bfx2_gt_mask(m, best[orow][oc][k]);
            bidx[orow][oc][k] = (bidx[orow][oc][k] & ~g) | (t & g);
            best[orow][oc][k] = __vmaxs2(best[orow][oc][k], m);
          }
        }
      }
    }
  }
#pragma unroll
  for (int orow = 0; orow < 2; ++orow) {
    if (EDGE && p0 + orow >= P) break;
#pragma unroll
    for (int oc = 0; oc < 2; ++oc) {
      if (EDGE && q0 + oc >= Q) break;
      // first window position inside the image: row 0 / column 0 of the window are padding for the first pooled row / column
      const uint32_t first = (uint32_t)((p0 + orow == 0 ? 3 : 0) + (q0 + oc == 0 ? 1 : 0)) * 0x00010001u;
      uint32_t o[4], t[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t pos = bfx2_gt_mask(best[orow][oc][k], 0u);
        o[k] = __vimax_s16x2_relu(best[orow][oc][k], best[orow][oc][k]);
        t[k] = (first & ~pos) | (bidx[orow][oc][k] & pos);
      }
      const size_t off = (((size_t)n * P + p0 + orow) * Q + q0 + oc) * C + c;
      *reinterpret_cast<uint4*>(y + off) = make_uint4(o[0], o[1], o[2], o[3]);
      *reinterpret_cast<uint2*>(idx + off) = make_uint2(__byte_perm(t[0], t[1], 0x6420), __byte_perm(t[2], t[3], 0x6420));
    }
  }
}

__global__ void __launch_bounds__(256, 2)
bn_relu_maxpool_fwd_bf16_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, uint8_t* __restrict__ idx, const float* __restrict__ gamma,
                                const float* __restrict__ beta, const float* __restrict__ mean, const float* __restrict__ invstd, int N,
                                int H, int W, int C, int P, int Q, int cg_log2, FastDiv div_q2, FastDiv div_p2) {
  const int cg = 1 << cg_log2;
  const int P2 = (P + 1) >> 1, Q2 = (Q + 1) >> 1;
  const int total = (N * P2 * Q2) << cg_log2;
  const int i0 = blockIdx.x * blockDim.x + threadIdx.x;
  const int c = (i0 & (cg - 1)) * 8;    // the grid stride is a multiple of cg: a thread keeps its channels
  float2 sc[4], sh[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {   // the expressions of bn_apply_kernel
    sc[k].x = invstd[c + 2 * k] * gamma[c + 2 * k];
    sc[k].y = invstd[c + 2 * k + 1] * gamma[c + 2 * k + 1];
    sh[k].x = beta[c + 2 * k] - mean[c + 2 * k] * sc[k].x;
    sh[k].y = beta[c + 2 * k + 1] - mean[c + 2 * k + 1] * sc[k].y;
  }
  for (int i = i0; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t pix = (uint32_t)i >> cg_log2;
    const uint32_t t1 = div_q2.div(pix);
    const int q0 = (int)(pix - t1 * (uint32_t)Q2) * 2;
    const uint32_t n = div_p2.div(t1);
    const int p0 = (int)(t1 - n * (uint32_t)P2) * 2;
    // interior tiles need no bounds checks: rows 2*p0-1 .. 2*p0+3 and columns 2*q0-1 .. 2*q0+3 inside the image
    const bool inside = p0 > 0 && q0 > 0 && 2 * p0 + 3 < H && 2 * q0 + 3 < W;
    if (inside) bn_pool_fwd_tile<false>(x, y, idx, sc, sh, (int)n, p0, q0, c, H, W, C, P, Q);
    else bn_pool_fwd_tile<true>(x, y, idx, sc, sh, (int)n, p0, q0, c, H, W, C, P, Q);
  }
}

// 16-bit lane masks that are 0xFFFF where the stored window position DIFFERS from `tap` (4 channels per index word).
// Positions are 0..8 (255 = no window), so (i ^ tap) < 0x80 except for "no window", and per byte ((i ^ tap) | 0x80) - 1 clears the
// top bit exactly where the two are equal; PRMT's sign-replicate mode spreads that bit over a 16-bit lane.
__device__ __forceinline__ void pool_ne_masks(uint32_t iw, uint32_t tap4, uint32_t& m01, uint32_t& m23) {
  const uint32_t e = ((iw ^ tap4) | 0x80808080u) - 0x01010101u;
  // prmt with bit 3 of a selector nibble set replicates the selected byte's sign (__byte_perm() masks that bit off)
  asm("prmt.b32 %0, %1, %1, 0x9988;" : "=r"(m01) : "r"(e));
  asm("prmt.b32 %0, %1, %1, 0xBBAA;" : "=r"(m23) : "r"(e));
}
__device__ __forceinline__ uint32_t bfx2_add(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hadd2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

// One thread = a 2x2 block of x (rows 2p, 2p+1; columns 2q, 2q+1) x 8 channels and the four windows (p+a, q+b) that can have
// their maximum in it.  Element (dh, dw) is window position (dh+1-2a, dw+1-2b) of window (a, b), a <= dh, b <= dw: one, two, two
// and four candidate windows.  One or two candidates are combined in bf16 (a single rounding, as the fp32 sum + rounding of
// maxpool_bwd_kernel gives: two bf16 values add exactly in fp32 unless they are > 2^16 apart, where both roundings return the
// larger one); four go through fp32 in maxpool_bwd_kernel's order.  dy is canonicalised (-0 -> +0) on load because that kernel's
// accumulator starts from +0.
template <int MODE>
__global__ void __launch_bounds__(256, 2)
bn_relu_maxpool_bwd_bf16_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, const uint8_t* __restrict__ idx, bf16* __restrict__ dx,
                                const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                                const float* __restrict__ invstd, const float* __restrict__ dgamma, const float* __restrict__ dbeta,
                                float* __restrict__ partial, float inv_m, int N, int H, int W, int C, int P, int Q, int cg_log2,
                                FastDiv div_w2, FastDiv div_h2) {
  __shared__ float red[MODE == 0 ? 2 * 256 * 8 : 1];
  const int cg = 1 << cg_log2;
  const int H2 = (H + 1) >> 1, W2 = (W + 1) >> 1;
  const int total = (N * H2 * W2) << cg_log2;
  const int i0 = blockIdx.x * blockDim.x + threadIdx.x;
  const int c = (i0 & (cg - 1)) * 8;
  float2 sc[4], sh[4], nmu[4], k2[4], k3[4], a1[4], a2[4];   // MODE 0: k2 holds invstd; sc doubles as k1 (= gamma * invstd)
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float is_[2], mu_[2], sc_[2], sh_[2], k2_[2], k3_[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int ch = c + 2 * k + j;
      is_[j] = invstd[ch];
      mu_[j] = mean[ch];
      sc_[j] = is_[j] * gamma[ch];
      sh_[j] = beta[ch] - mu_[j] * sc_[j];
      if (MODE == 1) {
        const float k1 = gamma[ch] * is_[j];
        k2_[j] = k1 * is_[j] * dgamma[ch] * inv_m;
        k3_[j] = k1 * dbeta[ch] * inv_m;
      } else {
        k2_[j] = is_[j];
        k3_[j] = 0.f;
      }
    }
    sc[k] = make_float2(sc_[0], sc_[1]); sh[k] = make_float2(sh_[0], sh_[1]);
    nmu[k] = make_float2(-mu_[0], -mu_[1]);
    k2[k] = make_float2(k2_[0], k2_[1]); k3[k] = make_float2(k3_[0], k3_[1]);
    a1[k] = make_float2(0.f, 0.f); a2[k] = make_float2(0.f, 0.f);
  }
  for (int i = i0; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t pix = (uint32_t)i >> cg_log2;
    const uint32_t t1 = div_w2.div(pix);
    const int q = (int)(pix - t1 * (uint32_t)W2);
    const int n = (int)div_h2.div(t1);
    const int p = (int)(t1 - (uint32_t)n * (uint32_t)H2);
    uint32_t g[2][2][4], ti[2][2][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        if ((p + a) < P && (q + b) < Q) {
          const size_t o = (((size_t)n * P + p + a) * Q + q + b) * C + c;
          const uint4 v = *reinterpret_cast<const uint4*>(dy + o);
          const uint2 t = *reinterpret_cast<const uint2*>(idx + o);
          g[a][b][0] = v.x; g[a][b][1] = v.y; g[a][b][2] = v.z; g[a][b][3] = v.w;
          ti[a][b][0] = t.x; ti[a][b][1] = t.y;
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) g[a][b][k] = 0u;
          ti[a][b][0] = ti[a][b][1] = 0xFFFFFFFFu;
        }
      }
    uint4 xin[2][2];
#pragma unroll
    for (int dh = 0; dh < 2; ++dh)
#pragma unroll
      for (int dw = 0; dw < 2; ++dw)
        if (2 * p + dh < H && 2 * q + dw < W)
          xin[dh][dw] = *reinterpret_cast<const uint4*>(x + (((size_t)n * H + 2 * p + dh) * W + 2 * q + dw) * C + c);
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int k = 0; k < 4; ++k) g[a][b][k] = bfx2_add(g[a][b][k], 0u);   // -0 -> +0
#pragma unroll
    for (int dh = 0; dh < 2; ++dh) {
      if (2 * p + dh >= H) break;
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        if (2 * q + dw >= W) break;
        // the pooled gradient that lands on this element, rounded to bf16 as the pool's backward would have stored it
        uint32_t sel[2][2][4];
#pragma unroll
        for (int a = 0; a <= dh; ++a)
#pragma unroll
          for (int b = 0; b <= dw; ++b) {
            const uint32_t tap4 = (uint32_t)((dh + 1 - 2 * a) * 3 + (dw + 1 - 2 * b)) * 0x01010101u;
#pragma unroll
            for (int hw = 0; hw < 2; ++hw) {
              uint32_t m01, m23;
              pool_ne_masks(ti[a][b][hw], tap4, m01, m23);
              sel[a][b][2 * hw] = g[a][b][2 * hw] & ~m01;
              sel[a][b][2 * hw + 1] = g[a][b][2 * hw + 1] & ~m23;
            }
          }
        const uint32_t xw[4] = {xin[dh][dw].x, xin[dh][dw].y, xin[dh][dw].z, xin[dh][dw].w};
        uint32_t outw[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          uint32_t dzb;
          if (dh == 0 && dw == 0) dzb = sel[0][0][k];
          else if (dh == 0) dzb = bfx2_add(sel[0][0][k], sel[0][1][k]);
          else if (dw == 0) dzb = bfx2_add(sel[0][0][k], sel[1][0][k]);
          else {
            float2 t = __fadd2_rn(bfx2_to_f2(sel[0][0][k]), bfx2_to_f2(sel[0][1][k]));
            t = __fadd2_rn(t, bfx2_to_f2(sel[1][0][k]));
            t = __fadd2_rn(t, bfx2_to_f2(sel[1][1][k]));
            dzb = f2_to_bfx2(t);
          }
          const float2 xf = bfx2_to_f2(xw[k]);
          const float2 v = __ffma2_rn(xf, sc[k], sh[k]);
          float2 dz = bfx2_to_f2(dzb);
          dz.x = v.x > 0.f ? dz.x : 0.f;
          dz.y = v.y > 0.f ? dz.y : 0.f;
          const float2 xc = __fadd2_rn(xf, nmu[k]);
          if (MODE == 0) {
            a1[k] = __fadd2_rn(a1[k], dz);
            a2[k] = __ffma2_rn(dz, __fmul2_rn(xc, k2[k]), a2[k]);
          } else {
            const float2 t = __ffma2_rn(k2[k], xc, k3[k]);
            outw[k] = f2_to_bfx2(__ffma2_rn(sc[k], dz, make_float2(-t.x, -t.y)));
          }
        }
        if (MODE == 1)
          *reinterpret_cast<uint4*>(dx + (((size_t)n * H + 2 * p + dh) * W + 2 * q + dw) * C + c) = make_uint4(outw[0], outw[1], outw[2], outw[3]);
      }
    }
  }
  if (MODE == 0) {
    const int tcg = threadIdx.x % cg, lanes = 256 / cg;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      red[threadIdx.x * 8 + 2 * k] = a1[k].x; red[threadIdx.x * 8 + 2 * k + 1] = a1[k].y;
      red[256 * 8 + threadIdx.x * 8 + 2 * k] = a2[k].x; red[256 * 8 + threadIdx.x * 8 + 2 * k + 1] = a2[k].y;
    }
    __syncthreads();
    if (threadIdx.x < cg) {
      float s1[8], s2[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) { s1[j] = 0.f; s2[j] = 0.f; }
      for (int l = 0; l < lanes; ++l)
#pragma unroll
        for (int j = 0; j < 8; ++j) { s1[j] += red[(l * cg + tcg) * 8 + j]; s2[j] += red[256 * 8 + (l * cg + tcg) * 8 + j]; }
      float* p1 = partial + ((size_t)blockIdx.x * 2) * C + tcg * 8;
#pragma unroll
      for (int j = 0; j < 8; ++j) { p1[j] = s1[j]; p1[C + j] = s2[j]; }
    }
  }
}

static int ilog2(int v) { int l = 0; while ((1 << l) < v) ++l; return l; }
static int bn_pool_blocks(int64_t cells_vec) {   // one wave of resident CTAs, a multiple-of-nothing grid (256 threads each)
  int64_t b = ceil_div64(cells_vec, 256);
  const int64_t cap = (int64_t)num_sms() * 4;
  return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

template <typename T>
static int bn_pool_fwd_t(const dv_bn_pool_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  const int P = (p->H - 1) / 2 + 1, Q = (p->W - 1) / 2 + 1;
  const int64_t total = (int64_t)p->N * P * ((Q + 1) / 2) * (p->C / V);
  if constexpr (std::is_same<T, bf16>::value) {
    const int64_t blocks2 = (int64_t)p->N * ((P + 1) / 2) * ((Q + 1) / 2) * (p->C / V);
    bn_relu_maxpool_fwd_bf16_kernel<<<bn_pool_blocks(blocks2), 256, 0, st>>>((const bf16*)p->x, (bf16*)p->y, p->idx, p->gamma, p->beta,
                                                                            p->save_mean, p->save_invstd, p->N, p->H, p->W, p->C, P, Q,
                                                                            ilog2(p->C / V), FastDiv((Q + 1) / 2), FastDiv((P + 1) / 2));
  } else {
    bn_relu_maxpool_fwd_kernel<T><<<bn_pool_blocks(total) * 2, 256, 0, st>>>((const T*)p->x, (T*)p->y, p->idx, p->gamma, p->beta, p->save_mean,
                                                                            p->save_invstd, p->N, p->H, p->W, p->C, P, Q);
  }
  DV_LAUNCH_OK("bn_relu_maxpool_fwd_kernel");
  return DV_OK;
}
template <typename T>
static int bn_pool_bwd_t(const dv_bn_pool_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  const int P = (p->H - 1) / 2 + 1, Q = (p->W - 1) / 2 + 1;
  const int64_t total = (int64_t)p->N * ((p->H + 1) / 2) * ((p->W + 1) / 2) * (p->C / V);
  const int nblk = bn_pool_blocks(total);
  const size_t need = (size_t)nblk * 2 * p->C * sizeof(float);
  DV_REQUIRE(p->workspace && p->workspace_bytes >= need, DV_ERR_WORKSPACE, "dv_bn_relu_maxpool_bwd: workspace %zu bytes required, %zu given",
             need, p->workspace_bytes);
  const float inv_m = 1.0f / (float)((int64_t)p->N * p->H * p->W);
  if constexpr (std::is_same<T, bf16>::value)
    bn_relu_maxpool_bwd_bf16_kernel<0><<<nblk, 256, 0, st>>>((const bf16*)p->x, (const bf16*)p->dy, p->idx, nullptr, p->gamma, p->beta,
                                                            p->save_mean, p->save_invstd, nullptr, nullptr, (float*)p->workspace, inv_m, p->N,
                                                            p->H, p->W, p->C, P, Q, ilog2(p->C / V), FastDiv((p->W + 1) / 2),
                                                            FastDiv((p->H + 1) / 2));
  else
    bn_relu_maxpool_bwd_kernel<T, 0><<<nblk, 256, 0, st>>>((const T*)p->x, (const T*)p->dy, p->idx, nullptr, p->gamma, p->beta, p->save_mean,
                                                          p->save_invstd, nullptr, nullptr, (float*)p->workspace, inv_m, p->N, p->H, p->W, p->C, P, Q);
  DV_LAUNCH_OK("bn_relu_maxpool_bwd_kernel<reduce>");
  bn_bwd_finalize_kernel<<<ceil_div(p->C, 32), 32 * kFinLanes, 0, st>>>((const float*)p->workspace, nblk, p->C, p->dgamma, p->dbeta);
  DV_LAUNCH_OK("bn_bwd_finalize_kernel");
  if constexpr (std::is_same<T, bf16>::value)
    bn_relu_maxpool_bwd_bf16_kernel<1><<<nblk, 256, 0, st>>>((const bf16*)p->x, (const bf16*)p->dy, p->idx, (bf16*)p->dx, p->gamma, p->beta,
                                                            p->save_mean, p->save_invstd, p->dgamma, p->dbeta, nullptr, inv_m, p->N, p->H,
                                                            p->W, p->C, P, Q, ilog2(p->C / V), FastDiv((p->W + 1) / 2), FastDiv((p->H + 1) / 2));
  else
    bn_relu_maxpool_bwd_kernel<T, 1><<<nblk, 256, 0, st>>>((const T*)p->x, (const T*)p->dy, p->idx, (T*)p->dx, p->gamma, p->beta, p->save_mean,
                                                          p->save_invstd, p->dgamma, p->dbeta, nullptr, inv_m, p->N, p->H, p->W, p->C, P, Q);
  DV_LAUNCH_OK("bn_relu_maxpool_bwd_kernel<apply>");
  return DV_OK;
}
static int bn_pool_check(const dv_bn_pool_params* p, const char* who) {
  DV_REQUIRE(p && p->x && p->idx && p->gamma && p->beta && p->save_mean && p->save_invstd && p->N > 0 && p->H > 0 && p->W > 0 && p->C > 0,
             DV_ERR_ARG, "%s: bad arguments", who);
  DV_REQUIRE(p->dtype == DV_F32 || p->dtype == DV_BF16, DV_ERR_ARG, "%s: bad dtype", who);
  const int V = p->dtype == DV_BF16 ? 8 : 4;
  DV_REQUIRE(p->C % V == 0 && 256 % (p->C / V) == 0, DV_ERR_UNSUPPORTED, "%s: C=%d must be %d times a power of two up to 256", who, p->C, V);
  DV_REQUIRE((int64_t)p->N * p->H * p->W * p->C < (1LL << 31), DV_ERR_UNSUPPORTED, "%s: tensor exceeds 2^31 elements", who);
  return DV_OK;
}

// ---- dv_bn_stats_static: the statistics pass for callers that must not allocate -------------------------------------
// The block partials of bn_reduce_kernel normally live in a caller-provided workspace.  Here they live in a fixed array
// inside the library image, so the call needs no scratch pointer and performs no allocation of any kind.
// Capacity: kStaticMaxBlocks partial rows of 2*C floats, C <= kStaticMaxC  ->  148 * 2 * 2048 floats = 2.3 MiB, resident on
// every GPU that loads the library until the process exits.
constexpr int kStaticMaxBlocks = 148;
constexpr int kStaticMaxC = 2048;
constexpr size_t kStaticFloats = (size_t)kStaticMaxBlocks * 2 * kStaticMaxC;
static __device__ __align__(256) float g_bn_static_scratch[kStaticFloats];

// In host code `g_bn_static_scratch` is only a handle for the symbol, NOT a device address: a kernel handed that
// name would read and write wherever the host-side shadow happens to live.  The device address is looked up once per
// device (with that device current) and cached.
static int bn_static_scratch(float** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static float* cached[kMaxDevices] = {nullptr};
  int dev = -1;
  DV_CUDA_OK(cudaGetDevice(&dev));
  DV_REQUIRE(dev >= 0 && dev < kMaxDevices, DV_ERR_UNSUPPORTED, "dv_bn_stats_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (cached[dev] == nullptr) {
    float* ptr = nullptr;
    DV_CUDA_OK(cudaGetSymbolAddress((void**)&ptr, g_bn_static_scratch));
    DV_REQUIRE(ptr != nullptr, DV_ERR_CUDA, "dv_bn_stats_static: cudaGetSymbolAddress returned a null address");
    cached[dev] = ptr;
  }
  *out = cached[dev];
  return DV_OK;
}

template <typename T>
static int stats_static_t(const dv_bn_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  DV_REQUIRE(p->C % V == 0, DV_ERR_UNSUPPORTED, "dv_bn_stats_static: C=%d must be a multiple of %d", p->C, V);
  DV_REQUIRE(p->C <= kStaticMaxC, DV_ERR_UNSUPPORTED,
             "dv_bn_stats_static: C=%d exceeds the built-in scratch (at most %d channels); use dv_bn_stats with a workspace", p->C,
             kStaticMaxC);
  Plan pl = make_plan(p->M, p->C, V);
  if (pl.nblk > kStaticMaxBlocks) {  // same kernel, fewer and longer row ranges so the partials fit the fixed array
    pl.rows_per_blk = ceil_div64(p->M, kStaticMaxBlocks);
    pl.nblk = (int)ceil_div64(p->M, pl.rows_per_blk);
  }
  DV_REQUIRE((size_t)pl.nblk * 2 * p->C <= kStaticFloats, DV_ERR_UNSUPPORTED, "dv_bn_stats_static: %d x 2 x %d partials exceed the built-in scratch",
             pl.nblk, p->C);
  float* scratch = nullptr;
  int rc = bn_static_scratch(&scratch);
  if (rc) return rc;
  // Every partial the finalize pass reads is overwritten by the reduce pass of the same call, so nothing in the shared
  // array has to be cleared between calls.
  bn_reduce_kernel<T, 0, 0><<<pl.nblk, kThreads, 0, st>>>((const T*)p->x, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, scratch,
                                                         p->M, p->C, pl.cgb, pl.rows_par, pl.rows_per_blk);
  DV_LAUNCH_OK("bn_reduce_kernel<stats,static>");
  bn_stats_finalize_kernel<T><<<ceil_div(p->C, 32), 32 * kFinLanes, 0, st>>>((const T*)p->x, scratch, pl.nblk, p->C, p->M, p->eps,
                                                                            p->momentum, p->save_mean, p->save_invstd, p->running_mean,
                                                                            p->running_var);
  DV_LAUNCH_OK("bn_stats_finalize_kernel<static>");
  return DV_OK;
}

}  // namespace
}  // namespace dv

using namespace dv;

extern "C" size_t dv_bn_workspace_bytes(int64_t M, int C) {
  // sized for the narrower vector (fp32, 4 channels/thread) so one buffer serves both dtypes
  Plan a = make_plan(M, C, 4), b = make_plan(M, C, 8);
  const int nb = a.nblk > b.nblk ? a.nblk : b.nblk;
  return (size_t)nb * 2 * C * sizeof(float);
}

extern "C" int dv_bn_stats(const dv_bn_params* p, void* stream) {
  DV_REQUIRE(p && p->x && p->save_mean && p->save_invstd && p->M > 0 && p->C > 0, DV_ERR_ARG, "dv_bn_stats: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  return p->dtype == DV_BF16 ? stats_t<bf16>(p, st) : stats_t<float>(p, st);
}
extern "C" int dv_bn_stats_from_sums(const dv_bn_params* p, float* sums, void* stream) {
  DV_REQUIRE(p && sums && p->save_mean && p->save_invstd && p->M > 0 && p->C > 0, DV_ERR_ARG, "dv_bn_stats_from_sums: bad arguments");
  bn_from_sums_kernel<<<ceil_div(p->C, 128), 128, 0, (cudaStream_t)stream>>>(sums, p->C, p->M, p->eps, p->momentum, p->save_mean,
                                                                             p->save_invstd, p->running_mean, p->running_var);
  DV_LAUNCH_OK("bn_from_sums_kernel");
  return DV_OK;
}
extern "C" int dv_bn_apply(const dv_bn_params* p, void* stream) {
  DV_REQUIRE(p && p->x && p->y && p->gamma && p->beta && p->M > 0 && p->C > 0, DV_ERR_ARG, "dv_bn_apply: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  return p->dtype == DV_BF16 ? apply_t<bf16>(p, st) : apply_t<float>(p, st);
}
extern "C" int dv_bn_bwd_reduce(const dv_bn_bwd_params* p, void* stream) {
  DV_REQUIRE(p && p->x && p->dy && p->save_mean && p->save_invstd && p->dgamma && p->dbeta && p->M > 0 && p->C > 0, DV_ERR_ARG,
             "dv_bn_bwd_reduce: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  return p->dtype == DV_BF16 ? bwd_reduce_t<bf16>(p, st) : bwd_reduce_t<float>(p, st);
}
extern "C" int dv_bn_bwd_apply(const dv_bn_bwd_params* p, void* stream) {
  DV_REQUIRE(p && p->x && p->dy && p->dx && p->gamma && p->save_mean && p->save_invstd && p->dgamma && p->dbeta && p->M > 0 && p->C > 0,
             DV_ERR_ARG, "dv_bn_bwd_apply: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  return p->dtype == DV_BF16 ? bwd_apply_t<bf16>(p, st) : bwd_apply_t<float>(p, st);
}

extern "C" int dv_relu_fwd(const void* x, void* y, int64_t n, int dtype, void* stream) {
  DV_REQUIRE(x && y && n > 0, DV_ERR_ARG, "dv_relu_fwd: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == DV_BF16) {
    DV_REQUIRE(n % 8 == 0, DV_ERR_UNSUPPORTED, "dv_relu_fwd: n must be a multiple of 8");
    relu_fwd_kernel<bf16><<<stream_grid(n / 8, 8), kThreads, 0, st>>>((const bf16*)x, (bf16*)y, n / 8);
  } else {
    DV_REQUIRE(n % 4 == 0, DV_ERR_UNSUPPORTED, "dv_relu_fwd: n must be a multiple of 4");
    relu_fwd_kernel<float><<<stream_grid(n / 4, 8), kThreads, 0, st>>>((const float*)x, (float*)y, n / 4);
  }
  DV_LAUNCH_OK("relu_fwd_kernel");
  return DV_OK;
}
extern "C" int dv_relu_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype, void* stream) {
  DV_REQUIRE(y && dy && dx && n > 0, DV_ERR_ARG, "dv_relu_bwd: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == DV_BF16) {
    DV_REQUIRE(n % 8 == 0, DV_ERR_UNSUPPORTED, "dv_relu_bwd: n must be a multiple of 8");
    relu_bwd_kernel<bf16><<<stream_grid(n / 8, 8), kThreads, 0, st>>>((const bf16*)y, (const bf16*)dy, (bf16*)dx, n / 8);
  } else {
    DV_REQUIRE(n % 4 == 0, DV_ERR_UNSUPPORTED, "dv_relu_bwd: n must be a multiple of 4");
    relu_bwd_kernel<float><<<stream_grid(n / 4, 8), kThreads, 0, st>>>((const float*)y, (const float*)dy, (float*)dx, n / 4);
  }
  DV_LAUNCH_OK("relu_bwd_kernel");
  return DV_OK;
}

extern "C" int dv_bn_stats_static(const dv_bn_params* p, void* stream) {
  DV_REQUIRE(p && p->x && p->save_mean && p->save_invstd && p->M > 0 && p->C > 0, DV_ERR_ARG, "dv_bn_stats_static: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  return p->dtype == DV_BF16 ? stats_static_t<bf16>(p, st) : stats_static_t<float>(p, st);
}

// ---- eval-mode BatchNorm folded into the preceding convolution (dv_conv_params.ep_*, dv_dwconv_params.ep_shift) -----------
namespace dv {
namespace {
__global__ void bn_fold_kernel(const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                               const float* __restrict__ var, float eps, float* __restrict__ scale, float* __restrict__ shift, int C,
                               const float* __restrict__ dw_in, float* __restrict__ dw_out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float sc = gamma[c] * rsqrtf(var[c] + eps);  // the expressions of bn_apply_kernel's eval mode
  scale[c] = sc;
  shift[c] = beta[c] - mean[c] * sc;
  if (dw_in != nullptr) {
#pragma unroll
    for (int t = 0; t < 9; ++t) dw_out[c * 9 + t] = dw_in[c * 9 + t] * sc;
  }
}
}  // namespace
}  // namespace dv

extern "C" int dv_bn_fold(const float* gamma, const float* beta, const float* running_mean, const float* running_var, float eps,
                          float* scale, float* shift, int C, const float* dw_w, float* dw_w_scaled, void* stream) {
  DV_REQUIRE(gamma && beta && running_mean && running_var && scale && shift && C > 0, DV_ERR_ARG, "dv_bn_fold: bad arguments");
  DV_REQUIRE((dw_w == nullptr) == (dw_w_scaled == nullptr), DV_ERR_ARG, "dv_bn_fold: dw_w and dw_w_scaled go together");
  dv::bn_fold_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(gamma, beta, running_mean, running_var, eps, scale, shift, C, dw_w,
                                                                       dw_w_scaled);
  DV_LAUNCH_OK("bn_fold_kernel");
  return DV_OK;
}

extern "C" size_t dv_bn_pool_workspace_bytes(const dv_bn_pool_params* p) {
  if (!p || p->N <= 0 || p->H <= 0 || p->W <= 0 || p->C <= 0) return 0;
  const int V = p->dtype == DV_BF16 ? 8 : 4;
  const int64_t total = (int64_t)p->N * ((p->H + 1) / 2) * ((p->W + 1) / 2) * (p->C / V);
  return (size_t)bn_pool_blocks(total) * 2 * p->C * sizeof(float);
}
extern "C" int dv_bn_relu_maxpool_fwd(const dv_bn_pool_params* p, void* stream) {
  int rc = bn_pool_check(p, "dv_bn_relu_maxpool_fwd");
  if (rc) return rc;
  DV_REQUIRE(p->y != nullptr, DV_ERR_ARG, "dv_bn_relu_maxpool_fwd: y is null");
  return p->dtype == DV_BF16 ? bn_pool_fwd_t<bf16>(p, (cudaStream_t)stream) : bn_pool_fwd_t<float>(p, (cudaStream_t)stream);
}
extern "C" int dv_bn_relu_maxpool_bwd(const dv_bn_pool_params* p, void* stream) {
  int rc = bn_pool_check(p, "dv_bn_relu_maxpool_bwd");
  if (rc) return rc;
  DV_REQUIRE(p->dy && p->dx && p->dgamma && p->dbeta, DV_ERR_ARG, "dv_bn_relu_maxpool_bwd: dy / dx / dgamma / dbeta is null");
  return p->dtype == DV_BF16 ? bn_pool_bwd_t<bf16>(p, (cudaStream_t)stream) : bn_pool_bwd_t<float>(p, (cudaStream_t)stream);
}
