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
