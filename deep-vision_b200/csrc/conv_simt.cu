// conv_simt.cu — CUDA-core implicit-GEMM convolution (fprop / dgrad / wgrad), any shape, fp32 or bf16 storage,
// fp32 accumulation.  This is the fp32 parity path (a single-pass TF32 tensor-core product cannot meet the 1e-4
// bound) and the path for shapes the tcgen05 kernels do not take (C=3 stems, C%64!=0, 3x3 stride-2
// dgrad).  It is a GPU kernel like the others: there is no host fallback anywhere in this library.
//
// One 64x64 output tile per CTA, 256 threads, 4x4 outputs per thread, K stepped by 16 through shared memory.
// The three passes differ only in how a GEMM coordinate maps to a tensor element:
//   fprop: M=(n,p,q)  N=k        K=(r,s,c)    A=x (gathered)   B=w
//   dgrad: M=(n,h,w)  N=c        K=(r,s,k)    A=dy (gathered)  B=w
//   wgrad: M=k        N=(r,s,c)  K=(n,p,q)    A=dy             B=x (gathered)   split over K, fp32 atomics
#include "common.cuh"

namespace dv {

struct ConvDims {
  int N, H, W, C, K, R, S, stride, pad, P, Q;
};

template <int PASS, typename T>
struct Operand {
  const T* a;
  const T* b;
  ConvDims d;

  __device__ __forceinline__ float A(int64_t m, int64_t k) const {
    if (PASS == DV_CONV_FPROP) {
      const int pq = d.P * d.Q;
      const int n = (int)(m / pq), rem = (int)(m % pq);
      const int p = rem / d.Q, q = rem % d.Q;
      const int tap = (int)(k / d.C), c = (int)(k % d.C);
      const int r = tap / d.S, s = tap % d.S;
      const int h = p * d.stride + r - d.pad, w = q * d.stride + s - d.pad;
      if ((unsigned)h >= (unsigned)d.H || (unsigned)w >= (unsigned)d.W) return 0.f;
      return to_f32(a[(((size_t)n * d.H + h) * d.W + w) * d.C + c]);
    } else if (PASS == DV_CONV_DGRAD) {
      const int hw = d.H * d.W;
      const int n = (int)(m / hw), rem = (int)(m % hw);
      const int h = rem / d.W, w = rem % d.W;
      const int tap = (int)(k / d.K), ko = (int)(k % d.K);
      const int r = tap / d.S, s = tap % d.S;
      const int hp = h + d.pad - r, wq = w + d.pad - s;
      if (hp < 0 || wq < 0 || hp % d.stride || wq % d.stride) return 0.f;
      const int p = hp / d.stride, q = wq / d.stride;
      if (p >= d.P || q >= d.Q) return 0.f;
      return to_f32(a[(((size_t)n * d.P + p) * d.Q + q) * d.K + ko]);
    } else {
      return to_f32(a[(size_t)k * d.K + m]);
    }
  }
  __device__ __forceinline__ float B(int64_t k, int64_t n) const {
    if (PASS == DV_CONV_FPROP) {
      return to_f32(b[(size_t)n * ((size_t)d.R * d.S * d.C) + k]);
    } else if (PASS == DV_CONV_DGRAD) {
      const int tap = (int)(k / d.K), ko = (int)(k % d.K);
      return to_f32(b[((size_t)ko * d.R * d.S + tap) * d.C + n]);
    } else {
      const int pq = d.P * d.Q;
      const int ni = (int)(k / pq), rem = (int)(k % pq);
      const int p = rem / d.Q, q = rem % d.Q;
      const int tap = (int)(n / d.C), c = (int)(n % d.C);
      const int r = tap / d.S, s = tap % d.S;
      const int h = p * d.stride + r - d.pad, w = q * d.stride + s - d.pad;
      if ((unsigned)h >= (unsigned)d.H || (unsigned)w >= (unsigned)d.W) return 0.f;
      return to_f32(b[(((size_t)ni * d.H + h) * d.W + w) * d.C + c]);
    }
  }
};

// A_KC / B_KC: whether the operand's memory-contiguous index is the GEMM K index (then consecutive threads walk K).
template <int PASS, typename T, typename TO>
__global__ void __launch_bounds__(256)
conv_simt_kernel(Operand<PASS, T> op, TO* __restrict__ out, int64_t M, int64_t Ng, int64_t Kg, int64_t k_per_split, int atomic) {
  constexpr bool A_KC = (PASS != DV_CONV_WGRAD);
  constexpr bool B_KC = (PASS == DV_CONV_FPROP);
  __shared__ __align__(16) float As[16][68];
  __shared__ __align__(16) float Bs[16][68];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * 64, n0 = (int64_t)blockIdx.y * 64;
  const int64_t kbeg = (int64_t)blockIdx.z * k_per_split;
  const int64_t kend = (kbeg + k_per_split < Kg) ? kbeg + k_per_split : Kg;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int64_t k0 = kbeg; k0 < kend; k0 += 16) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + i * 256;
      int kk, mm;
      if (A_KC) { kk = idx & 15; mm = idx >> 4; } else { mm = idx & 63; kk = idx >> 6; }
      const int64_t m = m0 + mm, k = k0 + kk;
      As[kk][mm] = (m < M && k < kend) ? op.A(m, k) : 0.f;
      int kb, nn;
      if (B_KC) { kb = idx & 15; nn = idx >> 4; } else { nn = idx & 63; kb = idx >> 6; }
      const int64_t n = n0 + nn, k2 = k0 + kb;
      Bs[kb][nn] = (n < Ng && k2 < kend) ? op.B(k2, n) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      const float4 av = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 bv = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float a[4] = {av.x, av.y, av.z, av.w};
      const float b[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t n = n0 + tx * 4 + j;
      if (n >= Ng) continue;
      if (PASS == DV_CONV_WGRAD) {
        if (atomic) atomicAdd(reinterpret_cast<float*>(out) + m * Ng + n, acc[i][j]);
        else reinterpret_cast<float*>(out)[m * Ng + n] = acc[i][j];
      } else if (atomic) {  // dgrad with accumulate: dx += result
        out[m * Ng + n] = from_f32<TO>(acc[i][j] + to_f32(out[m * Ng + n]));
      } else {
        out[m * Ng + n] = from_f32<TO>(acc[i][j]);
      }
    }
  }
}

template <int PASS, typename T, typename TO>
static int launch(const dv_conv_params* p, const ConvDims& d, cudaStream_t st) {
  Operand<PASS, T> op;
  op.d = d;
  int64_t M, Ng, Kg;
  void* out;
  if (PASS == DV_CONV_FPROP) {
    op.a = (const T*)p->x; op.b = (const T*)p->w; out = p->y;
    M = (int64_t)d.N * d.P * d.Q; Ng = d.K; Kg = (int64_t)d.R * d.S * d.C;
  } else if (PASS == DV_CONV_DGRAD) {
    op.a = (const T*)p->y; op.b = (const T*)p->w; out = p->x;
    M = (int64_t)d.N * d.H * d.W; Ng = d.C; Kg = (int64_t)d.R * d.S * d.K;
  } else {
    op.a = (const T*)p->y; op.b = (const T*)p->x; out = p->w;
    M = d.K; Ng = (int64_t)d.R * d.S * d.C; Kg = (int64_t)d.N * d.P * d.Q;
  }
  const int64_t gx = ceil_div64(M, 64), gy = ceil_div64(Ng, 64);
  int splits = 1;
  if (PASS == DV_CONV_WGRAD) {
    // fill ~4 waves of the machine; each split at least 256 reduction steps
    const int64_t want = (4LL * num_sms() + gx * gy - 1) / (gx * gy);
    const int64_t max_by_k = (Kg + 255) / 256;
    splits = (int)(want < max_by_k ? want : max_by_k);
    if (splits < 1) splits = 1;
    if (splits > 65535) splits = 65535;
  }
  int64_t k_per_split = ceil_div64(Kg, splits);
  k_per_split = ceil_div64(k_per_split, 16) * 16;
  splits = (int)ceil_div64(Kg, k_per_split);
  if (PASS == DV_CONV_WGRAD && splits > 1 && !p->accumulate) {
    DV_CUDA_OK(cudaMemsetAsync(out, 0, (size_t)M * Ng * sizeof(float), st));
  }
  DV_REQUIRE(gy <= 65535, DV_ERR_UNSUPPORTED, "conv_simt: N tiles %lld exceed grid.y", (long long)gy);
  dim3 grid((unsigned)gx, (unsigned)gy, (unsigned)splits);
  const int flag = (PASS == DV_CONV_WGRAD) ? ((splits > 1 || p->accumulate) ? 1 : 0) : (PASS == DV_CONV_DGRAD && p->accumulate ? 1 : 0);
  conv_simt_kernel<PASS, T, TO><<<grid, 256, 0, st>>>(op, (TO*)out, M, Ng, Kg, k_per_split, flag);
  DV_LAUNCH_OK("conv_simt_kernel");
  return DV_OK;
}

int conv_simt(const dv_conv_params* p, int pass, cudaStream_t st) {
  ConvDims d;
  d.N = p->N; d.H = p->H; d.W = p->W; d.C = p->C; d.K = p->K; d.R = p->R; d.S = p->S;
  d.stride = p->stride; d.pad = p->pad;
  d.P = (p->H + 2 * p->pad - p->R) / p->stride + 1;
  d.Q = (p->W + 2 * p->pad - p->S) / p->stride + 1;
  const bool bf = p->dtype == DV_BF16;
  switch (pass) {
    case DV_CONV_FPROP:
      if (!bf) return launch<DV_CONV_FPROP, float, float>(p, d, st);
      if (p->y_f32) return launch<DV_CONV_FPROP, bf16, float>(p, d, st);
      return launch<DV_CONV_FPROP, bf16, bf16>(p, d, st);
    case DV_CONV_DGRAD:
      return bf ? launch<DV_CONV_DGRAD, bf16, bf16>(p, d, st) : launch<DV_CONV_DGRAD, float, float>(p, d, st);
    case DV_CONV_WGRAD:
      return bf ? launch<DV_CONV_WGRAD, bf16, float>(p, d, st) : launch<DV_CONV_WGRAD, float, float>(p, d, st);
  }
  set_error("conv_simt: bad pass %d", pass);
  return DV_ERR_ARG;
}

}  // namespace dv
