// dwconv.cu — depth-wise 3x3 convolution (pad 1, stride 1|2, groups == C) for MobileNetV1's DepthwiseConv
// (MobileNet/pytorch/models/mobilenet_v1.py:109-133).  Arithmetic intensity is ~3.5 FLOP/B, so the design target is
// the HBM roof: NHWC, one 16-byte channel vector per thread, consecutive threads on consecutive channels (coalesced
// 128 B lines), the 9 taps of a channel slab cached in shared memory.  Each thread produces a STRIP of 4 neighbouring
// outputs along W and keeps the input columns they share in registers (18 loads instead of 36 for stride 1), all index
// arithmetic is 32-bit, the stride is a template constant, the FMAs are packed (fma.rn.f32x2) and grids are one wave of
// resident CTAs.  Measured 0.1-0.6 of the HBM roof (profiles/r01_layer_bench_v10_depthwise.md): the kernels are bound by
// instruction issue and load latency at 16-24 warps per SM, see DESIGN.md section 3.2 for what was tried.
//   fprop : y[n,p,q,c]  = sum_{r,s} x[n, p*st+r-1, q*st+s-1, c] * w[c,r,s]
//   dgrad : stride 1: the same kernel on dy with the taps flipped;
//           stride 2: one thread per 2x2 block of dx, which depends on exactly dy[i..i+1][j..j+1]
//   wgrad : dw[c,r,s]   = sum_{n,p,q} dy[n,p,q,c] * x[n, p*st+r-1, q*st+s-1, c]      (block partials -> fixed-order reduce)
#include "common.cuh"
#include <cuda.h>  // CUtensorMap (type only)
#ifndef DV_DW_RW
#define DV_DW_RW 1   // output rows per thread of the TMA-staged stride-1 kernel (2 measured the same: the kernel is bound by the fp32 pipe, not by loads)
#endif

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

template <int V>
__device__ __forceinline__ void load_taps(float (*ws)[128 * V], const float* __restrict__ w, int c0, int nch, bool flip) {
  for (int i = threadIdx.x; i < 9 * nch; i += kThreads) {
    const int ch = i / 9, tap = i - ch * 9;
    ws[flip ? 8 - tap : tap][ch] = w[(size_t)(c0 + ch) * 9 + tap];
  }
}

// Same taps as [tap][16-byte piece (V/4)][channel group][4 floats]: a lane's float4 reads sit 16 B from its neighbours'
// (conflict-free LDS.128); with [tap][channel] they sit 32 B apart for V = 8, a 2-way bank conflict on every read.
template <int V>
__device__ __forceinline__ void load_taps_split(float* ws, const float* __restrict__ w, int c0, int nch, bool flip, int cgb) {
  for (int i = threadIdx.x; i < 9 * nch; i += kThreads) {
    const int ch = i / 9, tap = i - ch * 9;
    const int t = flip ? 8 - tap : tap;
    const int g = ch / V, e = ch % V;
    ws[((t * (V / 4) + e / 4) * cgb + g) * 4 + (e & 3)] = w[(size_t)(c0 + ch) * 9 + tap];
  }
}

// out[n,oh,ow,c] = sum_{r,s} in[n, oh*ST+r-1, ow*ST+s-1, c] * taps[r][s]   (in: [N,IH,IW,C], out: [N,OH,OW,C])
// outputs per thread along W: 4 for stride 1 (6 input columns), 2 for stride 2 (5 input columns) and for wgrad, whose 72
// accumulators already fill the register budget
template <int ST> struct StripW { static constexpr int fwd = ST == 1 ? 4 : 2; static constexpr int wgrad = 2; };

// packed fp32 pairs: sm_100's fma.rn.f32x2 does two channels per instruction (ncu on the scalar version: issue slots 67 %
// busy at 43 % of the HBM roof — the depth-wise kernels are bound by instruction issue before they are bound by memory)
__device__ __forceinline__ void unpack2(const Vec16<float>& v, float2* f) { f[0] = make_float2(v.raw.x, v.raw.y); f[1] = make_float2(v.raw.z, v.raw.w); }
__device__ __forceinline__ void unpack2(const Vec16<bf16>& v, float2* f) {
  const uint32_t w[4] = {v.raw.x, v.raw.y, v.raw.z, v.raw.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) f[i] = make_float2(__uint_as_float(w[i] << 16), __uint_as_float(w[i] & 0xffff0000u));
}
__device__ __forceinline__ void pack2(Vec16<float>& v, const float2* f) { v.raw = make_float4(f[0].x, f[0].y, f[1].x, f[1].y); }
__device__ __forceinline__ void pack2(Vec16<bf16>& v, const float2* f) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_bfloat162 h = __floats2bfloat162_rn(f[i].x, f[i].y);
    w[i] = *reinterpret_cast<uint32_t*>(&h);
  }
  v.raw = make_uint4(w[0], w[1], w[2], w[3]);
}

// CTAs per SM: 2 for stride 1 (4 outputs per thread need the 128 registers), 3 for stride 2 (measured 10 % faster than 2;
// stride 1 with 2 outputs per thread at 3 or 4 CTAs per SM measured 12-25 % slower than this)
template <typename T, int ST>
__global__ void __launch_bounds__(kThreads, ST == 1 ? 2 : 3)
dwconv_strip_kernel(const T* __restrict__ in, const float* __restrict__ w, T* __restrict__ out, int N, int IH, int IW, int OH, int OW,
                    int C, int cgb, int lanes, int strips_per_blk, int flip, const float* __restrict__ ep_shift = nullptr, int ep_relu = 0) {
  constexpr int V = Vec16<T>::N;
  constexpr int TW = StripW<ST>::fwd;
  constexpr int NCOL = (TW - 1) * ST + 3;  // input columns a strip touches
  __shared__ __align__(16) float ws[9 * 128 * V];
  const int cg0 = blockIdx.y * cgb;
  load_taps_split<V>(ws, w, cg0 * V, min(cgb * V, C - cg0 * V), flip != 0, cgb);
  __syncthreads();
  const int tcg = threadIdx.x % cgb, lane = threadIdx.x / cgb;
  if (lane >= lanes || (cg0 + tcg) * V >= C) return;
  const int c = (cg0 + tcg) * V;
  const int strips_w = (OW + TW - 1) / TW;
  const int nstrips = N * OH * strips_w;
  const int s0 = blockIdx.x * strips_per_blk;
  const int s1 = min(nstrips, s0 + strips_per_blk);
  constexpr int H2 = V / 2;
  float2 bias[H2];  // folded-BatchNorm shift (inference) or zero
#pragma unroll
  for (int e = 0; e < H2; ++e) bias[e] = ep_shift != nullptr ? make_float2(ep_shift[c + 2 * e], ep_shift[c + 2 * e + 1]) : make_float2(0.f, 0.f);
  for (int sidx = s0 + lane; sidx < s1; sidx += lanes) {
    const int sw = sidx % strips_w;
    const int t = sidx / strips_w;
    const int oh = t % OH, n = t / OH;
    const int ow0 = sw * TW;
    float2 acc[TW][H2];
#pragma unroll
    for (int j = 0; j < TW; ++j)
#pragma unroll
      for (int e = 0; e < H2; ++e) acc[j][e] = bias[e];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int ih = oh * ST + r - 1;
      if ((unsigned)ih >= (unsigned)IH) continue;
      const T* row = in + ((size_t)n * IH + ih) * IW * C + c;
      Vec16<T> raw[NCOL];
#pragma unroll
      for (int k = 0; k < NCOL; ++k) {
        const int iw = ow0 * ST - 1 + k;
        if ((unsigned)iw < (unsigned)IW) raw[k].load(row + (size_t)iw * C);
      }
      float2 xv[NCOL][H2];
#pragma unroll
      for (int k = 0; k < NCOL; ++k) {
        const int iw = ow0 * ST - 1 + k;
        if ((unsigned)iw < (unsigned)IW) unpack2(raw[k], xv[k]);
        else {
#pragma unroll
          for (int e = 0; e < H2; ++e) xv[k][e] = make_float2(0.f, 0.f);
        }
      }
      // one shared-memory read of each tap's weights per strip (not per output): the taps are the outer loop
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        float2 wv[H2];
#pragma unroll
        for (int e = 0; e < V; e += 4) {
          const float4 q4 = *reinterpret_cast<const float4*>(ws + (((r * 3 + s) * (V / 4) + e / 4) * cgb + tcg) * 4);
          wv[e / 2] = make_float2(q4.x, q4.y); wv[e / 2 + 1] = make_float2(q4.z, q4.w);
        }
#pragma unroll
        for (int j = 0; j < TW; ++j) {
#pragma unroll
          for (int e = 0; e < H2; ++e) acc[j][e] = __ffma2_rn(xv[j * ST + s][e], wv[e], acc[j][e]);
        }
      }
    }
    T* orow = out + (((size_t)n * OH + oh) * OW + ow0) * C + c;
    if (ep_relu) {
#pragma unroll
      for (int j = 0; j < TW; ++j)
#pragma unroll
        for (int e = 0; e < H2; ++e) acc[j][e] = make_float2(fmaxf(acc[j][e].x, 0.f), fmaxf(acc[j][e].y, 0.f));
    }
#pragma unroll
    for (int j = 0; j < TW; ++j) {
      if (ow0 + j < OW) { Vec16<T> o; pack2(o, acc[j]); o.store(orow + (size_t)j * C); }
    }
  }
}

// ---- TMA-staged variant (bf16, stride 1) -----------------------------------------------------------------------------
// The strip kernel above keeps its 18 input vectors per strip in REGISTERS while they are in flight: at 128 registers only 16
// warps are resident and ncu shows them waiting on global loads with the HBM half idle.  Here the input halo tile of a
// (TH x 16 pixel, CB channel) output tile is brought into shared memory by one TMA box load per tile — a 4-D tiled map over
// [N][H][W][C] whose out-of-range rows / columns read as zero, so padding costs nothing — through a ring of STAGES tiles:
// the bytes in flight live in shared memory, not in registers, and the arithmetic reads 16-byte vectors from the tile
// (a quarter-warp reads one pixel's 128 contiguous bytes: conflict-free).  One thread per strip of 4 outputs x 8 channels.
__device__ __forceinline__ uint32_t dw_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void dw_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dw_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void dw_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(dw_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void dw_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(dw_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void dw_mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = dw_smem_u32(bar);
  uint32_t done = 0;
  long long t0 = 0;
  for (uint32_t spin = 0;; ++spin) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    if (done) return;
    if ((spin & 0x3ff) == 0x3ff) {   // a protocol bug must trap, not hang the GPU
      const long long now = clock64();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 4000000000LL) { printf("dv dwconv: mbarrier wait timed out (block %d,%d)\n", blockIdx.x, blockIdx.y); __trap(); }
    }
  }
}
__device__ __forceinline__ void dw_tma_load_4d(void* smem, const void* tm, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
                   dw_smem_u32(smem)),
               "l"(reinterpret_cast<uint64_t>(tm)), "r"(dw_smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

struct DwTmaArgs {
  int N, OH, OW, C;        // output extents (== input extents for stride 1)
  int tiles_h, tiles_w;
  int flip, ep_relu;
  const float* ep_shift;
  float* stat_sums;        // fprop, optional: [2][C] += per-channel sum and sum of squares of the bf16-rounded outputs (BatchNorm statistics)
  // stride-1 dgrad, optional (template BNB): the output dx is the gradient w.r.t. relu(bn(x0)); the kernel ADDS that BatchNorm's
  // backward sums  dbeta += sum dz,  dgamma += invstd * sum dz * (x0 - mean),  dz = dx * (bn(x0) > 0)  (as dv_conv_params.bnb_*)
  const bf16* bnb_x;
  const float* bnb_gamma;
  const float* bnb_beta;
  const float* bnb_mean;
  const float* bnb_invstd;
  float* bnb_dgamma;
  float* bnb_dbeta;
};

template <int CB, int ST = 1, int RW = 1>
struct DwTile {
  static constexpr int GROUPS = CB / 8;            // 16-byte channel groups per pixel
  static constexpr int NSTRIP = kThreads / GROUPS;  // strips (4 outputs wide, RW rows high) per tile
  static constexpr int TW = 16, TH = RW * NSTRIP / 4;    // output tile
  static constexpr int IW = (TW - 1) * ST + 3, IH = (TH - 1) * ST + 3;    // input halo tile
  static constexpr int STAGE_BYTES = ((IH * IW * CB * 2 + 127) / 128) * 128;
  static constexpr int LOAD_BYTES = IH * IW * CB * 2;
};

// RW = output rows per thread (stride 1 only): with two rows a thread reads 4 input rows for 8 outputs instead of 6, and the strip
// kernel's instruction count per output, which is what bounds the stride-1 case, drops by a third.
template <int CB, int ST, int RW, int STAGES, bool BNB = false>
__global__ void __launch_bounds__(kThreads, (ST == 1 && RW == 1) ? 2 : 1)
dwconv_tma_fwd_kernel(const __grid_constant__ CUtensorMap tmX, const float* __restrict__ w, bf16* __restrict__ out, const DwTmaArgs g) {
  static_assert(RW == 1 || ST == 1, "two rows per thread exist for stride 1 only");
  using TL = DwTile<CB, ST, RW>;
  constexpr int V = 8, H2 = 4, NCOL = 3 * ST + 3;   // input columns under a strip of 4 outputs
  extern __shared__ __align__(128) uint8_t dw_smem[];
  uint8_t* tiles = dw_smem;
  float* ws = reinterpret_cast<float*>(dw_smem + STAGES * TL::STAGE_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ws + 9 * CB);
  uint64_t* empty_bar = full_bar + STAGES;
  const int c0 = blockIdx.y * CB;
  load_taps_split<V>(ws, w, c0, CB, g.flip != 0, TL::GROUPS);
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { dw_mbar_init(&full_bar[i], 1); dw_mbar_init(&empty_bar[i], kThreads / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int per_img = g.tiles_h * g.tiles_w;
  const int ntiles = g.N * per_img;
  const int my_tiles = blockIdx.x < ntiles ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  auto issue = [&](int it) {   // thread 0: tile `it` of this CTA into stage it % STAGES
    const int tile = blockIdx.x + it * gridDim.x;
    const int n = tile / per_img, rem = tile - n * per_img;
    const int th = rem / g.tiles_w, tw = rem - th * g.tiles_w;
    const int s = it % STAGES;
    dw_mbar_expect_tx(&full_bar[s], TL::LOAD_BYTES);
    dw_tma_load_4d(tiles + s * TL::STAGE_BYTES, &tmX, &full_bar[s], c0, tw * TL::TW * ST - 1, th * TL::TH * ST - 1, n);
  };
  if (threadIdx.x == 0)
    for (int it = 0; it < STAGES && it < my_tiles; ++it) issue(it);

  const int tcg = threadIdx.x % TL::GROUPS, strip = threadIdx.x / TL::GROUPS;
  const int row = strip >> 2, col0 = (strip & 3) * 4;
  const int c = c0 + tcg * V;
  float2 bias[H2];
#pragma unroll
  for (int e = 0; e < H2; ++e) bias[e] = g.ep_shift != nullptr ? make_float2(g.ep_shift[c + 2 * e], g.ep_shift[c + 2 * e + 1]) : make_float2(0.f, 0.f);
  float2 st1[H2], st2[H2];   // running sum / sum of squares of what this thread stored  (BNB: sum dz / sum dz * (x0 - mean))
#pragma unroll
  for (int e = 0; e < H2; ++e) { st1[e] = make_float2(0.f, 0.f); st2[e] = make_float2(0.f, 0.f); }
  float2 bsc[BNB ? H2 : 1], bsh[BNB ? H2 : 1], bnmu[BNB ? H2 : 1];   // bn_apply's expressions: the mask bits agree with its output
  if (BNB) {
#pragma unroll
    for (int e = 0; e < H2; ++e) {
      const int ch = c + 2 * e < g.C ? c + 2 * e : 0;
      const float mu0 = g.bnb_mean[ch], mu1 = g.bnb_mean[ch + 1];
      bsc[e] = make_float2(g.bnb_invstd[ch] * g.bnb_gamma[ch], g.bnb_invstd[ch + 1] * g.bnb_gamma[ch + 1]);
      bsh[e] = make_float2(g.bnb_beta[ch] - mu0 * bsc[e].x, g.bnb_beta[ch + 1] - mu1 * bsc[e].y);
      bnmu[e] = make_float2(-mu0, -mu1);
    }
  }

  for (int it = 0; it < my_tiles; ++it) {
    const int s = it % STAGES;
    const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
    if (BNB) {   // the x0 vectors this thread will need behind its stores: start them towards the L2 before the tile's arithmetic
      const int tile_ = blockIdx.x + it * gridDim.x;
      const int n_ = tile_ / per_img, rem_ = tile_ - n_ * per_img;
      const int th_ = rem_ / g.tiles_w, tw_ = rem_ - th_ * g.tiles_w;
      const int oh_ = th_ * TL::TH + row * RW, ow_ = tw_ * TL::TW + col0;
      if (oh_ < g.OH && c < g.C) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (ow_ + j < g.OW)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(g.bnb_x + (((size_t)n_ * g.OH + oh_) * g.OW + ow_ + j) * g.C + c));
      }
    }
    dw_mbar_wait(&full_bar[s], ph);
    const uint8_t* tile_in = tiles + s * TL::STAGE_BYTES;
    float2 acc[RW][4][H2];
#pragma unroll
    for (int o = 0; o < RW; ++o)
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < H2; ++e) acc[o][j][e] = bias[e];
#pragma unroll
    for (int ir = 0; ir < 2 + RW; ++ir) {   // input rows under this thread's RW output rows
      float2 xv[NCOL][H2];
#pragma unroll
      for (int k = 0; k < NCOL; ++k) {
        Vec16<bf16> raw;
        raw.raw = *reinterpret_cast<const uint4*>(tile_in + (((row * RW * ST + ir) * TL::IW + col0 * ST + k) * CB + tcg * V) * 2);
        unpack2(raw, xv[k]);
      }
#pragma unroll
      for (int o = 0; o < RW; ++o) {
        const int r = ir - o;                // filter row this input row meets for output row o
        if (r < 0 || r > 2) continue;
#pragma unroll
        for (int sx = 0; sx < 3; ++sx) {
          float2 wv[H2];
#pragma unroll
          for (int e = 0; e < V; e += 4) {
            const float4 q4 = *reinterpret_cast<const float4*>(ws + (((r * 3 + sx) * (V / 4) + e / 4) * TL::GROUPS + tcg) * 4);
            wv[e / 2] = make_float2(q4.x, q4.y); wv[e / 2 + 1] = make_float2(q4.z, q4.w);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < H2; ++e) acc[o][j][e] = __ffma2_rn(xv[j * ST + sx][e], wv[e], acc[o][j][e]);
        }
      }
    }
    // this warp has read everything it needs from the stage: hand it back, and let thread 0 refill it for tile it + STAGES
    __syncwarp();
    if ((threadIdx.x & 31) == 0) dw_mbar_arrive(&empty_bar[s]);
    if (threadIdx.x == 0 && it + STAGES < my_tiles) {
      dw_mbar_wait(&empty_bar[s], ph);
      issue(it + STAGES);
    }
    const int tile = blockIdx.x + it * gridDim.x;
    const int n = tile / per_img, rem = tile - n * per_img;
    const int th = rem / g.tiles_w, tw = rem - th * g.tiles_w;
    const int ow0 = tw * TL::TW + col0;
#pragma unroll
    for (int o = 0; o < RW; ++o) {
      const int oh = th * TL::TH + row * RW + o;
      if (oh < g.OH && c < g.C) {
        if (g.ep_relu) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < H2; ++e) acc[o][j][e] = make_float2(fmaxf(acc[o][j][e].x, 0.f), fmaxf(acc[o][j][e].y, 0.f));
        }
        bf16* orow = out + (((size_t)n * g.OH + oh) * g.OW + ow0) * g.C + c;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (ow0 + j < g.OW) {
            Vec16<bf16> ov;
            pack2(ov, acc[o][j]);
            ov.store(orow + (size_t)j * g.C);
            if (BNB) {   // dz from the rounded dx (what bn_bwd_apply will read back) and the matching x0 vector
              Vec16<bf16> xr;
              xr.load(g.bnb_x + (((size_t)n * g.OH + oh) * g.OW + ow0 + j) * g.C + c);
              float2 rv[H2], x0[H2];
              unpack2(ov, rv);
              unpack2(xr, x0);
#pragma unroll
              for (int e = 0; e < H2; ++e) {
                const float2 pre = __ffma2_rn(x0[e], bsc[e], bsh[e]);
                const float2 z = make_float2(pre.x > 0.f ? rv[e].x : 0.f, pre.y > 0.f ? rv[e].y : 0.f);
                st1[e] = __fadd2_rn(st1[e], z);
                st2[e] = __ffma2_rn(z, __fadd2_rn(x0[e], bnmu[e]), st2[e]);
              }
            } else if (g.stat_sums != nullptr) {   // from the rounded values: what the BatchNorm will read back
              float2 rv[H2];
              unpack2(ov, rv);
#pragma unroll
              for (int e = 0; e < H2; ++e) { st1[e] = __fadd2_rn(st1[e], rv[e]); st2[e] = __ffma2_rn(rv[e], rv[e], st2[e]); }
            }
          }
      }
    }
  }
  if (BNB || g.stat_sums != nullptr) {   // CTA-wide: strips of one channel group -> one pair of atomics per channel
    __syncthreads();
    float* red = reinterpret_cast<float*>(dw_smem);
#pragma unroll
    for (int e = 0; e < H2; ++e) {
      red[threadIdx.x * 16 + 2 * e] = st1[e].x; red[threadIdx.x * 16 + 2 * e + 1] = st1[e].y;
      red[threadIdx.x * 16 + 8 + 2 * e] = st2[e].x; red[threadIdx.x * 16 + 8 + 2 * e + 1] = st2[e].y;
    }
    __syncthreads();
    if (threadIdx.x < TL::GROUPS * 16) {
      const int grp = threadIdx.x / 16, k = threadIdx.x % 16;   // k < 8: sum of channel k, else sum of squares of channel k - 8
      float sum = 0.f;
      for (int l = 0; l < TL::NSTRIP; ++l) sum += red[(l * TL::GROUPS + grp) * 16 + k];
      const int ch = c0 + grp * V + (k & 7);
      if (ch < g.C) {
        if (BNB) {
          if (k < 8) atomicAdd(g.bnb_dbeta + ch, sum);
          else atomicAdd(g.bnb_dgamma + ch, sum * g.bnb_invstd[ch]);
        } else {
          atomicAdd(g.stat_sums + (k < 8 ? 0 : g.C) + ch, sum);
        }
      }
    }
  }
}

// TMA-staged stride-2 dgrad: the dy tile of an (8 x 8 cell) block — a cell (i, j) is the 2x2 block dx[2i..2i+1][2j..2j+1], which
// depends on dy[i..i+1][j..j+1] exactly (see dwconv_dgrad_s2_kernel below) — comes in by one TMA box (9 x 9 pixels, rows / columns
// past the tensor read as zero); a thread owns two neighbouring cells of one channel group: 6 shared-memory vectors in, 8 stores out.
template <int CB>
struct DwS2Tile {
  static constexpr int GROUPS = CB / 8, NSTRIP = kThreads / GROUPS;   // a strip = 2 cells
  static constexpr int CW = 8, CH = NSTRIP / 4;                        // cells per tile (4 strips per cell row)
  static constexpr int IW = CW + 1, IH = CH + 1;
  static constexpr int LOAD_BYTES = IH * IW * CB * 2, STAGE_BYTES = ((LOAD_BYTES + 127) / 128) * 128;
};

template <int CB, int STAGES>
__global__ void __launch_bounds__(kThreads, 2)
dwconv_tma_dgrad_s2_kernel(const __grid_constant__ CUtensorMap tmDY, const float* __restrict__ w, bf16* __restrict__ dx, int N, int H, int W,
                           int C, int tiles_h, int tiles_w) {
  using TL = DwS2Tile<CB>;
  constexpr int V = 8;
  extern __shared__ __align__(128) uint8_t dw_smem[];
  uint8_t* tiles = dw_smem;
  float* ws = reinterpret_cast<float*>(dw_smem + STAGES * TL::STAGE_BYTES);   // [9][CB] taps, tap-major
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ws + 9 * CB);
  uint64_t* empty_bar = full_bar + STAGES;
  const int c0 = blockIdx.y * CB;
  for (int i = threadIdx.x; i < 9 * CB; i += kThreads) {
    const int ch = i / 9, tap = i - ch * 9;
    ws[tap * CB + ch] = w[(size_t)(c0 + ch) * 9 + tap];
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { dw_mbar_init(&full_bar[i], 1); dw_mbar_init(&empty_bar[i], kThreads / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int per_img = tiles_h * tiles_w;
  const int ntiles = N * per_img;
  const int my_tiles = blockIdx.x < ntiles ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  auto issue = [&](int it) {
    const int tile = blockIdx.x + it * gridDim.x;
    const int n = tile / per_img, rem = tile - n * per_img;
    const int th = rem / tiles_w, tw = rem - th * tiles_w;
    const int s = it % STAGES;
    dw_mbar_expect_tx(&full_bar[s], TL::LOAD_BYTES);
    dw_tma_load_4d(tiles + s * TL::STAGE_BYTES, &tmDY, &full_bar[s], c0, tw * TL::CW, th * TL::CH, n);
  };
  if (threadIdx.x == 0)
    for (int it = 0; it < STAGES && it < my_tiles; ++it) issue(it);
  const int tcg = threadIdx.x % TL::GROUPS, strip = threadIdx.x / TL::GROUPS;
  const int crow = strip >> 2, ccol0 = (strip & 3) * 2;
  const int c = c0 + tcg * V;
  const float* wt = ws + tcg * V;
  for (int it = 0; it < my_tiles; ++it) {
    const int s = it % STAGES;
    const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
    dw_mbar_wait(&full_bar[s], ph);
    const uint8_t* t_in = tiles + s * TL::STAGE_BYTES;
    float g[2][3][V];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        Vec16<bf16> raw;
        raw.raw = *reinterpret_cast<const uint4*>(t_in + (((crow + a) * TL::IW + ccol0 + b) * CB + tcg * V) * 2);
        raw.unpack(g[a][b]);
      }
    __syncwarp();
    if ((threadIdx.x & 31) == 0) dw_mbar_arrive(&empty_bar[s]);
    if (threadIdx.x == 0 && it + STAGES < my_tiles) {
      dw_mbar_wait(&empty_bar[s], ph);
      issue(it + STAGES);
    }
    const int tile = blockIdx.x + it * gridDim.x;
    const int n = tile / per_img, rem = tile - n * per_img;
    const int th = rem / tiles_w, tw = rem - th * tiles_w;
    const int h = 2 * (th * TL::CH + crow);
#pragma unroll
    for (int cell = 0; cell < 2; ++cell) {
      const int x = 2 * (tw * TL::CW + ccol0 + cell);
      if (h >= H || x >= W || c >= C) continue;
      float o00[V], o01[V], o10[V], o11[V];
#pragma unroll
      for (int e = 0; e < V; ++e) {
        const float w00 = wt[0 * CB + e], w01 = wt[1 * CB + e], w02 = wt[2 * CB + e];
        const float w10 = wt[3 * CB + e], w11 = wt[4 * CB + e], w12 = wt[5 * CB + e];
        const float w20 = wt[6 * CB + e], w21 = wt[7 * CB + e], w22 = wt[8 * CB + e];
        const float g00 = g[0][cell][e], g01 = g[0][cell + 1][e], g10 = g[1][cell][e], g11 = g[1][cell + 1][e];
        o00[e] = g00 * w11;
        o01[e] = fmaf(g00, w12, g01 * w10);
        o10[e] = fmaf(g00, w21, g10 * w01);
        o11[e] = fmaf(g00, w22, fmaf(g01, w20, fmaf(g10, w02, g11 * w00)));
      }
      bf16* base = dx + (((size_t)n * H + h) * W + x) * C + c;
      Vec16<bf16> o;
      o.pack(o00); o.store(base);
      if (x + 1 < W) { o.pack(o01); o.store(base + C); }
      if (h + 1 < H) {
        o.pack(o10); o.store(base + (size_t)W * C);
        if (x + 1 < W) { o.pack(o11); o.store(base + (size_t)W * C + C); }
      }
    }
  }
}

// TMA-staged wgrad (bf16, stride 1 or 2): per output tile the x halo tile AND the dy tile come in by TMA (out-of-range dy reads
// as zero, so ragged edges need no tests); a thread keeps the 9 x 8 fp32 sums of its channel group over all of its CTA's tiles,
// the CTA reduces its strips once at the end and writes one partial row per tap.
template <int CB, int ST>
struct DwWgTile {
  static constexpr int GROUPS = CB / 8, NSTRIP = kThreads / GROUPS, TW = 16, TH = NSTRIP / 4;
  static constexpr int IW = (TW - 1) * ST + 3, IH = (TH - 1) * ST + 3;
  static constexpr int X_BYTES = ((IH * IW * CB * 2 + 127) / 128) * 128, DY_BYTES = TH * TW * CB * 2;
  static constexpr int STAGE_BYTES = X_BYTES + DY_BYTES;
};

template <int CB, int ST, int STAGES>
__global__ void __launch_bounds__(kThreads, 1)
dwconv_tma_wgrad_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDY, float* __restrict__ partial,
                        int N, int tiles_h, int tiles_w, int C) {
  using TL = DwWgTile<CB, ST>;
  constexpr int V = 8, H2 = 4, NCOL = 3 * ST + 3;   // input columns a strip of 4 outputs touches: 6 (stride 1) or 9 (stride 2)
  extern __shared__ __align__(128) uint8_t dw_smem[];
  uint8_t* tiles = dw_smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(dw_smem + STAGES * TL::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  const int c0 = blockIdx.y * CB;
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { dw_mbar_init(&full_bar[i], 1); dw_mbar_init(&empty_bar[i], kThreads / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int per_img = tiles_h * tiles_w;
  const int ntiles = N * per_img;
  const int my_tiles = blockIdx.x < ntiles ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  auto issue = [&](int it) {
    const int tile = blockIdx.x + it * gridDim.x;
    const int n = tile / per_img, rem = tile - n * per_img;
    const int th = rem / tiles_w, tw = rem - th * tiles_w;
    const int s = it % STAGES;
    uint8_t* dst = tiles + s * TL::STAGE_BYTES;
    dw_mbar_expect_tx(&full_bar[s], TL::IH * TL::IW * CB * 2 + TL::DY_BYTES);
    dw_tma_load_4d(dst, &tmX, &full_bar[s], c0, tw * TL::TW * ST - 1, th * TL::TH * ST - 1, n);
    dw_tma_load_4d(dst + TL::X_BYTES, &tmDY, &full_bar[s], c0, tw * TL::TW, th * TL::TH, n);
  };
  if (threadIdx.x == 0)
    for (int it = 0; it < STAGES && it < my_tiles; ++it) issue(it);
  const int tcg = threadIdx.x % TL::GROUPS, strip = threadIdx.x / TL::GROUPS;
  const int row = strip >> 2, col0 = (strip & 3) * 4;
  float2 acc[9][H2];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int e = 0; e < H2; ++e) acc[t][e] = make_float2(0.f, 0.f);
  for (int it = 0; it < my_tiles; ++it) {
    const int s = it % STAGES;
    const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
    dw_mbar_wait(&full_bar[s], ph);
    const uint8_t* tx = tiles + s * TL::STAGE_BYTES;
    const uint8_t* tdy = tx + TL::X_BYTES;
    float2 gv[4][H2];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      Vec16<bf16> raw;
      raw.raw = *reinterpret_cast<const uint4*>(tdy + ((row * TL::TW + col0 + j) * CB + tcg * V) * 2);
      unpack2(raw, gv[j]);
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      float2 xv[NCOL][H2];
#pragma unroll
      for (int k = 0; k < NCOL; ++k) {
        Vec16<bf16> raw;
        raw.raw = *reinterpret_cast<const uint4*>(tx + (((row * ST + r) * TL::IW + col0 * ST + k) * CB + tcg * V) * 2);
        unpack2(raw, xv[k]);
      }
#pragma unroll
      for (int sx = 0; sx < 3; ++sx)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int e = 0; e < H2; ++e) acc[r * 3 + sx][e] = __ffma2_rn(xv[j * ST + sx][e], gv[j][e], acc[r * 3 + sx][e]);
    }
    __syncwarp();
    if ((threadIdx.x & 31) == 0) dw_mbar_arrive(&empty_bar[s]);
    if (threadIdx.x == 0 && it + STAGES < my_tiles) {
      dw_mbar_wait(&empty_bar[s], ph);
      issue(it + STAGES);
    }
  }
  // reduce the CTA's strips (fixed order), one tap at a time, through the (now idle) tile ring
  __syncthreads();
  float* red = reinterpret_cast<float*>(dw_smem);
#pragma unroll
  for (int t = 0; t < 9; ++t) {
#pragma unroll
    for (int e = 0; e < H2; ++e) { red[threadIdx.x * V + 2 * e] = acc[t][e].x; red[threadIdx.x * V + 2 * e + 1] = acc[t][e].y; }
    __syncthreads();
    if (strip == 0 && c0 + tcg * V < C) {
      float sum[V];
#pragma unroll
      for (int e = 0; e < V; ++e) sum[e] = 0.f;
      for (int l = 0; l < TL::NSTRIP; ++l)
#pragma unroll
        for (int e = 0; e < V; ++e) sum[e] += red[(l * TL::GROUPS + tcg) * V + e];
      float* dst = partial + ((size_t)blockIdx.x * 9 + t) * C + c0 + tcg * V;
#pragma unroll
      for (int e = 0; e < V; ++e) dst[e] = sum[e];
    }
    __syncthreads();
  }
}

// stride-2 dgrad: dx[2i+a, 2j+b] for a,b in {0,1} from dy[i..i+1][j..j+1]
//   (2i  ,2j  ) = dy[i,j]*w11
//   (2i  ,2j+1) = dy[i,j]*w12 + dy[i,j+1]*w10
//   (2i+1,2j  ) = dy[i,j]*w21 + dy[i+1,j]*w01
//   (2i+1,2j+1) = dy[i,j]*w22 + dy[i,j+1]*w20 + dy[i+1,j]*w02 + dy[i+1,j+1]*w00
template <typename T>
__global__ void __launch_bounds__(kThreads)
dwconv_dgrad_s2_kernel(const T* __restrict__ dy, const float* __restrict__ w, T* __restrict__ dx, int N, int H, int W, int P, int Q,
                       int C, int cgb, int lanes, int cells_per_blk) {
  constexpr int V = Vec16<T>::N;
  __shared__ __align__(16) float ws[9][128 * V];
  const int cg0 = blockIdx.y * cgb;
  load_taps<V>(ws, w, cg0 * V, min(cgb * V, C - cg0 * V), false);
  __syncthreads();
  const int tcg = threadIdx.x % cgb, lane = threadIdx.x / cgb;
  if (lane >= lanes || (cg0 + tcg) * V >= C) return;
  const int c = (cg0 + tcg) * V;
  const int ch = (H + 1) / 2, cw = (W + 1) / 2;
  const int ncells = N * ch * cw;
  const int c0 = blockIdx.x * cells_per_blk;
  const int c1 = min(ncells, c0 + cells_per_blk);
  for (int cell = c0 + lane; cell < c1; cell += lanes) {
    const int j = cell % cw;
    const int t = cell / cw;
    const int i = t % ch, n = t / ch;
    float g[2][2][V];
    Vec16<T> raw[2][2];
    bool ok[2][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        ok[a][b] = (i + a < P) && (j + b < Q);
        if (ok[a][b]) raw[a][b].load(dy + (((size_t)n * P + i + a) * Q + j + b) * C + c);
      }
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        if (ok[a][b]) raw[a][b].unpack(g[a][b]);
        else {
#pragma unroll
          for (int e = 0; e < V; ++e) g[a][b][e] = 0.f;
        }
      }
    const float* wt = &ws[0][tcg * V];
    constexpr int WS = 128 * V;  // stride between taps in ws
    float o00[V], o01[V], o10[V], o11[V];
#pragma unroll
    for (int e = 0; e < V; ++e) {
      const float w00 = wt[0 * WS + e], w01 = wt[1 * WS + e], w02 = wt[2 * WS + e];
      const float w10 = wt[3 * WS + e], w11 = wt[4 * WS + e], w12 = wt[5 * WS + e];
      const float w20 = wt[6 * WS + e], w21 = wt[7 * WS + e], w22 = wt[8 * WS + e];
      o00[e] = g[0][0][e] * w11;
      o01[e] = fmaf(g[0][0][e], w12, g[0][1][e] * w10);
      o10[e] = fmaf(g[0][0][e], w21, g[1][0][e] * w01);
      o11[e] = fmaf(g[0][0][e], w22, fmaf(g[0][1][e], w20, fmaf(g[1][0][e], w02, g[1][1][e] * w00)));
    }
    const int h = 2 * i, x = 2 * j;
    T* base = dx + (((size_t)n * H + h) * W + x) * C + c;
    Vec16<T> o;
    o.pack(o00); o.store(base);
    if (x + 1 < W) { o.pack(o01); o.store(base + C); }
    if (h + 1 < H) {
      o.pack(o10); o.store(base + (size_t)W * C);
      if (x + 1 < W) { o.pack(o11); o.store(base + (size_t)W * C + C); }
    }
  }
}

template <typename T, int ST>
__global__ void __launch_bounds__(kThreads, 2)
dwconv_wgrad_kernel(const T* __restrict__ x, const T* __restrict__ dy, float* __restrict__ partial, int N, int H, int W, int P, int Q,
                    int C, int cgb, int lanes, int strips_per_blk) {
  constexpr int V = Vec16<T>::N;
  constexpr int TW = StripW<ST>::wgrad;
  constexpr int NCOL = (TW - 1) * ST + 3;
  __shared__ float red[kThreads * V];
  const int cg0 = blockIdx.y * cgb;
  const int tcg = threadIdx.x % cgb, lane = threadIdx.x / cgb;
  const bool active = lane < lanes && (cg0 + tcg) * V < C;
  const int c = (cg0 + tcg) * V;
  float acc[9][V];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int e = 0; e < V; ++e) acc[t][e] = 0.f;
  const int strips_w = (Q + TW - 1) / TW;
  const int nstrips = N * P * strips_w;
  const int s0 = blockIdx.x * strips_per_blk;
  const int s1 = min(nstrips, s0 + strips_per_blk);
  if (active) {
    for (int sidx = s0 + lane; sidx < s1; sidx += lanes) {
      const int sw = sidx % strips_w;
      const int t = sidx / strips_w;
      const int p = t % P, n = t / P;
      const int q0 = sw * TW;
      float g[TW][V];
      const T* gy = dy + (((size_t)n * P + p) * Q + q0) * C + c;
#pragma unroll
      for (int j = 0; j < TW; ++j) {
        if (q0 + j < Q) { Vec16<T> v; v.load(gy + (size_t)j * C); v.unpack(g[j]); }
        else {
#pragma unroll
          for (int e = 0; e < V; ++e) g[j][e] = 0.f;
        }
      }
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int h = p * ST + r - 1;
        if ((unsigned)h >= (unsigned)H) continue;
        const T* row = x + ((size_t)n * H + h) * W * C + c;
        Vec16<T> raw[NCOL];
#pragma unroll
        for (int k = 0; k < NCOL; ++k) {
          const int iw = q0 * ST - 1 + k;
          if ((unsigned)iw < (unsigned)W) raw[k].load(row + (size_t)iw * C);
        }
#pragma unroll
        for (int k = 0; k < NCOL; ++k) {
          const int iw = q0 * ST - 1 + k;
          if ((unsigned)iw >= (unsigned)W) continue;
          float xv[V];
          raw[k].unpack(xv);
#pragma unroll
          for (int j = 0; j < TW; ++j) {
            const int s = k - j * ST;
            if (s < 0 || s > 2) continue;
#pragma unroll
            for (int e = 0; e < V; ++e) acc[r * 3 + s][e] = fmaf(xv[e], g[j][e], acc[r * 3 + s][e]);
          }
        }
      }
    }
  }
  // reduce the strip lanes of this block, one tap at a time
#pragma unroll
  for (int t = 0; t < 9; ++t) {
#pragma unroll
    for (int e = 0; e < V; ++e) red[threadIdx.x * V + e] = acc[t][e];
    __syncthreads();
    if (lane == 0 && active) {
      float s[V];
#pragma unroll
      for (int e = 0; e < V; ++e) s[e] = 0.f;
      for (int l = 0; l < lanes; ++l)
#pragma unroll
        for (int e = 0; e < V; ++e) s[e] += red[(l * cgb + tcg) * V + e];
      float* dst = partial + ((size_t)blockIdx.x * 9 + t) * C + c;
#pragma unroll
      for (int e = 0; e < V; ++e) dst[e] = s[e];
    }
    __syncthreads();
  }
}

// dw[c][tap] = sum_blk partial[blk][tap][c].  32 consecutive (tap, c) columns x 8 row lanes per CTA: coalesced 128-byte reads,
// nblk/8 additions per thread, fixed order (ncu on the one-thread-per-column version: 51 us for a 2.7 MB read, 5 CTAs).
__global__ void __launch_bounds__(256) dwconv_wgrad_finalize_kernel(const float* __restrict__ partial, float* __restrict__ dw, int nblk, int C) {
  __shared__ double red[8][32];
  const int col = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + col;  // over 9*C, c fastest
  double s = 0.0;
  if (i < 9 * C)
    for (int b = rl; b < nblk; b += 8) s += (double)partial[(size_t)b * 9 * C + i];
  red[rl][col] = s;
  __syncthreads();
  if (rl == 0 && i < 9 * C) {
#pragma unroll
    for (int k = 1; k < 8; ++k) s += red[k][col];
    const int t = i / C, c = i % C;
    dw[(size_t)c * 9 + t] = (float)s;
  }
}

static DwDims dims_of(const dv_dwconv_params* p) {
  DwDims d;
  d.N = p->N; d.H = p->H; d.W = p->W; d.C = p->C; d.stride = p->stride;
  d.P = (p->H + 2 - 3) / p->stride + 1;
  d.Q = (p->W + 2 - 3) / p->stride + 1;
  return d;
}

// blocks over `units` work items (strips / cells) with `lanes` lanes per block, capped at `cap_per_sm` blocks per SM
static void split_units(int64_t units, int lanes, int per_lane, int cap_per_sm, int* nblk, int* per_blk) {
  int64_t nb = ceil_div64(units, (int64_t)lanes * per_lane);
  const int64_t cap = (int64_t)num_sms() * cap_per_sm;
  if (nb > cap) nb = cap;
  if (nb < 1) nb = 1;
  const int64_t pb = ceil_div64(units, nb);
  *per_blk = (int)pb;
  *nblk = (int)ceil_div64(units, pb);
}

// one wave: two CTAs per SM are resident (128 registers x 256 threads); caps of 4, 6, 8 measured 5-35 % slower
static int env_cap() {
  static const int v = getenv("DV_DW_WGRAD_CAP") ? atoi(getenv("DV_DW_WGRAD_CAP")) : 2;
  return v > 0 ? v : 2;
}

static int wgrad_blocks(const DwDims& d, int lanes, int* per_blk) {
  const int64_t strips = (int64_t)d.N * d.P * ceil_div(d.Q, 2);  // StripW::wgrad
  int nb;
  split_units(strips, lanes, 8, env_cap(), &nb, per_blk);  // CTAs per SM the grid is capped at (2 are resident)
  return nb;
}

// TMA-staged forward-shaped kernel: `in` [N][H][W][C] -> `out` [N][OH][OW][C], bf16 (fprop: x -> y, stride 1 or 2; stride-1 dgrad:
// dy -> dx with flipped taps)
template <int CB, int ST>
static int launch_dw_tma_fwd(const void* in, const float* w, void* out, int N, int H, int W, int OH, int OW, int C, int flip,
                             const float* ep_shift, int ep_relu, cudaStream_t st, float* stat_sums = nullptr,
                             const dv_dwconv_params* bnb = nullptr) {
  constexpr int RW = ST == 1 ? DV_DW_RW : 1;
  using TL = DwTile<CB, ST, RW>;
  constexpr int STAGES = (ST == 1 && RW == 1) ? 3 : (ST == 1 ? 3 : 2);
  alignas(64) CUtensorMap tm;
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  const uint64_t strides[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
  const uint32_t box[4] = {(uint32_t)CB, (uint32_t)TL::IW, (uint32_t)TL::IH, 1};
  int rc = make_tiled_map_bf16(&tm, in, 4, dims, strides, box, 0);
  if (rc) return rc;
  DwTmaArgs g{};
  g.N = N; g.OH = OH; g.OW = OW; g.C = C;
  g.tiles_h = ceil_div(OH, TL::TH); g.tiles_w = ceil_div(OW, TL::TW);
  g.flip = flip; g.ep_relu = ep_relu; g.ep_shift = ep_shift; g.stat_sums = stat_sums;
  if (bnb != nullptr) {
    g.bnb_x = (const bf16*)bnb->bnb_x; g.bnb_gamma = bnb->bnb_gamma; g.bnb_beta = bnb->bnb_beta; g.bnb_mean = bnb->bnb_mean;
    g.bnb_invstd = bnb->bnb_invstd; g.bnb_dgamma = bnb->bnb_dgamma; g.bnb_dbeta = bnb->bnb_dbeta;
  }
  constexpr int smem = STAGES * TL::STAGE_BYTES + 9 * CB * 4 + 2 * STAGES * 8;
  static_assert(smem <= 227 * 1024, "depth-wise tile ring exceeds shared memory");
  DV_CUDA_OK(cudaFuncSetAttribute(dwconv_tma_fwd_kernel<CB, ST, RW, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int cblocks = C / CB;
  const int64_t ntiles = (int64_t)N * g.tiles_h * g.tiles_w;
  DV_REQUIRE(ntiles < (1LL << 31), DV_ERR_UNSUPPORTED, "dwconv: tensor too large");
  int gx = ((ST == 1 && RW == 1) ? 2 : 1) * num_sms() / cblocks;
  if (gx < 1) gx = 1;
  if (gx > ntiles) gx = (int)ntiles;
  if constexpr (ST == 1) {
    if (bnb != nullptr) {
      DV_CUDA_OK(cudaFuncSetAttribute(dwconv_tma_fwd_kernel<CB, ST, RW, STAGES, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      dwconv_tma_fwd_kernel<CB, ST, RW, STAGES, true><<<dim3((unsigned)gx, (unsigned)cblocks), kThreads, smem, st>>>(tm, w, (bf16*)out, g);
      DV_LAUNCH_OK("dwconv_tma_fwd_kernel<bnb>");
      return DV_OK;
    }
  }
  dwconv_tma_fwd_kernel<CB, ST, RW, STAGES><<<dim3((unsigned)gx, (unsigned)cblocks), kThreads, smem, st>>>(tm, w, (bf16*)out, g);
  DV_LAUNCH_OK("dwconv_tma_fwd_kernel");
  return DV_OK;
}
static bool dw_tma_ok(const dv_dwconv_params* p) {
  static const int enabled = getenv("DV_DW_TMA") ? atoi(getenv("DV_DW_TMA")) : 1;
  return enabled && p->dtype == DV_BF16 && p->C % 32 == 0 && p->W >= 14 && p->H >= 8 &&
         ((reinterpret_cast<uintptr_t>(p->x) | reinterpret_cast<uintptr_t>(p->y)) & 15) == 0;
}

template <typename T>
static int run_fprop(const dv_dwconv_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  DV_REQUIRE(p->C % V == 0, DV_ERR_UNSUPPORTED, "dwconv: C=%d must be a multiple of %d", p->C, V);
  DwDims d = dims_of(p);
  if (sizeof(T) == 2 && dw_tma_ok(p) && d.Q >= 8) {
    const float* w = (const float*)p->w;
    if (d.stride == 1)
      return p->C % 64 == 0 ? launch_dw_tma_fwd<64, 1>(p->x, w, p->y, d.N, d.H, d.W, d.P, d.Q, d.C, 0, p->ep_shift, p->ep_relu, st, p->stat_sums)
                            : launch_dw_tma_fwd<32, 1>(p->x, w, p->y, d.N, d.H, d.W, d.P, d.Q, d.C, 0, p->ep_shift, p->ep_relu, st, p->stat_sums);
    return p->C % 64 == 0 ? launch_dw_tma_fwd<64, 2>(p->x, w, p->y, d.N, d.H, d.W, d.P, d.Q, d.C, 0, p->ep_shift, p->ep_relu, st, p->stat_sums)
                          : launch_dw_tma_fwd<32, 2>(p->x, w, p->y, d.N, d.H, d.W, d.P, d.Q, d.C, 0, p->ep_shift, p->ep_relu, st, p->stat_sums);
  }
  DV_REQUIRE(p->stat_sums == nullptr, DV_ERR_UNSUPPORTED, "dwconv fprop: stat_sums is honoured by the TMA-staged kernels only (ask dv_dwconv_fuses_stats)");
  DwPlan pl = dw_plan(p->C, V);
  const int64_t strips = (int64_t)d.N * d.P * ceil_div(d.Q, d.stride == 1 ? StripW<1>::fwd : StripW<2>::fwd);
  DV_REQUIRE(strips < (1LL << 31), DV_ERR_UNSUPPORTED, "dwconv: tensor too large");
  int nb, per;
  // one wave of resident CTAs (2 per SM for stride 1, 3 for stride 2): every CTA loads the tap table once and streams
  static const int fwd_cap = getenv("DV_DW_FWD_CAP") ? atoi(getenv("DV_DW_FWD_CAP")) : 0;
  split_units(strips, pl.lanes, 4, fwd_cap > 0 ? fwd_cap : (d.stride == 1 ? 2 : 3), &nb, &per);
  dim3 grid((unsigned)nb, (unsigned)pl.cchunks);
  if (d.stride == 1)
    dwconv_strip_kernel<T, 1><<<grid, kThreads, 0, st>>>((const T*)p->x, (const float*)p->w, (T*)p->y, d.N, d.H, d.W, d.P, d.Q, d.C, pl.cgb,
                                                       pl.lanes, per, 0, p->ep_shift, p->ep_relu);
  else
    dwconv_strip_kernel<T, 2><<<grid, kThreads, 0, st>>>((const T*)p->x, (const float*)p->w, (T*)p->y, d.N, d.H, d.W, d.P, d.Q, d.C, pl.cgb,
                                                       pl.lanes, per, 0, p->ep_shift, p->ep_relu);
  DV_LAUNCH_OK("dwconv_strip_kernel");
  return DV_OK;
}

template <int CB>
static int launch_dw_tma_dgrad_s2(const dv_dwconv_params* p, const DwDims& d, cudaStream_t st) {
  using TL = DwS2Tile<CB>;
  constexpr int STAGES = 4;
  alignas(64) CUtensorMap tm;
  const uint64_t dims[4] = {(uint64_t)d.C, (uint64_t)d.Q, (uint64_t)d.P, (uint64_t)d.N};
  const uint64_t strides[3] = {(uint64_t)d.C * 2, (uint64_t)d.Q * d.C * 2, (uint64_t)d.P * d.Q * d.C * 2};
  const uint32_t box[4] = {(uint32_t)CB, (uint32_t)TL::IW, (uint32_t)TL::IH, 1};
  int rc = make_tiled_map_bf16(&tm, p->y, 4, dims, strides, box, 0);
  if (rc) return rc;
  const int tiles_h = ceil_div((d.H + 1) / 2, TL::CH), tiles_w = ceil_div((d.W + 1) / 2, TL::CW);
  const int64_t ntiles = (int64_t)d.N * tiles_h * tiles_w;
  DV_REQUIRE(ntiles < (1LL << 31), DV_ERR_UNSUPPORTED, "dwconv: tensor too large");
  constexpr int smem = STAGES * TL::STAGE_BYTES + 9 * CB * 4 + 2 * STAGES * 8;
  DV_CUDA_OK(cudaFuncSetAttribute(dwconv_tma_dgrad_s2_kernel<CB, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int cblocks = d.C / CB;
  int gx = 2 * num_sms() / cblocks;
  if (gx < 1) gx = 1;
  if (gx > ntiles) gx = (int)ntiles;
  dwconv_tma_dgrad_s2_kernel<CB, STAGES><<<dim3((unsigned)gx, (unsigned)cblocks), kThreads, smem, st>>>(tm, (const float*)p->w, (bf16*)p->x, d.N,
                                                                                                    d.H, d.W, d.C, tiles_h, tiles_w);
  DV_LAUNCH_OK("dwconv_tma_dgrad_s2_kernel");
  return DV_OK;
}

template <typename T>
static int run_dgrad(const dv_dwconv_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  DV_REQUIRE(p->C % V == 0, DV_ERR_UNSUPPORTED, "dwconv: C=%d must be a multiple of %d", p->C, V);
  DwDims d = dims_of(p);
  if (sizeof(T) == 2 && d.stride == 1 && dw_tma_ok(p)) {
    const dv_dwconv_params* bnb = p->bnb_x != nullptr ? p : nullptr;
    if (bnb != nullptr)
      DV_REQUIRE(p->bnb_gamma && p->bnb_beta && p->bnb_mean && p->bnb_invstd && p->bnb_dgamma && p->bnb_dbeta && ((uintptr_t)p->bnb_x & 15) == 0,
                 DV_ERR_ARG, "dwconv dgrad: bnb_x given but one of bnb_gamma/beta/mean/invstd/dgamma/dbeta is null (or bnb_x is not 16-byte aligned)");
    return p->C % 64 == 0 ? launch_dw_tma_fwd<64, 1>(p->y, (const float*)p->w, p->x, d.N, d.H, d.W, d.H, d.W, d.C, 1, nullptr, 0, st, nullptr, bnb)
                          : launch_dw_tma_fwd<32, 1>(p->y, (const float*)p->w, p->x, d.N, d.H, d.W, d.H, d.W, d.C, 1, nullptr, 0, st, nullptr, bnb);
  }
  DV_REQUIRE(p->bnb_x == nullptr, DV_ERR_UNSUPPORTED,
             "dwconv dgrad: bnb_* (BatchNorm-backward sums) is honoured by the TMA-staged stride-1 kernel only (ask dv_dwconv_fuses_bn_bwd)");
  DwPlan pl = dw_plan(p->C, V);
  int nb, per;
  if (d.stride == 1) {
    // dx = "fprop" of dy with the taps flipped (pad 1 both ways)
    const int64_t strips = (int64_t)d.N * d.H * ceil_div(d.W, StripW<1>::fwd);
    DV_REQUIRE(strips < (1LL << 31), DV_ERR_UNSUPPORTED, "dwconv: tensor too large");
    static const int dg_cap = getenv("DV_DW_FWD_CAP") ? atoi(getenv("DV_DW_FWD_CAP")) : 0;
    split_units(strips, pl.lanes, 4, dg_cap > 0 ? dg_cap : 2, &nb, &per);
    dim3 grid((unsigned)nb, (unsigned)pl.cchunks);
    dwconv_strip_kernel<T, 1><<<grid, kThreads, 0, st>>>((const T*)p->y, (const float*)p->w, (T*)p->x, d.N, d.P, d.Q, d.H, d.W, d.C, pl.cgb,
                                                       pl.lanes, per, 1);
  } else if (sizeof(T) == 2 && dw_tma_ok(p) && d.Q >= 8) {
    return p->C % 64 == 0 ? launch_dw_tma_dgrad_s2<64>(p, d, st) : launch_dw_tma_dgrad_s2<32>(p, d, st);
  } else {
    const int64_t cells = (int64_t)d.N * ((d.H + 1) / 2) * ((d.W + 1) / 2);
    DV_REQUIRE(cells < (1LL << 31), DV_ERR_UNSUPPORTED, "dwconv: tensor too large");
    static const int s2_cap = getenv("DV_DW_S2_CAP") ? atoi(getenv("DV_DW_S2_CAP")) : 16;
    split_units(cells, pl.lanes, 8, s2_cap, &nb, &per);
    dim3 grid((unsigned)nb, (unsigned)pl.cchunks);
    dwconv_dgrad_s2_kernel<T><<<grid, kThreads, 0, st>>>((const T*)p->y, (const float*)p->w, (T*)p->x, d.N, d.H, d.W, d.P, d.Q, d.C, pl.cgb,
                                                       pl.lanes, per);
  }
  DV_LAUNCH_OK("dwconv dgrad");
  return DV_OK;
}

static int dw_tma_wgrad_blocks(const DwDims& d, int CB, int TH) {   // CTAs along x (one per SM over all channel blocks)
  const int64_t ntiles = (int64_t)d.N * ceil_div(d.P, TH) * ceil_div(d.Q, 16);
  int gx = num_sms() / (d.C / CB);
  if (gx < 1) gx = 1;
  if (gx > ntiles) gx = (int)ntiles;
  return gx;
}
void dwconv_wgrad_finalize_launch(const float* partial, float* dw, int nblk, int C, cudaStream_t st);

template <int CB, int ST>
static int launch_dw_tma_wgrad(const dv_dwconv_params* p, const DwDims& d, cudaStream_t st) {
  using TL = DwWgTile<CB, ST>;
  constexpr int STAGES = ST == 1 ? 3 : 2;
  const int gx = dw_tma_wgrad_blocks(d, CB, TL::TH);
  const size_t need = (size_t)gx * 9 * d.C * sizeof(float);
  DV_REQUIRE(p->workspace && p->workspace_bytes >= need, DV_ERR_WORKSPACE, "dwconv_wgrad: workspace %zu bytes required, %zu given", need,
             p->workspace_bytes);
  alignas(64) CUtensorMap tmX, tmDY;
  {
    const uint64_t dims[4] = {(uint64_t)d.C, (uint64_t)d.W, (uint64_t)d.H, (uint64_t)d.N};
    const uint64_t strides[3] = {(uint64_t)d.C * 2, (uint64_t)d.W * d.C * 2, (uint64_t)d.H * d.W * d.C * 2};
    const uint32_t box[4] = {(uint32_t)CB, (uint32_t)TL::IW, (uint32_t)TL::IH, 1};
    int rc = make_tiled_map_bf16(&tmX, p->x, 4, dims, strides, box, 0);
    if (rc) return rc;
  }
  {
    const uint64_t dims[4] = {(uint64_t)d.C, (uint64_t)d.Q, (uint64_t)d.P, (uint64_t)d.N};
    const uint64_t strides[3] = {(uint64_t)d.C * 2, (uint64_t)d.Q * d.C * 2, (uint64_t)d.P * d.Q * d.C * 2};
    const uint32_t box[4] = {(uint32_t)CB, (uint32_t)TL::TW, (uint32_t)TL::TH, 1};
    int rc = make_tiled_map_bf16(&tmDY, p->y, 4, dims, strides, box, 0);
    if (rc) return rc;
  }
  constexpr int smem = STAGES * TL::STAGE_BYTES + 2 * STAGES * 8 + 128;
  static_assert(smem <= 227 * 1024, "depth-wise wgrad ring exceeds shared memory");
  DV_CUDA_OK(cudaFuncSetAttribute(dwconv_tma_wgrad_kernel<CB, ST, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  dwconv_tma_wgrad_kernel<CB, ST, STAGES><<<dim3((unsigned)gx, (unsigned)(d.C / CB)), kThreads, smem, st>>>(
      tmX, tmDY, (float*)p->workspace, d.N, ceil_div(d.P, TL::TH), ceil_div(d.Q, TL::TW), d.C);
  DV_LAUNCH_OK("dwconv_tma_wgrad_kernel");
  dwconv_wgrad_finalize_launch((const float*)p->workspace, (float*)p->w, gx, d.C, st);
  DV_LAUNCH_OK("dwconv_wgrad_finalize_kernel");
  return DV_OK;
}

template <typename T>
static int run_wgrad(const dv_dwconv_params* p, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  DV_REQUIRE(p->C % V == 0, DV_ERR_UNSUPPORTED, "dwconv: C=%d must be a multiple of %d", p->C, V);
  DwDims d = dims_of(p);
  if (sizeof(T) == 2 && dw_tma_ok(p) && d.Q >= 8) {
    if (d.stride == 1) return p->C % 64 == 0 ? launch_dw_tma_wgrad<64, 1>(p, d, st) : launch_dw_tma_wgrad<32, 1>(p, d, st);
    return p->C % 64 == 0 ? launch_dw_tma_wgrad<64, 2>(p, d, st) : launch_dw_tma_wgrad<32, 2>(p, d, st);
  }
  DwPlan pl = dw_plan(p->C, V);
  int per;
  const int nb = wgrad_blocks(d, pl.lanes, &per);
  const size_t need = (size_t)nb * 9 * p->C * sizeof(float);
  DV_REQUIRE(p->workspace && p->workspace_bytes >= need, DV_ERR_WORKSPACE, "dwconv_wgrad: workspace %zu bytes required, %zu given", need,
             p->workspace_bytes);
  dim3 grid((unsigned)nb, (unsigned)pl.cchunks);
  if (d.stride == 1)
    dwconv_wgrad_kernel<T, 1><<<grid, kThreads, 0, st>>>((const T*)p->x, (const T*)p->y, (float*)p->workspace, d.N, d.H, d.W, d.P, d.Q, d.C,
                                                       pl.cgb, pl.lanes, per);
  else
    dwconv_wgrad_kernel<T, 2><<<grid, kThreads, 0, st>>>((const T*)p->x, (const T*)p->y, (float*)p->workspace, d.N, d.H, d.W, d.P, d.Q, d.C,
                                                       pl.cgb, pl.lanes, per);
  DV_LAUNCH_OK("dwconv_wgrad_kernel");
  dwconv_wgrad_finalize_kernel<<<ceil_div(9 * p->C, 32), 256, 0, st>>>((const float*)p->workspace, (float*)p->w, nb, p->C);
  DV_LAUNCH_OK("dwconv_wgrad_finalize_kernel");
  return DV_OK;
}

void dwconv_wgrad_finalize_launch(const float* partial, float* dw, int nblk, int C, cudaStream_t st) {
  dwconv_wgrad_finalize_kernel<<<ceil_div(9 * C, 32), 256, 0, st>>>(partial, dw, nblk, C);
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
  if (!p || p->C <= 0 || (p->stride != 1 && p->stride != 2)) return 0;
  DwDims d = dims_of(p);
  size_t best = 0;
  for (int V : {4, 8}) {  // sized for whichever dtype's plan needs more blocks
    if (p->C % V) continue;
    DwPlan pl = dw_plan(p->C, V);
    int per;
    const size_t nb = (size_t)wgrad_blocks(d, pl.lanes, &per);
    const size_t need = nb * 9 * p->C * sizeof(float);
    if (need > best) best = need;
  }
  if (p->C % 32 == 0) {   // the TMA-staged kernels: one partial row per CTA along x
    const int CB = p->C % 64 == 0 ? 64 : 32;
    const size_t need = (size_t)dw_tma_wgrad_blocks(d, CB, 256 / (CB / 8) / 4) * 9 * p->C * sizeof(float);
    if (need > best) best = need;
  }
  return best;
}
extern "C" int dv_dwconv_fuses_stats(const dv_dwconv_params* p) {
  if (!p || (p->stride != 1 && p->stride != 2) || p->dtype != DV_BF16 || !dw_tma_ok(p)) return 0;
  return dims_of(p).Q >= 8 ? 1 : 0;
}
extern "C" int dv_dwconv_fuses_bn_bwd(const dv_dwconv_params* p) {
  if (!p || p->stride != 1 || p->dtype != DV_BF16 || !dw_tma_ok(p)) return 0;
  return 1;
}
extern "C" int dv_dwconv_fprop(const dv_dwconv_params* p, void* stream) {
  int rc = check(p, "dv_dwconv_fprop");
  if (rc) return rc;
  DV_REQUIRE(p->bnb_x == nullptr, DV_ERR_UNSUPPORTED, "dv_dwconv_fprop: bnb_* belongs to dgrad");
  return p->dtype == DV_BF16 ? run_fprop<bf16>(p, (cudaStream_t)stream) : run_fprop<float>(p, (cudaStream_t)stream);
}
extern "C" int dv_dwconv_dgrad(const dv_dwconv_params* p, void* stream) {
  int rc = check(p, "dv_dwconv_dgrad");
  if (rc) return rc;
  DV_REQUIRE(p->ep_shift == nullptr && p->ep_relu == 0, DV_ERR_UNSUPPORTED, "dv_dwconv_dgrad: the folded epilogue is fprop-only");
  return p->dtype == DV_BF16 ? run_dgrad<bf16>(p, (cudaStream_t)stream) : run_dgrad<float>(p, (cudaStream_t)stream);
}
extern "C" int dv_dwconv_wgrad(const dv_dwconv_params* p, void* stream) {
  int rc = check(p, "dv_dwconv_wgrad");
  if (rc) return rc;
  DV_REQUIRE(p->ep_shift == nullptr && p->ep_relu == 0, DV_ERR_UNSUPPORTED, "dv_dwconv_wgrad: the folded epilogue is fprop-only");
  DV_REQUIRE(p->bnb_x == nullptr, DV_ERR_UNSUPPORTED, "dv_dwconv_wgrad: bnb_* belongs to dgrad");
  return p->dtype == DV_BF16 ? run_wgrad<bf16>(p, (cudaStream_t)stream) : run_wgrad<float>(p, (cudaStream_t)stream);
}
