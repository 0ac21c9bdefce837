// conv_tc.cu — dense convolution on the Blackwell tensor cores: NHWC implicit GEMM, bf16 operands staged into
// 128B-swizzled shared memory by TMA (tiled 2-D maps for filters and 1x1 activations, im2col maps for 3x3 /
// strided activations), tcgen05.mma with the fp32 accumulator in TMEM, tcgen05.ld epilogue.
//
//   gemm kernels (fprop; dgrad as "fprop of dy" with the KRSC filters read MN-major and the taps flipped by index; 1x1 stride-s
//   dgrad = GEMM + strided row scatter; 3x3 stride-2 dgrad = four parity-class stride-1 sub-convolutions):
//       D[128 pixels x BN channels] += A[128 px x 64 ch] (K-major) * B[BN x 64 ch], K = taps * C/64 blocks
//       persistent, accumulator double-buffered in TMEM; a CTA-pair variant (cta_group::2) takes the 256-wide layers.
//       Epilogue: TMEM -> bf16 -> swizzled shared tile -> TMA store / TMA reduce-add, optional BatchNorm sums, optional folded
//       BatchNorm (scale, shift, identity by TMA, ReLU) for inference.
//   wgrad kernel:
//       D[128 k x BNC c] += dY[64 px x 128 k]^T (MN-major) * Xtap[64 px x BNC c] (MN-major), reduction over pixels split
//       across one wave of CTAs; tiles added into dW by TMA reduce-add (fp32, staged in the idle operand ring), or written as
//       partials and reduced in a fixed order (deterministic mode).
//
// Warp roles per CTA (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer (both run warp-uniform loops and
// issue from the lane elect.sync picks), warps 2..5 = epilogue (each owns the 32 TMEM lanes of its warp_id%4 quarter).  smem
// ring of STAGES stages, full/empty mbarriers, tcgen05.commit releases a stage and finally signals the epilogue.
#include "common.cuh"
#include <cuda.h>  // CUtensorMap + enums only; the two encode entry points are resolved at run time (no -lcuda)
#include <mutex>
#include <stdlib.h>
#include <string.h>
#include <unordered_map>

namespace dv {
namespace {

// ------------------------------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bounded wait: a protocol bug must trap, not hang the GPU (a hung box is a lost box).
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done = 0;
  long long t0 = 0;
  for (uint32_t spin = 0;; ++spin) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (done) return;
    if ((spin & 0x3ff) == 0x3ff) {
      const long long now = clock64();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 4000000000LL) {  // ~2 s at 2 GHz
        printf("dv conv_tc: mbarrier wait timed out (block %d,%d thread %d parity %u)\n", blockIdx.x, blockIdx.y,
               threadIdx.x, parity);
        __trap();
      }
    }
  }
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ void tma_load_2d(void* smem, const CUtensorMap* tm, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_im2col(void* smem, const CUtensorMap* tm, uint64_t* bar, int c, int w, int h,
                                                int n, uint16_t off_w, uint16_t off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(smem_u32(smem)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
      : "memory");
}
// smem tile -> global through the TMA unit (whole 128-byte lines, clipped at the tensor edge, no LSU traffic)
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, uint32_t smem_addr, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(reinterpret_cast<uint64_t>(tm)),
               "r"(smem_addr), "r"(c0), "r"(c1)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// the same, but the tile is ADDED to global memory (element-wise bf16 add performed at the L2)
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* tm, uint32_t smem_addr, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(tm)),
               "r"(smem_addr), "r"(c0), "r"(c1)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// fp32 tile ADDED to a 3-D tensor (wgrad: dW [K][taps][C]); elements outside the tensor are clipped
__device__ __forceinline__ void tma_reduce_add_3d(const CUtensorMap* tm, uint32_t smem_addr, int c0, int c1, int c2) {
  asm volatile("cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(tm)),
               "r"(smem_addr), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tm)) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc]
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// all previously issued MMAs of this thread arrive on `bar` when they complete (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// 32 lanes x 32 consecutive fp32 columns: thread i of the warp receives row (lane base + i)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- CTA-pair (cta_group::2) helpers -------------------------------------------------------------------------
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // clears the CTA-rank bit of a shared::cluster address: "the same offset in CTA 0"

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 remote;\n\t"
      "mapa.shared::cluster.u32 remote, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [remote];\n\t}" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}
// TMA loads issued by either CTA of a pair: data lands in the issuing CTA's shared memory, the bytes are counted on
// CTA 0's barrier (the MMA issuer waits there for both halves)
__device__ __forceinline__ void tma_load_2d_pair(void* smem, const CUtensorMap* tm, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_im2col_pair(void* smem, const CUtensorMap* tm, uint64_t* bar, int c, int w, int h, int n,
                                                     uint16_t off_w, uint16_t off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.im2col.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(smem_u32(smem)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// one MMA over both SMs of the pair: D[256 x N] (128 rows in each SM's TMEM) += A (128 rows per SM) * B (N/2 rows per SM)
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// completion of all prior MMAs arrives on the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}

// ---- descriptors ---------------------------------------------------------------------------------------------
// Shared-memory matrix descriptor (PTX ISA "tcgen05 matrix descriptor"): start>>4 [0,14), LBO>>4 [16,30),
// SBO>>4 [32,46), version=1 [46,48), layout type [61,64) with 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// The issue loop runs on ONE thread and every instruction in it is on the tensor pipe's critical path (ncu: ~95 SASS
// instructions per 64-wide K block cost ~1000 cycles against 512 cycles of MMA work).  So descriptors are not rebuilt
// per MMA: the constant high word is a literal and the low word (address >> 4 | LBO) only takes an integer add.
// Role loops are executed by the WHOLE warp with warp-uniform control flow (the warp index comes through a shuffle so the
// compiler knows it is uniform) and only the issue itself is predicated on elect.sync: then ptxas keeps descriptors in
// uniform registers and emits the UTCHMMA / UTMALDG instructions back to back.  Issuing from `if (lane == 0)` instead
// makes it wrap every such instruction in an ELECT + BRA.U.ANY loop with R2UR moves (10 SASS instructions per MMA).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ int uniform_warp_idx() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }

constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);          // SBO 1024 B, version 1, SWIZZLE_128B
__device__ __forceinline__ uint32_t desc_lo_kmajor(uint32_t saddr) { return ((saddr & 0x3FFFFu) >> 4) | (1u << 16); }
__device__ __forceinline__ uint32_t desc_lo_mnmajor(uint32_t saddr) { return ((saddr & 0x3FFFFu) >> 4) | ((8192u >> 4) << 16); }
__device__ __forceinline__ uint64_t desc_pack(uint32_t lo) {
  uint64_t d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(kDescHi));
  return d;
}

// Instruction descriptor for kind::f16: D=f32 (bit 4), A=B=bf16 (bits 7,10), majors (15,16; 1 = MN-major),
// N>>3 at [17,23), M>>4 at [24,29).
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // bf16 elements = one 128-byte swizzle row
constexpr int kABytes = kBlockM * kBlockK * 2;
constexpr int kThreads = 192;
constexpr int kEpiStageBytes = 4 * 32 * 128;  // epilogue transpose buffers

// ------------------------------------------------------------------------------------------------------------
// GEMM-shaped kernel: fprop / dgrad
// ------------------------------------------------------------------------------------------------------------
struct GemmArgs {
  bf16* out;
  int M;            // output rows (pixels enumerated n,p,q)
  int Nout;         // output channels
  int kb_per_tap;   // ceil(C_in / 64); a partial last block is zero-filled by TMA on the A side
  int Cin;          // reduction channels per tap
  int taps, S;
  int a_im2col;     // 0: A is a 2-D [M][C] map, 1: im2col 4-D map
  int P, Q, stride, pad;
  int scatter;      // 1: row (n,p,q) is written at pixel (n, p*os, q*os) of an [*, Ho, Wo, Nout] tensor
  int Ho, Wo, os;
  int oa, ob;       // scatter: row/column offset added to (p*os, q*os) — the parity class of a stride-2 dgrad
  int use_btap;     // 1: the filter tap of reduction step `tap` is btap[tap] (sub-sampled taps of a parity class)
  signed char btap[12];
  int out_f32;      // 1: `out` is float* (fp32 logits of the classifier)
  int accumulate;   // 1: out += result (bf16 read-modify-write by the owning thread)
  float* stat_sums; // fprop, optional: [2][Nout] fp32, += per-channel sum and sum of squares of the bf16-rounded output
  // fprop, optional (inference with BatchNorm folded away): out = relu(acc * ep_scale[n] + ep_shift[n] + ep_res[m][n])
  const float* ep_scale;
  const float* ep_shift;
  const bf16* ep_res;
  int ep_relu;
  // dgrad, optional: ep_scale == NULL, out = acc + (bit(ep_mask, m, n) ? ep_res[m][n] : 0) — the shortcut's share of a residual
  // block's input gradient, ep_res = dy of the block's last ReLU and ep_mask that ReLU's one-bit-per-element mask (Nout/8 bytes per row)
  const uint8_t* ep_mask;
  // dgrad, optional (EPI == 2): the output dx is the gradient w.r.t. a = relu(bn(x0)).  While a dx tile is still on chip the
  // epilogue fetches the matching x0 tile (tensor map tmR), recomputes the ReLU mask exactly as bn_apply computed it and adds
  //   dbeta[c] += sum_m dz,   dgamma[c] += invstd[c] * sum_m dz * (x0 - mean[c]),   dz = dx * (relu'(bn(x0)))
  // which is all that BatchNorm's backward reduce pass (4 bytes per element of HBM traffic) would have produced.
  const float* bnb_gamma;
  const float* bnb_beta;
  const float* bnb_mean;
  const float* bnb_invstd;
  float* bnb_dgamma;
  float* bnb_dbeta;
};

// BatchNorm-backward sums over one staged [32 rows][64 columns] tile (EPI == 2).  The lane owns columns 2*lane, 2*lane+1 of the
// group: `xw` holds its two x0 values of every row (read from the staging buffer before dy overwrote it), dy is read back from
// the staging buffer as it was rounded to bf16 — the values bn_bwd_apply will read.  Sums are returned through a/b.
__device__ __forceinline__ void bnb_tile_sums(const uint8_t* my_stage, const uint32_t* xw, int lane, float sc0, float sh0, float mu0, float sc1,
                                              float sh1, float mu1, float& a0, float& a1, float& b0, float& b1) {
  const int wch = lane >> 2, wof = (lane & 3) * 4;
  // two columns per lane = one packed fp32 pair per row: fma.rn.f32x2 / add.rn.f32x2 (IEEE, the same bits as the scalar forms)
  const float2 sc = make_float2(sc0, sc1), sh = make_float2(sh0, sh1), nmu = make_float2(-mu0, -mu1);
  float2 a = make_float2(0.f, 0.f), b = make_float2(0.f, 0.f);
#pragma unroll   // fully: xw[] must stay in registers (constant indices)
  for (int r = 0; r < 32; ++r) {
    const uint32_t word = *reinterpret_cast<const uint32_t*>(my_stage + r * 128 + ((wch ^ (r & 7)) * 16) + wof);
    const float2 d = make_float2(__uint_as_float(word << 16), __uint_as_float(word & 0xffff0000u));
    const float2 x = make_float2(__uint_as_float(xw[r] << 16), __uint_as_float(xw[r] & 0xffff0000u));
    const float2 pre = __ffma2_rn(x, sc, sh);   // the expression of bn_apply_kernel / bn_reduce_kernel<.,1,2>: fmaf(x, sc, sh) > 0
    const float2 z = make_float2(pre.x > 0.f ? d.x : 0.f, pre.y > 0.f ? d.y : 0.f);
    a = __fadd2_rn(a, z);
    b = __ffma2_rn(z, __fadd2_rn(x, nmu), b);
  }
  a0 += a.x; a1 += a.y;
  b0 += b.x; b1 += b.y;
}

// Folded-BatchNorm epilogue on the 32 accumulator columns [col0, col0+32) of the output row held by one lane.  The identity
// tile was brought into the warp's staging buffer by TMA in the layout the result is written back in ([32 rows][128 B],
// 16-byte chunks XOR-swizzled by row), so a lane finds its row's residual exactly where its result will go: `res_row` is the
// shared address of the lane's row (0 = no residual), `chunk0` the first of the four 16-byte chunks these columns occupy.
__device__ __forceinline__ void apply_folded_epilogue(uint32_t* v, const GemmArgs& g, int col0, const uint32_t* rr, int m) {
  if (col0 >= g.Nout) return;  // Nout is a multiple of 8 and col0 of 32: a partial group ends on an 8-column boundary
  const bool scaled = g.ep_scale != nullptr;
  uint32_t bits = 0xFFFFFFFFu;   // this row's mask bits of the 32 columns from col0 (col0 is a multiple of 32: one aligned word)
  if (g.ep_mask != nullptr) bits = m < g.M ? __ldg(reinterpret_cast<const uint32_t*>(g.ep_mask + (size_t)m * (g.Nout >> 3) + (col0 >> 3))) : 0u;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (col0 + j * 8 >= g.Nout) break;
    float sc[8] = {1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f};
    float sh[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (scaled) {
      const float4 sa = __ldg(reinterpret_cast<const float4*>(g.ep_scale + col0 + j * 8));
      const float4 sb = __ldg(reinterpret_cast<const float4*>(g.ep_scale + col0 + j * 8 + 4));
      const float4 ha = __ldg(reinterpret_cast<const float4*>(g.ep_shift + col0 + j * 8));
      const float4 hb = __ldg(reinterpret_cast<const float4*>(g.ep_shift + col0 + j * 8 + 4));
      sc[0] = sa.x; sc[1] = sa.y; sc[2] = sa.z; sc[3] = sa.w; sc[4] = sb.x; sc[5] = sb.y; sc[6] = sb.z; sc[7] = sb.w;
      sh[0] = ha.x; sh[1] = ha.y; sh[2] = ha.z; sh[3] = ha.w; sh[4] = hb.x; sh[5] = hb.y; sh[6] = hb.z; sh[7] = hb.w;
    }
    float r[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (rr != nullptr) {
#pragma unroll
      for (int e = 0; e < 4; ++e) { r[2 * e] = __uint_as_float(rr[j * 4 + e] << 16); r[2 * e + 1] = __uint_as_float(rr[j * 4 + e] & 0xffff0000u); }
    }
    const uint32_t b8 = bits >> (j * 8);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float acc = __uint_as_float(v[j * 8 + e]);
      float y = (scaled ? fmaf(acc, sc[e], sh[e]) : acc) + (((b8 >> e) & 1u) ? r[e] : 0.f);
      if (g.ep_relu) y = fmaxf(y, 0.f);
      v[j * 8 + e] = __float_as_uint(y);
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// GEMM-shaped kernel (fprop / dgrad), persistent: one CTA per SM slot walks output tiles (n fastest, so CTAs running side by side share the same
// activation tile through L2); the smem ring runs on across tiles and the accumulator is DOUBLE-BUFFERED in TMEM
// (2 x BN columns), so the epilogue of tile i drains TMEM while the tensor core already works on tile i+1.
// ------------------------------------------------------------------------------------------------------------
// STAGED: the epilogue converts the accumulator to bf16 into a per-warp, 128B-swizzled shared-memory tile ([32 rows][64
// columns]) and hands it to the TMA unit (cp.async.bulk.tensor store): the SM writes whole 128-byte lines instead of the
// 16-byte-per-row pieces a TMEM load naturally yields (ncu: those doubled the L2 write sectors and held the memory-bound
// 1x1 layers at ~50 % of the HBM roof).  The same tile feeds the optional BatchNorm sums.  !STAGED (row-per-lane stores
// from registers) serves the variants TMA cannot express: strided scatter, read-modify-write accumulate, fp32 logits.
// B_MN: the B operand is taken as [64 reduction rows][64 output columns] boxes straight from KRSC filters and consumed
// MN-major — the dgrad case, where the reduction runs over K (filter rows) and the output over C (contiguous), with the
// tap flipped by index.  No transposed copy of the filters is needed.
// EPI: 0 = plain, 1 = folded BatchNorm (inference fprop) with an identity tile, 3 = the same without one, 2 = BatchNorm-backward sums of
// the producing layer (dgrad)
template <int BN, int STAGES, bool STATS, bool B_MN, int EPI>
__global__ void __launch_bounds__(kThreads)
conv_tc_gemm_persistent_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                               const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmR, const GemmArgs g) {
  constexpr bool FOLD = EPI == 1 || EPI == 3, BNB = EPI == 2, AUX = EPI == 1 || EPI == 2;   // AUX: the epilogue reads a tile (identity / x0)
  static_assert(!BNB || STATS, "the BatchNorm-backward sums ride on the staged (TMA store) epilogue");
  constexpr int ACC = 2;
  constexpr int kBBytes = BN * kBlockK * 2;
  constexpr int kStageBytes = kABytes + kBBytes;
  constexpr uint32_t kTmemCols = ACC * BN;  // 128, 256 or 512: powers of two
  // SWIZZLE_128B tiles need 1024-byte alignment; the kernel has no static shared memory, so the dynamic window starts
  // at the CTA's (1 KB-granular) base.  Checked below rather than paid for with a kilobyte of slack: with it, two
  // CTAs of the BN<=128 variants would no longer fit one SM.
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0) {
    if (threadIdx.x == 0) printf("dv conv_tc: dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }
  uint8_t* stage_out = smem + STAGES * kStageBytes;  // STATS: 4 epilogue warps x [32 rows][128 B], 16-byte chunks XOR-swizzled by row
  // AUX: a second set of per-warp buffers for the tile the epilogue reads (identity / x0), fetched one column group ahead
  uint8_t* stage_aux = stage_out + (STATS ? kEpiStageBytes : 0);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(stage_aux + (AUX ? kEpiStageBytes : 0));
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + ACC;
  uint64_t* res_bar = tempty_bar + ACC;  // one per epilogue warp: the identity tile of a folded-BN epilogue has landed
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(res_bar + 4);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int num_n = (g.Nout + BN - 1) / BN;
  const int num_m = (g.M + kBlockM - 1) / kBlockM;
  const int total = num_m * num_n;
  const int num_kb = g.taps * g.kb_per_tap;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    if (STATS) prefetch_tmap(&tmO);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < ACC; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], 4);  // one arrival per epilogue warp
    }
    for (int i = 0; i < 4; ++i) mbar_init(&res_bar[i], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_ptr, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
      const int n_tile = tile % num_n, m_tile = tile / num_n;
      const int m0 = m_tile * kBlockM;
      int n_img = 0, h0 = 0, w0 = 0;
      if (g.a_im2col) {
        const int pq = g.P * g.Q;
        n_img = m0 / pq;
        const int rem = m0 - n_img * pq;
        const int p0 = rem / g.Q, q0 = rem - p0 * g.Q;
        h0 = p0 * g.stride - g.pad;
        w0 = q0 * g.stride - g.pad;
      }
      int tap = 0, cb = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          uint8_t* sA = smem + stage * kStageBytes;
          uint8_t* sB = sA + kABytes;
          mbar_expect_tx(&full_bar[stage], kStageBytes);
          if (g.a_im2col) {
            const int r = tap / g.S, s = tap - r * g.S;
            tma_load_im2col(sA, &tmA, &full_bar[stage], cb * kBlockK, w0, h0, n_img, (uint16_t)s, (uint16_t)r);
          } else {
            tma_load_2d(sA, &tmA, &full_bar[stage], cb * kBlockK, m0);
          }
          if (B_MN) {
            // filters [K][taps*Nout]: rows = reduction channels cb*64.., columns = flipped tap's slice of the output channels
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sB + j * 8192, &tmB, &full_bar[stage], (g.use_btap ? g.btap[tap] : g.taps - 1 - tap) * g.Nout + n_tile * BN + j * 64,
                          cb * kBlockK);
          } else {
            tma_load_2d(sB, &tmB, &full_bar[stage], (g.use_btap ? g.btap[tap] : tap) * g.Cin + cb * kBlockK, n_tile * BN);
          }
        }
        __syncwarp();
        if (++cb == g.kb_per_tap) { cb = 0; ++tap; }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    {
      constexpr uint32_t idesc = make_idesc(kBlockM, BN, 0, B_MN ? 1 : 0);
      constexpr uint32_t kStageStep = kStageBytes >> 4;            // descriptor units (16 B) between ring stages
      constexpr uint32_t kBStep = B_MN ? (2048u >> 4) : 2u;       // one K=16 step inside a stage
      const uint32_t a_lo0 = desc_lo_kmajor(smem_u32(smem));
      const uint32_t b_lo0 = B_MN ? desc_lo_mnmajor(smem_u32(smem) + kABytes) : desc_lo_kmajor(smem_u32(smem) + kABytes);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);  // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_lo = a_lo0 + (uint32_t)stage * kStageStep;
            const uint32_t b_lo = b_lo0 + (uint32_t)stage * kStageStep;
            umma_f16(d_tmem, desc_pack(a_lo), desc_pack(b_lo), idesc, kb ? 1u : 0u);
#pragma unroll
            for (int k = 1; k < kBlockK / 16; ++k)
              umma_f16(d_tmem, desc_pack(a_lo + k * 2u), desc_pack(b_lo + k * kBStep), idesc, 1u);
            umma_commit(&empty_bar[stage]);
            if (kb == num_kb - 1) umma_commit(&tfull_bar[acc]);  // same thread as the MMAs it tracks
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc == ACC) { acc = 0; acc_phase ^= 1; }
      }
    }
    __syncwarp();
  } else {
    // Epilogue.  A TMEM load hands each lane one ROW (32 fp32 columns); storing that directly would make every store
    // instruction touch 32 different 128-byte lines.  Instead the warp transposes through its private 4 KB of shared
    // memory, 64 columns at a time: lanes write their row's 8 x 16 B chunks (XOR-swizzled by row), then read back so that
    // 8 consecutive lanes hold one row's contiguous 128 B, and each store instruction covers 4 whole lines.
    const int quarter = warp & 3;
    uint8_t* my_stage = stage_out + quarter * (32 * 128);
    const uint32_t my_stage_u32 = smem_u32(my_stage);
    uint8_t* my_aux = stage_aux + quarter * (32 * 128);
    const uint32_t my_aux_u32 = smem_u32(my_aux);
    (void)my_aux_u32;
    const bool aux_wanted = AUX && ((FOLD && g.ep_res != nullptr) || BNB);
    // one elected lane asks TMA for the [32 rows][64 columns] tile of the identity / x0 tensor that column group `cgp` will read
    auto aux_request = [&](int n_tile_, int cgp_, int m_base_) {
      mbar_expect_tx(&res_bar[quarter], 32 * 128);
      tma_load_2d(my_aux, &tmR, &res_bar[quarter], n_tile_ * BN + cgp_ * 64, m_base_);
    };
    const int sub_row = lane >> 3, chunk = lane & 7;  // read-back role: row (it*4 + sub_row), 16-byte chunk `chunk`
    int acc = 0;
    uint32_t acc_phase = 0, res_phase = 0;
    // running per-channel sums of this CTA's tiles (lane owns columns 2*lane, 2*lane+1 of every 64-column group)
    float s1[BN / 64][2], s2[BN / 64][2];
    int stat_n_tile = -1;
#pragma unroll
    for (int i = 0; i < BN / 64; ++i) { s1[i][0] = s1[i][1] = s2[i][0] = s2[i][1] = 0.f; }
    auto flush_stats = [&]() {
      if ((BNB ? (const void*)g.bnb_dgamma : (const void*)g.stat_sums) == nullptr || stat_n_tile < 0) return;
#pragma unroll
      for (int i = 0; i < BN / 64; ++i) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int col = stat_n_tile * BN + i * 64 + 2 * lane + e;
          if (col < g.Nout) {
            if (BNB) {
              atomicAdd(g.bnb_dbeta + col, s1[i][e]);
              atomicAdd(g.bnb_dgamma + col, s2[i][e] * __ldg(g.bnb_invstd + col));
            } else {
              atomicAdd(g.stat_sums + col, s1[i][e]);
              atomicAdd(g.stat_sums + g.Nout + col, s2[i][e]);
            }
          }
          s1[i][e] = 0.f;
          s2[i][e] = 0.f;
        }
      }
    };
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
      const int n_tile = tile % num_n, m_tile = tile / num_n;
      if (n_tile != stat_n_tile) { flush_stats(); stat_n_tile = n_tile; }
      const int m_base = m_tile * kBlockM + quarter * 32;
      // global row offsets of the 8 rows this lane stores (rows it*4 + sub_row)
      size_t row_off[8];
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int m = m_base + it * 4 + sub_row;
        size_t off = 0;
        if (m < g.M) {
          if (g.scatter) {
            const int pq = g.P * g.Q;
            const int n_img = m / pq, rem = m - n_img * pq;
            const int pp = rem / g.Q, qq = rem - pp * g.Q;
            off = (((size_t)n_img * g.Ho + (size_t)pp * g.os + g.oa) * g.Wo + (size_t)qq * g.os + g.ob) * g.Nout;
          } else {
            off = (size_t)m * g.Nout;
          }
        }
        row_off[it] = off;
      }
      const int n_left = g.Nout - n_tile * BN;
      if (AUX && aux_wanted && lane == 0 && m_base < g.M) aux_request(n_tile, 0, m_base);   // lands while the tensor core still works
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN);
      if (!STATS && !g.out_f32) {
        const int m = m_base + lane;
        size_t ro = 0;
        if (m < g.M) {
          if (g.scatter) {
            const int pq = g.P * g.Q;
            const int n_img = m / pq, rem = m - n_img * pq;
            const int pp = rem / g.Q, qq = rem - pp * g.Q;
            ro = (((size_t)n_img * g.Ho + (size_t)pp * g.os + g.oa) * g.Wo + (size_t)qq * g.os + g.ob) * g.Nout;
          } else {
            ro = (size_t)m * g.Nout;
          }
        }
        bf16* orow = g.out + ro + (size_t)n_tile * BN;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(t_addr + (uint32_t)(c * 32), v);
          tmem_ld_wait();
          if (m < g.M) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (c * 32 + j * 8 < n_left) {
                uint32_t w[4];
                if (g.accumulate) {
                  const uint4 old = *reinterpret_cast<const uint4*>(orow + c * 32 + j * 8);
                  const uint32_t ow[4] = {old.x, old.y, old.z, old.w};
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    v[j * 8 + 2 * e] = __float_as_uint(__uint_as_float(v[j * 8 + 2 * e]) + __uint_as_float(ow[e] << 16));
                    v[j * 8 + 2 * e + 1] = __float_as_uint(__uint_as_float(v[j * 8 + 2 * e + 1]) + __uint_as_float(ow[e] & 0xffff0000u));
                  }
                }
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[j * 8 + 2 * e]), __uint_as_float(v[j * 8 + 2 * e + 1]));
                  w[e] = *reinterpret_cast<uint32_t*>(&h);
                }
                *reinterpret_cast<uint4*>(orow + c * 32 + j * 8) = make_uint4(w[0], w[1], w[2], w[3]);
              }
            }
          }
        }
      } else if (!STATS) {
        // fp32 logits (classifier only, tiny): direct row-per-lane stores
        const int m = m_base + lane;
        float* orow32 = reinterpret_cast<float*>(g.out) + (size_t)m * g.Nout + (size_t)n_tile * BN;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(t_addr + (uint32_t)(c * 32), v);
          tmem_ld_wait();
          if (m < g.M) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (c * 32 + j * 4 < n_left)
                *reinterpret_cast<uint4*>(orow32 + c * 32 + j * 4) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          }
        }
      } else {
#pragma unroll
        for (int cgp = 0; cgp < BN / 64; ++cgp) {
          // the TMA store of the previous column group must have finished READING the staging tile
          if (lane == 0) tma_store_wait_read();
          __syncwarp();
          const bool group_live = cgp * 64 < n_left && m_base < g.M;
          const bool with_res = AUX && FOLD && g.ep_res != nullptr && group_live;
          const bool with_x0 = BNB && group_live;
          uint32_t xw[AUX ? 32 : 1];   // BNB: the lane's two columns of x0, 32 rows.  FOLD: the lane's row of the identity tile, 64 columns
          if (with_res || with_x0) {       // (warp-uniform) the tile was requested one column group earlier
            mbar_wait(&res_bar[quarter], res_phase);
            res_phase ^= 1;
            if (BNB) {
              const int wch = lane >> 2, wof = (lane & 3) * 4;
#pragma unroll
              for (int r = 0; r < 32; ++r) xw[r] = *reinterpret_cast<const uint32_t*>(my_aux + r * 128 + ((wch ^ (r & 7)) * 16) + wof);
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                asm volatile("ld.volatile.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(xw[4 * j]), "=r"(xw[4 * j + 1]), "=r"(xw[4 * j + 2]), "=r"(xw[4 * j + 3])
                             : "r"(my_aux_u32 + (uint32_t)(lane * 128 + ((j ^ (lane & 7)) * 16))) : "memory");
            }
            // Every lane has its copy, so the buffer may take the next group's tile while this one is worked on — but the refill is
            // a write by the async proxy, which nothing orders after these generic-proxy reads except a proxy fence: without it ptxas
            // sank the last loads of the row below the TMA issue and they returned the next tile's bytes.
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && cgp + 1 < BN / 64 && (cgp + 1) * 64 < n_left) aux_request(n_tile, cgp + 1, m_base);
          }
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint32_t v[32];
            tmem_ld_32x32(t_addr + (uint32_t)(cgp * 64 + half * 32), v);
            tmem_ld_wait();
            if (FOLD) apply_folded_epilogue(v, g, n_tile * BN + cgp * 64 + half * 32, with_res ? xw + half * 16 : nullptr, m_base + lane);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint32_t w[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[j * 8 + 2 * e]), __uint_as_float(v[j * 8 + 2 * e + 1]));
                w[e] = *reinterpret_cast<uint32_t*>(&h);
              }
              const int ch = (half * 4 + j) ^ (lane & 7);  // SWIZZLE_128B: 16-byte chunk index XOR (row mod 8)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_stage_u32 + lane * 128 + ch * 16), "r"(w[0]), "r"(w[1]),
                           "r"(w[2]), "r"(w[3])
                           : "memory");
            }
          }
          fence_proxy_async_smem();  // make the generic-proxy writes visible to the TMA (async proxy)
          __syncwarp();
          if (lane == 0 && cgp * 64 < n_left && m_base < g.M) {
            if (g.accumulate) tma_reduce_add_2d(&tmO, my_stage_u32, n_tile * BN + cgp * 64, m_base);
            else tma_store_2d(&tmO, my_stage_u32, n_tile * BN + cgp * 64, m_base);
          }
          if (BNB) {
            if (with_x0) {
              const int col = n_tile * BN + cgp * 64 + 2 * lane;   // Nout is a multiple of 8: col and col+1 are both in or both out
              float sc0 = 0.f, sh0 = 0.f, mu0 = 0.f, sc1 = 0.f, sh1 = 0.f, mu1 = 0.f;
              if (col < g.Nout) {
                mu0 = __ldg(g.bnb_mean + col); mu1 = __ldg(g.bnb_mean + col + 1);
                sc0 = __ldg(g.bnb_invstd + col) * __ldg(g.bnb_gamma + col); sc1 = __ldg(g.bnb_invstd + col + 1) * __ldg(g.bnb_gamma + col + 1);
                sh0 = __ldg(g.bnb_beta + col) - mu0 * sc0; sh1 = __ldg(g.bnb_beta + col + 1) - mu1 * sc1;
              }
              float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
              bnb_tile_sums(my_stage, xw, lane, sc0, sh0, mu0, sc1, sh1, mu1, a0, a1, b0, b1);
              s1[cgp][0] += a0; s1[cgp][1] += a1;
              s2[cgp][0] += b0; s2[cgp][1] += b1;
            }
          } else if (g.stat_sums != nullptr) {
            // column sums over this warp's 32 rows from the bf16-rounded values (what BN will read back)
            const int wch = lane >> 2, wof = (lane & 3) * 4;
            float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
#pragma unroll 8
            for (int r = 0; r < 32; ++r) {
              const uint32_t word = *reinterpret_cast<const uint32_t*>(my_stage + r * 128 + ((wch ^ (r & 7)) * 16) + wof);
              const float x0 = __uint_as_float(word << 16), x1 = __uint_as_float(word & 0xffff0000u);
              a0 += x0; a1 += x1;
              b0 = fmaf(x0, x0, b0); b1 = fmaf(x1, x1, b1);
            }
            s1[cgp][0] += a0; s1[cgp][1] += a1;
            s2[cgp][0] += b0; s2[cgp][1] += b1;
          }
        }
      }
      // all of this warp's TMEM reads are complete: hand the accumulator back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      if (++acc == ACC) { acc = 0; acc_phase ^= 1; }
    }
    flush_stats();
    if (STATS && lane == 0) tma_store_wait_read();  // shared memory must outlive the last bulk store's reads
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------------------------------------------------
// CTA-pair variant (BN = 256 layers): two CTAs of a cluster own 256 output pixels x 256 channels.  Each stages ITS 128
// activation rows and HALF of the filter tile; one thread of CTA 0 issues tcgen05.mma.cta_group::2 (M=256, N=256), which
// reads A from each SM's own shared memory and the two B halves from both.  Per SM that is 32 KB of TMA fill and 32 KB of
// operand reads per 512 tensor cycles = 128 B/clk, exactly the shared-memory port, where the single-CTA kernel needs
// 192 B/clk (ncu: 67 % tensor-pipe ceiling).  Full barriers live in CTA 0 (both CTAs' loads count there); empty and
// accumulator-ready barriers are signalled in both CTAs by a multicast commit; the peer's epilogue hands accumulators
// back by a remote arrive.
// ------------------------------------------------------------------------------------------------------------
// EW: epilogue warps per CTA, 4 or 8.  A warp reads the TMEM lanes of quarter (warp % 4); with 8 warps two share a quarter and take
// alternate 64-column groups, each with its own staging / tile buffers, so two warps per scheduler hide each other's latencies.
template <int STAGES, bool B_MN, int EPI, int EW>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(64 + 32 * EW)
conv_tc_gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmR, const GemmArgs g) {
  constexpr bool FOLD = EPI == 1 || EPI == 3, BNB = EPI == 2, AUX = EPI == 1 || EPI == 2;   // AUX: the epilogue reads a tile (identity / x0)
  constexpr int BN = 256, ACC = 2;
  constexpr int kBHalfBytes = (BN / 2) * kBlockK * 2;  // 16 KB: this CTA's half of the filter tile
  constexpr int kStageBytes = kABytes + kBHalfBytes;    // 32 KB
  constexpr uint32_t kTmemCols = ACC * BN;              // 512
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0) {
    if (threadIdx.x == 0) printf("dv conv_tc: dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }
  uint8_t* stage_out = smem + STAGES * kStageBytes;
  static_assert(EW == 4 || EW == 8, "4 or 8 epilogue warps");
  constexpr int kGroupStep = EW / 4;                 // column groups are dealt round-robin to the warps of a quarter
  constexpr int kMyGroups = (BN / 64) / kGroupStep;
  uint8_t* stage_aux = stage_out + EW * 4096;        // AUX: the tile the epilogue reads (identity / x0), one column group ahead
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(stage_aux + (AUX ? EW * 4096 : 0));
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + ACC;
  uint64_t* res_bar = tempty_bar + ACC;  // one per epilogue warp: the identity tile of a folded-BN epilogue has landed
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(res_bar + EW);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int num_n = (g.Nout + BN - 1) / BN;
  const int num_m = (g.M + kBlockM - 1) / kBlockM;
  const int total_pairs = ((num_m + 1) / 2) * num_n;
  const int num_kb = g.taps * g.kb_per_tap;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmO);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 2);   // CTA 0's expect_tx arrive + the peer's remote arrive
      mbar_init(&empty_bar[i], 1);  // multicast commit
    }
    for (int i = 0; i < ACC; ++i) {
      mbar_init(&tfull_bar[i], 1);   // multicast commit
      mbar_init(&tempty_bar[i], 2 * EW);  // the epilogue warps of both CTAs (used in CTA 0 only)
    }
    for (int i = 0; i < EW; ++i) mbar_init(&res_bar[i], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_pair(tmem_ptr, kTmemCols);
  tc_fence_before();
  cluster_sync_all();  // barriers of both CTAs are initialised before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int pair = cluster_id; pair < total_pairs; pair += num_clusters) {
      const int n_tile = pair % num_n, m_tile = 2 * (pair / num_n) + (int)rank;
      const int m0 = m_tile * kBlockM;  // may lie past M for the odd last pair: TMA zero-fills, the epilogue skips
      int n_img = 0, h0 = 0, w0 = 0;
      if (g.a_im2col) {
        const int pq = g.P * g.Q;
        n_img = m0 / pq;
        const int rem = m0 - n_img * pq;
        const int p0 = rem / g.Q, q0 = rem - p0 * g.Q;
        h0 = p0 * g.stride - g.pad;
        w0 = q0 * g.stride - g.pad;
      }
      int tap = 0, cb = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          uint8_t* sA = smem + stage * kStageBytes;
          uint8_t* sB = sA + kABytes;
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * kStageBytes);
          if (g.a_im2col) {
            const int r = tap / g.S, s = tap - r * g.S;
            tma_load_im2col_pair(sA, &tmA, &full_bar[stage], cb * kBlockK, w0, h0, n_img, (uint16_t)s, (uint16_t)r);
          } else {
            tma_load_2d_pair(sA, &tmA, &full_bar[stage], cb * kBlockK, m0);
          }
          if (B_MN) {
            // this CTA's 128 output columns = two [64 k][64 c] boxes of the KRSC filters (flipped tap)
#pragma unroll
            for (int j = 0; j < 2; ++j)
              tma_load_2d_pair(sB + j * 8192, &tmB, &full_bar[stage],
                               (g.taps - 1 - tap) * g.Nout + n_tile * BN + (int)rank * 128 + j * 64, cb * kBlockK);
          } else {
            tma_load_2d_pair(sB, &tmB, &full_bar[stage], tap * g.Cin + cb * kBlockK, n_tile * BN + (int)rank * 128);
          }
          if (rank != 0) mbar_arrive_remote(&full_bar[stage], 0);
        }
        __syncwarp();
        if (++cb == g.kb_per_tap) { cb = 0; ++tap; }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc(256, BN, 0, B_MN ? 1 : 0);
      constexpr uint32_t kStageStep = kStageBytes >> 4;
      constexpr uint32_t kBStep = B_MN ? (2048u >> 4) : 2u;
      const uint32_t a_lo0 = desc_lo_kmajor(smem_u32(smem));
      const uint32_t b_lo0 = B_MN ? desc_lo_mnmajor(smem_u32(smem) + kABytes) : desc_lo_kmajor(smem_u32(smem) + kABytes);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int pair = cluster_id; pair < total_pairs; pair += num_clusters) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);  // both CTAs' epilogues have drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_lo = a_lo0 + (uint32_t)stage * kStageStep;
            const uint32_t b_lo = b_lo0 + (uint32_t)stage * kStageStep;
            umma_f16_pair(d_tmem, desc_pack(a_lo), desc_pack(b_lo), idesc, kb ? 1u : 0u);
#pragma unroll
            for (int k = 1; k < kBlockK / 16; ++k)
              umma_f16_pair(d_tmem, desc_pack(a_lo + k * 2u), desc_pack(b_lo + k * kBStep), idesc, 1u);
            umma_commit_pair(&empty_bar[stage]);
            if (kb == num_kb - 1) umma_commit_pair(&tfull_bar[acc]);
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc == ACC) { acc = 0; acc_phase ^= 1; }
      }
    }
    __syncwarp();
  } else {
    const int quarter = warp & 3;          // TMEM lanes 32*quarter .. +31: rows of the tile
    const int ew = warp - 2;               // buffers and barriers are per warp
    const int cg_first = ew >> 2;          // first column group of this warp (0, or 0/1 with eight warps)
    uint8_t* my_stage = stage_out + ew * (32 * 128);
    const uint32_t my_stage_u32 = smem_u32(my_stage);
    uint8_t* my_aux = stage_aux + ew * (32 * 128);
    const uint32_t my_aux_u32 = smem_u32(my_aux);
    (void)my_aux_u32;
    const bool aux_wanted = AUX && ((FOLD && g.ep_res != nullptr) || BNB);
    // one elected lane asks TMA for the [32 rows][64 columns] tile of the identity / x0 tensor that column group `cgp` will read
    auto aux_request = [&](int n_tile_, int cgp_, int m_base_) {
      mbar_expect_tx(&res_bar[ew], 32 * 128);
      tma_load_2d(my_aux, &tmR, &res_bar[ew], n_tile_ * BN + cgp_ * 64, m_base_);
    };
    int acc = 0;
    uint32_t acc_phase = 0, res_phase = 0;
    float s1[kMyGroups][2], s2[kMyGroups][2];   // entry i: column group cg_first + i * kGroupStep
    int stat_n_tile = -1;
#pragma unroll
    for (int i = 0; i < kMyGroups; ++i) { s1[i][0] = s1[i][1] = s2[i][0] = s2[i][1] = 0.f; }
    auto flush_stats = [&]() {
      if ((BNB ? (const void*)g.bnb_dgamma : (const void*)g.stat_sums) == nullptr || stat_n_tile < 0) return;
#pragma unroll
      for (int i = 0; i < kMyGroups; ++i) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int col = stat_n_tile * BN + (cg_first + i * kGroupStep) * 64 + 2 * lane + e;
          if (col < g.Nout) {
            if (BNB) {
              atomicAdd(g.bnb_dbeta + col, s1[i][e]);
              atomicAdd(g.bnb_dgamma + col, s2[i][e] * __ldg(g.bnb_invstd + col));
            } else {
              atomicAdd(g.stat_sums + col, s1[i][e]);
              atomicAdd(g.stat_sums + g.Nout + col, s2[i][e]);
            }
          }
          s1[i][e] = 0.f;
          s2[i][e] = 0.f;
        }
      }
    };
    for (int pair = cluster_id; pair < total_pairs; pair += num_clusters) {
      const int n_tile = pair % num_n, m_tile = 2 * (pair / num_n) + (int)rank;
      if (n_tile != stat_n_tile) { flush_stats(); stat_n_tile = n_tile; }
      const int m_base = m_tile * kBlockM + quarter * 32;
      const int n_left = g.Nout - n_tile * BN;
      if (AUX && aux_wanted && lane == 0 && m_base < g.M && cg_first * 64 < n_left) aux_request(n_tile, cg_first, m_base);   // lands while the tensor core still works
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll
      for (int gi = 0; gi < kMyGroups; ++gi) {
        const int cgp = cg_first + gi * kGroupStep;
        if (lane == 0) tma_store_wait_read();
        __syncwarp();
        const bool group_live = cgp * 64 < n_left && m_base < g.M;
        const bool with_res = AUX && FOLD && g.ep_res != nullptr && group_live;
        const bool with_x0 = BNB && group_live;
        uint32_t xw[AUX ? 32 : 1];   // BNB: the lane's two columns of x0, 32 rows.  FOLD: the lane's row of the identity tile, 64 columns
        if (with_res || with_x0) {       // (warp-uniform) the tile was requested one column group earlier
          mbar_wait(&res_bar[ew], res_phase);
          res_phase ^= 1;
          if (BNB) {
            const int wch = lane >> 2, wof = (lane & 3) * 4;
#pragma unroll
            for (int r = 0; r < 32; ++r) xw[r] = *reinterpret_cast<const uint32_t*>(my_aux + r * 128 + ((wch ^ (r & 7)) * 16) + wof);
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              asm volatile("ld.volatile.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(xw[4 * j]), "=r"(xw[4 * j + 1]), "=r"(xw[4 * j + 2]), "=r"(xw[4 * j + 3])
                           : "r"(my_aux_u32 + (uint32_t)(lane * 128 + ((j ^ (lane & 7)) * 16))) : "memory");
          }
          fence_proxy_async_smem();   // orders the reads above before the async-proxy write that refills the buffer (see the persistent kernel)
          __syncwarp();
          if (lane == 0 && gi + 1 < kMyGroups && (cgp + kGroupStep) * 64 < n_left) aux_request(n_tile, cgp + kGroupStep, m_base);
        }
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          tmem_ld_32x32(t_addr + (uint32_t)(cgp * 64 + half * 32), v);
          tmem_ld_wait();
          if (FOLD) apply_folded_epilogue(v, g, n_tile * BN + cgp * 64 + half * 32, with_res ? xw + half * 16 : nullptr, m_base + lane);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t w[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[j * 8 + 2 * e]), __uint_as_float(v[j * 8 + 2 * e + 1]));
              w[e] = *reinterpret_cast<uint32_t*>(&h);
            }
            const int ch = (half * 4 + j) ^ (lane & 7);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_stage_u32 + lane * 128 + ch * 16), "r"(w[0]), "r"(w[1]),
                         "r"(w[2]), "r"(w[3])
                         : "memory");
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0 && cgp * 64 < n_left && m_base < g.M) {
          if (g.accumulate) tma_reduce_add_2d(&tmO, my_stage_u32, n_tile * BN + cgp * 64, m_base);
          else tma_store_2d(&tmO, my_stage_u32, n_tile * BN + cgp * 64, m_base);
        }
        if (BNB) {
          if (with_x0) {
            const int col = n_tile * BN + cgp * 64 + 2 * lane;
            float sc0 = 0.f, sh0 = 0.f, mu0 = 0.f, sc1 = 0.f, sh1 = 0.f, mu1 = 0.f;
            if (col < g.Nout) {
              mu0 = __ldg(g.bnb_mean + col); mu1 = __ldg(g.bnb_mean + col + 1);
              sc0 = __ldg(g.bnb_invstd + col) * __ldg(g.bnb_gamma + col); sc1 = __ldg(g.bnb_invstd + col + 1) * __ldg(g.bnb_gamma + col + 1);
              sh0 = __ldg(g.bnb_beta + col) - mu0 * sc0; sh1 = __ldg(g.bnb_beta + col + 1) - mu1 * sc1;
            }
            float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
            bnb_tile_sums(my_stage, xw, lane, sc0, sh0, mu0, sc1, sh1, mu1, a0, a1, b0, b1);
            s1[gi][0] += a0; s1[gi][1] += a1;
            s2[gi][0] += b0; s2[gi][1] += b1;
          }
        } else if (g.stat_sums != nullptr) {
          const int wch = lane >> 2, wof = (lane & 3) * 4;
          float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
#pragma unroll 8
          for (int r = 0; r < 32; ++r) {
            const uint32_t word = *reinterpret_cast<const uint32_t*>(my_stage + r * 128 + ((wch ^ (r & 7)) * 16) + wof);
            const float x0 = __uint_as_float(word << 16), x1 = __uint_as_float(word & 0xffff0000u);
            a0 += x0; a1 += x1;
            b0 = fmaf(x0, x0, b0); b1 = fmaf(x1, x1, b1);
          }
          s1[gi][0] += a0; s1[gi][1] += a1;
          s2[gi][0] += b0; s2[gi][1] += b1;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (rank == 0) mbar_arrive(&tempty_bar[acc]);
        else mbar_arrive_remote(&tempty_bar[acc], 0);
      }
      if (++acc == ACC) { acc = 0; acc_phase ^= 1; }
    }
    flush_stats();
    if (lane == 0) tma_store_wait_read();
  }
  tc_fence_before();
  cluster_sync_all();  // neither CTA may retire (or free TMEM) while the pair still issues MMAs / remote arrives
  if (warp == 1) tmem_dealloc_pair(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------------------------------------------------
// wgrad kernel
// ------------------------------------------------------------------------------------------------------------
struct WgradArgs {
  float* dw_atomic; // non-null: add the tile straight into dw [K][taps][C] with vector reductions (no partials, no 2nd pass)
  float* ws;        // [splits][K][taps][C] fp32 partial sums
  int Mpix;         // N*P*Q
  int K, C, taps, S;
  int P, Q, stride, pad;
  int x_im2col;     // 0: x is a 2-D [Mpix][C] map (1x1 stride 1), 1: im2col 4-D map
  int kb_per_split; // 64-pixel blocks per split
  int kb_total;
  int c_tiles;
  int tma_reduce;   // accumulate mode: 1 = fp32 tiles go to dW through TMA reduce-add, 0 = per-lane red.global.add
};

// MH = number of 128-row halves of the output tile (1 or 2).  With MH == 2 a CTA owns 256 output channels: two
// accumulators in TMEM, two MMAs per K step that share ONE staged x tile, so the activation is fetched from L2 once per
// 256 (not 128) output channels.
template <int BNC, int STAGES, int MH>
__global__ void __launch_bounds__(kThreads)
conv_tc_wgrad_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                     const __grid_constant__ CUtensorMap tmDW, const WgradArgs g) {
  constexpr int kPix = 64;                       // pixels per stage (= 4 MMAs of K=16)
  constexpr int kChunk = kPix * 128;             // one [64 px][64 ch] box = 8 KB
  constexpr int kABytesW = MH * 2 * kChunk;      // 128*MH output channels of dY
  constexpr int kBBytesW = (BNC / 64) * kChunk;  // BNC input channels of x
  constexpr int kStageBytes = kABytesW + kBBytesW;
  constexpr int kAccCols = MH * BNC;
  constexpr uint32_t kTmemCols = kAccCols <= 32 ? 32 : kAccCols <= 64 ? 64 : kAccCols <= 128 ? 128 : kAccCols <= 256 ? 256 : 512;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * kStageBytes);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  // tile index: tap fastest so the CTAs that share one dY tile are neighbours in launch order (L2 reuse)
  int t = blockIdx.x;
  const int tap = t % g.taps; t /= g.taps;
  const int c_tile = t % g.c_tiles;
  const int k_tile = t / g.c_tiles;
  const int split = blockIdx.y;
  const int kb_begin = split * g.kb_per_split;
  const int kb_end = min(g.kb_total, kb_begin + g.kb_per_split);
  const int num_kb = kb_end - kb_begin;  // >= 1 by construction

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmDY);
    prefetch_tmap(&tmX);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_ptr, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    const int r = tap / g.S, s = tap - r * g.S;
    const int pq = g.P * g.Q;
    int stage = 0;
    uint32_t phase = 0;
    for (int i = 0; i < num_kb; ++i) {
      const int pix0 = (kb_begin + i) * kPix;
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (elect_one()) {
        uint8_t* sA = smem + stage * kStageBytes;
        uint8_t* sB = sA + kABytesW;
        mbar_expect_tx(&full_bar[stage], kStageBytes);
#pragma unroll
        for (int j = 0; j < 2 * MH; ++j) tma_load_2d(sA + j * kChunk, &tmDY, &full_bar[stage], k_tile * (128 * MH) + j * 64, pix0);
        if (g.x_im2col) {
          const int n_img = pix0 / pq, rem = pix0 - n_img * pq;
          const int p0 = rem / g.Q, q0 = rem - p0 * g.Q;
          const int h0 = p0 * g.stride - g.pad, w0 = q0 * g.stride - g.pad;
#pragma unroll
          for (int j = 0; j < BNC / 64; ++j)
            tma_load_im2col(sB + j * kChunk, &tmX, &full_bar[stage], c_tile * BNC + j * 64, w0, h0, n_img, (uint16_t)s, (uint16_t)r);
        } else {
#pragma unroll
          for (int j = 0; j < BNC / 64; ++j) tma_load_2d(sB + j * kChunk, &tmX, &full_bar[stage], c_tile * BNC + j * 64, pix0);
        }
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(128, BNC, 1, 1);
    constexpr uint32_t kStageStep = kStageBytes >> 4;
    // MN-major SWIZZLE_128B: a row is one pixel (128 B = 64 channels); 8-pixel groups are 1024 B apart (SBO), 64-channel
    // chunks are kChunk apart (LBO); a 16-pixel K step is +2048 B; the second 128-row half of dY starts 2*kChunk further.
    static_assert(kChunk == 8192, "desc_lo_mnmajor encodes LBO = 8192");
    const uint32_t a_lo0 = desc_lo_mnmajor(smem_u32(smem));
    const uint32_t b_lo0 = desc_lo_mnmajor(smem_u32(smem) + kABytesW);
    int stage = 0;
    uint32_t phase = 0;
    for (int i = 0; i < num_kb; ++i) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t a_lo = a_lo0 + (uint32_t)stage * kStageStep;
        const uint32_t b_lo = b_lo0 + (uint32_t)stage * kStageStep;
#pragma unroll
        for (int k = 0; k < kPix / 16; ++k) {
#pragma unroll
          for (int h = 0; h < MH; ++h)
            umma_f16(tmem_base + (uint32_t)(h * BNC), desc_pack(a_lo + (uint32_t)(h * 2 * kChunk + k * 2048) / 16u),
                     desc_pack(b_lo + (uint32_t)(k * 2048) / 16u), idesc, (k == 0 && i == 0) ? 0u : 1u);
        }
        umma_commit(&empty_bar[stage]);
        if (i == num_kb - 1) umma_commit(tmem_full_bar);
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else {
    const int quarter = warp & 3;
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    if (g.dw_atomic != nullptr && g.tma_reduce) {
      // Every MMA has completed, so the operand ring is free: each warp stages its [32 filters][32 channels] fp32 chunks there
      // (row = lane, 128 B, 16-byte pieces XOR-swizzled by row) and hands them to the TMA unit, which ADDS whole 128-byte lines
      // into dW at the L2 (cp.reduce.async.bulk) and clips what lies outside [K][taps][C].  The per-lane red.global.add.v4.f32 it
      // replaces touched 32 different lines per instruction, half a sector each.
      uint8_t* tiles = smem + (size_t)quarter * (MH * (BNC / 32)) * 4096;
#pragma unroll 1
      for (int h = 0; h < MH; ++h) {
#pragma unroll 1
        for (int c = 0; c < BNC / 32; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(h * BNC + c * 32), v);
          tmem_ld_wait();
          const uint32_t tile = smem_u32(tiles + (size_t)(h * (BNC / 32) + c) * 4096);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(tile + lane * 128 + ((j ^ (lane & 7)) * 16)), "r"(v[4 * j]),
                         "r"(v[4 * j + 1]), "r"(v[4 * j + 2]), "r"(v[4 * j + 3])
                         : "memory");
          fence_proxy_async_smem();
          __syncwarp();
          const int c0 = c_tile * BNC + c * 32, k0 = k_tile * (128 * MH) + h * 128 + quarter * 32;
          if (lane == 0 && c0 < g.C && k0 < g.K) tma_reduce_add_3d(&tmDW, tile, c0, tap, k0);
        }
      }
      if (lane == 0) tma_store_wait_read();   // the bulk reductions have read their tiles before shared memory goes away
      __syncwarp();
    } else
#pragma unroll 1
    for (int h = 0; h < MH; ++h) {
      const int k = k_tile * (128 * MH) + h * 128 + quarter * 32 + lane;
      const size_t kk = (size_t)(k < g.K ? k : 0);
      float* orow = (g.dw_atomic != nullptr ? g.dw_atomic + (kk * g.taps + tap) * g.C
                                            : g.ws + (((size_t)split * g.K + kk) * g.taps + tap) * g.C) + (size_t)c_tile * BNC;
#pragma unroll 1
      for (int c = 0; c < BNC / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(h * BNC + c * 32), v);
        tmem_ld_wait();
        if (k < g.K) {
          const int c_left = g.C - c_tile * BNC - c * 32;  // valid input channels from this chunk on
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (j * 4 < c_left) {
              if (g.dw_atomic != nullptr)
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(orow + c * 32 + j * 4), "f"(__uint_as_float(v[4 * j])),
                             "f"(__uint_as_float(v[4 * j + 1])), "f"(__uint_as_float(v[4 * j + 2])), "f"(__uint_as_float(v[4 * j + 3]))
                             : "memory");
              else
                *reinterpret_cast<uint4*>(orow + c * 32 + j * 4) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// dw[i] = sum over splits of ws[split][i]   (fixed order => bit-reproducible)
__global__ void wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw, int64_t n4, int splits) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  float4 acc = reinterpret_cast<const float4*>(ws)[i];
  for (int s = 1; s < splits; ++s) {
    const float4 v = reinterpret_cast<const float4*>(ws)[(int64_t)s * n4 + i];
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  reinterpret_cast<float4*>(dw)[i] = acc;
}

// ------------------------------------------------------------------------------------------------------------
// host side: tensor maps
// ------------------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct DriverApi {
  EncodeTiledFn tiled = nullptr;
  EncodeIm2colFn im2col = nullptr;
  int driver_version = 0;
  bool ok = false;
};

static DriverApi& driver_api() {
  static DriverApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f1 = nullptr;
    void* f2 = nullptr;
    cudaDriverEntryPointQueryResult q1, q2;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f1, cudaEnableDefault, &q1) != cudaSuccess ||
        q1 != cudaDriverEntryPointSuccess)
      return;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &f2, cudaEnableDefault, &q2) != cudaSuccess ||
        q2 != cudaDriverEntryPointSuccess)
      return;
    api.tiled = (EncodeTiledFn)f1;
    api.im2col = (EncodeIm2colFn)f2;
    cudaDriverGetVersion(&api.driver_version);
    api.ok = true;
  });
  return api;
}

// Descriptor cache: a layer is launched with the same pointers and geometry step after step (the caching
// allocator hands an eager forward the same blocks again, a captured step never changes them), so an encoded CUtensorMap is
// kept under the arguments it was encoded from instead of being rebuilt by the driver three or four times per launch.
// Per host thread (no lock), bounded: when full it is simply emptied.
struct MapKey {
  uint64_t v[12];
  bool operator==(const MapKey& o) const { return memcmp(v, o.v, sizeof(v)) == 0; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    uint64_t h = 1469598103934665603ull;
    for (uint64_t x : k.v) { h ^= x; h *= 1099511628211ull; }
    return (size_t)h;
  }
};
static std::unordered_map<MapKey, CUtensorMap, MapKeyHash>& map_cache() {
  thread_local std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  return cache;
}
static bool map_cache_get(const MapKey& key, CUtensorMap* tm) {
  static const bool enabled = getenv("DV_TC_MAP_CACHE") == nullptr || atoi(getenv("DV_TC_MAP_CACHE")) != 0;
  if (!enabled) return false;
  auto& c = map_cache();
  auto it = c.find(key);
  if (it == c.end()) return false;
  *tm = it->second;
  return true;
}
static void map_cache_put(const MapKey& key, const CUtensorMap& tm) {
  auto& c = map_cache();
  if (c.size() >= 8192) c.clear();
  c.emplace(key, tm);
}

// [rows][cols] bf16 row-major, box = box_rows x 64 columns, 128B swizzle, OOB reads give zeros
static int make_map_2d(CUtensorMap* tm, const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  const MapKey key{{1, (uint64_t)(uintptr_t)ptr, rows, cols, box_rows}};
  if (map_cache_get(key, tm)) return DV_OK;
  DriverApi& api = driver_api();
  DV_REQUIRE(api.ok, DV_ERR_CUDA, "conv_tc: cuTensorMapEncode* entry points not available from the driver");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = api.tiled(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DV_REQUIRE(r == CUDA_SUCCESS, DV_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) rows=%llu cols=%llu box_rows=%u", (int)r,
             (unsigned long long)rows, (unsigned long long)cols, box_rows);
  map_cache_put(key, *tm);
  return DV_OK;
}

// NHWC activation, im2col traversal of a R x S window with the given stride / padding; one load = `pixels` output
// positions (walking q, then p, then n) x 64 channels.  Coordinates passed to the load are the input coordinate of
// the window's top-left tap for the first output position: (c, q*stride - pad, p*stride - pad, n), offsets = (s, r).
// dW as a 3-D fp32 tensor {C, taps, K} with a box of 32 channels x 1 tap x 32 filters (one epilogue warp's chunk)
static int make_map_dw(CUtensorMap* tm, float* dw, int K, int taps, int C) {
  const MapKey key{{2, (uint64_t)(uintptr_t)dw, (uint64_t)K, (uint64_t)taps, (uint64_t)C}};
  if (map_cache_get(key, tm)) return DV_OK;
  DriverApi& api = driver_api();
  DV_REQUIRE(api.ok, DV_ERR_CUDA, "conv_tc: cuTensorMapEncode* entry points not available from the driver");
  cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)taps, (cuuint64_t)K};
  cuuint64_t strides[2] = {(cuuint64_t)C * 4, (cuuint64_t)taps * C * 4};
  cuuint32_t box[3] = {32, 1, 32};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = api.tiled(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, dw, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DV_REQUIRE(r == CUDA_SUCCESS, DV_ERR_CUDA, "cuTensorMapEncodeTiled(dW) failed (%d) K=%d taps=%d C=%d", (int)r, K, taps, C);
  map_cache_put(key, *tm);
  return DV_OK;
}

// corners != nullptr: {lower_w, lower_h, upper_w, upper_h} given explicitly (window origins run over [lower, dim - 1 + upper])
static int make_map_im2col(CUtensorMap* tm, const void* ptr, int N, int H, int W, int C, int R, int S, int stride, int pad,
                           uint32_t pixels, const int* corners = nullptr) {
  MapKey key{{3, (uint64_t)(uintptr_t)ptr, ((uint64_t)(uint32_t)N << 32) | (uint32_t)H, ((uint64_t)(uint32_t)W << 32) | (uint32_t)C,
              ((uint64_t)(uint32_t)R << 32) | (uint32_t)S, ((uint64_t)(uint32_t)stride << 32) | (uint32_t)pad, pixels, corners ? 1u : 0u}};
  if (corners) {
    key.v[8] = ((uint64_t)(uint32_t)corners[0] << 32) | (uint32_t)corners[1];
    key.v[9] = ((uint64_t)(uint32_t)corners[2] << 32) | (uint32_t)corners[3];
  }
  if (map_cache_get(key, tm)) return DV_OK;
  DriverApi& api = driver_api();
  DV_REQUIRE(api.ok, DV_ERR_CUDA, "conv_tc: cuTensorMapEncode* entry points not available from the driver");
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  // bounding box of window origins: [-pad, dim - 1 + pad - (filter - 1)]  (W first, then H)
  int lower[2] = {-pad, -pad};
  int upper[2] = {pad - (S - 1), pad - (R - 1)};
  if (corners) { lower[0] = corners[0]; lower[1] = corners[1]; upper[0] = corners[2]; upper[1] = corners[3]; }
  cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  CUresult r = api.im2col(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, lower, upper, 64,
                          pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DV_REQUIRE(r == CUDA_SUCCESS, DV_ERR_CUDA, "cuTensorMapEncodeIm2col failed (%d) NHWC=%d,%d,%d,%d RS=%d,%d stride=%d pad=%d",
             (int)r, N, H, W, C, R, S, stride, pad);
  // Drivers up to 13.1 set a descriptor bit that breaks im2col loads of tensors smaller than 128 KiB; clear it
  // (the same correction CUTLASS applies in make_im2col_tma_copy_desc).
  if (api.driver_version <= 13010 && (size_t)N * H * W * C * 2 < 131072) {
    reinterpret_cast<uint64_t*>(tm)[1] &= ~(1ull << 21);
  }
  map_cache_put(key, *tm);
  return DV_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int STAGES, bool B_MN, int EPI, int EW>
static int launch_gemm_pair_ew(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmO, const CUtensorMap& tmR, const GemmArgs& g,
                               cudaStream_t st) {
  constexpr int smem = STAGES * (kABytes + 128 * kBlockK * 2) + EW * 4096 * ((EPI == 1 || EPI == 2) ? 2 : 1) + 256;
  static_assert(smem <= 227 * 1024, "pair kernel ring exceeds shared memory");
  DV_CUDA_OK(cudaFuncSetAttribute(conv_tc_gemm_pair_kernel<STAGES, B_MN, EPI, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int num_n = ceil_div(g.Nout, 256);
  const int pairs = ceil_div(ceil_div(g.M, kBlockM), 2) * num_n;
  int clusters = num_sms() / 2;
  if (clusters > num_n) clusters -= clusters % num_n;
  if (clusters > pairs) clusters = pairs;
  conv_tc_gemm_pair_kernel<STAGES, B_MN, EPI, EW><<<2 * clusters, 64 + 32 * EW, smem, st>>>(tmA, tmB, tmO, tmR, g);  // __cluster_dims__(2,1,1)
  DV_LAUNCH_OK("conv_tc_gemm_pair_kernel");
  return DV_OK;
}

static int env_int(const char* name, int dflt);
// STAGES is the ring depth with four epilogue warps and no epilogue tile; eight warps (DV_TC_EW, default 8) double the staging
// buffers, an epilogue tile doubles them again, and the ring gives up a stage when the 227 KB would be exceeded.
template <int STAGES, bool B_MN, int EPI>
static int launch_gemm_pair(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmO, const CUtensorMap& tmR, const GemmArgs& g,
                            cudaStream_t st) {
  static const int ew = env_int("DV_TC_EW", 8);
  if (ew == 8) return launch_gemm_pair_ew<((EPI == 1 || EPI == 2) ? STAGES - 1 : STAGES), B_MN, EPI, 8>(tmA, tmB, tmO, tmR, g, st);
  return launch_gemm_pair_ew<STAGES, B_MN, EPI, 4>(tmA, tmB, tmO, tmR, g, st);
}

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}

template <int BN, int STAGES, bool STATS, bool B_MN, int EPI>
static int launch_gemm_persistent(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmO, const CUtensorMap& tmR, const GemmArgs& g,
                                  cudaStream_t st) {
  constexpr int smem = STAGES * (kABytes + BN * kBlockK * 2) + (STATS ? kEpiStageBytes : 0) + ((EPI == 1 || EPI == 2) ? kEpiStageBytes : 0) + 256;
  static_assert(smem <= 227 * 1024, "persistent kernel ring exceeds shared memory");
  DV_CUDA_OK(cudaFuncSetAttribute(conv_tc_gemm_persistent_kernel<BN, STAGES, STATS, B_MN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  // ask for the whole L1/shared array as shared memory: otherwise the driver sizes the carve-out for ONE CTA of this
  // kernel and a second CTA never becomes resident (measured: occupancy 1 -> 2 for the 96 KB variants)
  DV_CUDA_OK(cudaFuncSetAttribute(conv_tc_gemm_persistent_kernel<BN, STAGES, STATS, B_MN, EPI>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
  // CTAs per SM the grid is sized for: two when two rings (<= 113 KB each) and two accumulator pairs (2 x 2 x BN <= 512 TMEM
  // columns) fit, else one.  cudaOccupancyMaxActiveBlocksPerMultiprocessor answers 1 for these kernels at ANY shared
  // memory size, but grids sized for two per SM measured 40-60 % faster on the BN <= 128 layers
  // (profiles/r01_layer_bench_v2 vs v4), so the answer is not used.
  static int occ_cached = 0;
  if (occ_cached == 0) {
    int occ = 0;
    DV_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, conv_tc_gemm_persistent_kernel<BN, STAGES, STATS, B_MN, EPI>, kThreads, smem));
    occ_cached = (smem <= 115712 && 4 * BN <= 512) ? 2 : 1;
    const int forced = env_int("DV_TC_OCC", 0);  // experiment knob
    if (forced > 0) occ_cached = forced;
    if (env_int("DV_TC_DEBUG", 0)) {
      cudaFuncAttributes fa;
      cudaFuncGetAttributes(&fa, conv_tc_gemm_persistent_kernel<BN, STAGES, STATS, B_MN, EPI>);
      int sweep[4] = {0, 0, 0, 0};
      const int sizes[4] = {32 << 10, 64 << 10, 96 << 10, 110 << 10};
      for (int i = 0; i < 4; ++i)
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&sweep[i], conv_tc_gemm_persistent_kernel<BN, STAGES, STATS, B_MN, EPI>, kThreads, sizes[i]);
      int dev = 0, smem_sm = 0, smem_res = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
      cudaDeviceGetAttribute(&smem_res, cudaDevAttrReservedSharedMemoryPerBlock, dev);
      fprintf(stderr,
              "dv conv_tc: gemm<BN=%d,STAGES=%d,STATS=%d> smem=%d occupancy(api)=%d used=%d regs=%d static_smem=%zu local=%zu maxdyn=%d "
              "occ@32K/64K/96K/110K=%d/%d/%d/%d smem_per_sm=%d reserved=%d\n",
              BN, STAGES, (int)STATS + 2 * (int)B_MN, smem, occ, occ_cached, fa.numRegs, fa.sharedSizeBytes, fa.localSizeBytes,
              fa.maxDynamicSharedSizeBytes, sweep[0], sweep[1], sweep[2], sweep[3], smem_sm, smem_res);
    }
  }
  const int num_n = ceil_div(g.Nout, BN);
  const int total = num_n * ceil_div(g.M, kBlockM);
  int slots = num_sms() * occ_cached;
  if (slots > num_n) slots -= slots % num_n;  // a CTA then keeps one n-tile for its whole life (running BN sums stay put)
  const int grid = total < slots ? total : slots;
  conv_tc_gemm_persistent_kernel<BN, STAGES, STATS, B_MN, EPI><<<grid, kThreads, smem, st>>>(tmA, tmB, tmO, tmR, g);
  DV_LAUNCH_OK("conv_tc_gemm_persistent_kernel");
  return DV_OK;
}

template <int BNC, int STAGES, int MH>
static int launch_wgrad(const CUtensorMap& tmDY, const CUtensorMap& tmX, const CUtensorMap& tmDW, const WgradArgs& g, int tiles, int splits,
                        cudaStream_t st) {
  constexpr int smem = STAGES * (2 * MH + BNC / 64) * 8192 + 1024 + 256;
  static_assert(smem <= 227 * 1024, "wgrad ring exceeds shared memory");
  DV_CUDA_OK(cudaFuncSetAttribute(conv_tc_wgrad_kernel<BNC, STAGES, MH>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  DV_CUDA_OK(cudaFuncSetAttribute(conv_tc_wgrad_kernel<BNC, STAGES, MH>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
  dim3 grid((unsigned)tiles, (unsigned)splits);
  static_assert(MH == 2 || 4 * (BNC / 32) * 4096 <= STAGES * (2 * MH + BNC / 64) * 8192, "epilogue tiles must fit the operand ring");
  conv_tc_wgrad_kernel<BNC, STAGES, MH><<<grid, kThreads, smem, st>>>(tmDY, tmX, tmDW, g);
  DV_LAUNCH_OK("conv_tc_wgrad_kernel");
  return DV_OK;
}

struct WgradPlan {
  int bnc, mh, c_tiles, k_tiles, tiles, kb_total, kb_per_split, splits;
};

static WgradPlan plan_wgrad(const dv_conv_params* p) {
  WgradPlan w;
  const int P = (p->H + 2 * p->pad - p->R) / p->stride + 1, Q = (p->W + 2 * p->pad - p->S) / p->stride + 1;
  const int64_t mpix = (int64_t)p->N * P * Q;
  w.bnc = p->C >= 256 ? 256 : ceil_div(p->C, 64) * 64;  // 64, 128, 192 or 256 input channels per tile
  w.c_tiles = ceil_div(p->C, w.bnc);
  static const int mh_env = env_int("DV_TC_WGRAD_MH", 1);  // 2 = 256 output channels per CTA (measured slower: more split partials)
  w.mh = (mh_env == 2 && p->K > 128) ? 2 : 1;
  w.k_tiles = ceil_div(p->K, 128 * w.mh);
  w.tiles = w.k_tiles * w.c_tiles * p->R * p->S;
  w.kb_total = (int)ceil_div64(mpix, 64);
  // ONE wave of CTAs: two per SM while two rings fit (MH == 1, BNC <= 128), else one; at least 4 stages of work each.
  // (Measured against "2 x SMs CTAs whatever the tile" and against one per SM: this split is 10-40 % faster.)
  const int slots = num_sms() * ((w.mh == 1 && w.bnc <= 128) ? 2 : 1);
  int want = slots / w.tiles;
  int max_by_len = w.kb_total / 4 > 0 ? w.kb_total / 4 : 1;
  int splits = want < max_by_len ? want : max_by_len;
  if (splits < 1) splits = 1;
  w.kb_per_split = ceil_div(w.kb_total, splits);
  w.splits = ceil_div(w.kb_total, w.kb_per_split);
  return w;
}

}  // namespace

int make_tiled_map_bf16(void* tm, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                        int swizzle) {
  DriverApi& api = driver_api();
  DV_REQUIRE(api.ok, DV_ERR_CUDA, "cuTensorMapEncode* entry points not available from the driver");
  DV_REQUIRE(rank >= 2 && rank <= 5, DV_ERR_ARG, "make_tiled_map_bf16: rank %d", rank);
  MapKey key{};
  key.v[0] = 16 + (uint64_t)rank * 8 + (uint64_t)swizzle;
  key.v[1] = (uint64_t)(uintptr_t)ptr;
  for (int i = 0; i < rank; ++i) key.v[2 + i] = dims[i] | ((uint64_t)box[i] << 40);
  for (int i = 0; i + 1 < rank; ++i) key.v[7 + i] = strides_bytes[i];
  if (map_cache_get(key, (CUtensorMap*)tm)) return DV_OK;
  cuuint64_t d[5], st[4];
  cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; e[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
  const CUtensorMapSwizzle sw = swizzle == 3 ? CU_TENSOR_MAP_SWIZZLE_128B : swizzle == 2 ? CU_TENSOR_MAP_SWIZZLE_64B
                              : swizzle == 1 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = api.tiled((CUtensorMap*)tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), d, st, b, e,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DV_REQUIRE(r == CUDA_SUCCESS, DV_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) rank=%d dims0=%llu box0=%u", (int)r, rank,
             (unsigned long long)dims[0], box[0]);
  map_cache_put(key, *(CUtensorMap*)tm);
  return DV_OK;
}

// ------------------------------------------------------------------------------------------------------------
// dispatch
// ------------------------------------------------------------------------------------------------------------
bool conv_tc_supported(const dv_conv_params* p, int pass, const char** why) {
  static const char* ok = "";
  *why = ok;
  if (p->dtype != DV_BF16) { *why = "tensor-core path is bf16 only"; return false; }
  if (p->y_f32 && pass != DV_CONV_FPROP) { *why = "fp32 output exists for fprop only"; return false; }
  if (p->y_f32 && p->K % 4) { *why = "fp32 output needs K % 4 == 0"; return false; }
  if (p->stride < 1 || p->stride > 8 || p->pad > 64 || p->R > 16 || p->S > 16) { *why = "stride/pad/filter out of TMA im2col range"; return false; }
  const int P = (p->H + 2 * p->pad - p->R) / p->stride + 1, Q = (p->W + 2 * p->pad - p->S) / p->stride + 1;
  if ((int64_t)p->N * p->H * p->W >= (1LL << 31) || (int64_t)p->N * P * Q >= (1LL << 31)) { *why = "pixel count exceeds int32"; return false; }
  // 2-D tiled maps zero-fill a partial last 64-channel block; the im2col maps are only used with whole blocks
  const bool plain = (p->R == 1 && p->S == 1 && p->pad == 0);
  if (pass == DV_CONV_FPROP) {
    if (p->C % 8 || p->C < 16) { *why = "C must be a multiple of 8 (and >= 16)"; return false; }
    if (!(plain && p->stride == 1) && p->C % 64) { *why = "C must be a multiple of 64 for RxS>1 or strided convolutions"; return false; }
    if (p->K % 8) { *why = "K must be a multiple of 8"; return false; }
    return true;
  }
  if (pass == DV_CONV_DGRAD) {
    if (p->K % 8 || p->K < 16) { *why = "K must be a multiple of 8 (and >= 16)"; return false; }
    if (p->C % 8) { *why = "C must be a multiple of 8"; return false; }
    if (!plain && p->K % 64) { *why = "K must be a multiple of 64 for RxS>1 dgrad"; return false; }
    if (p->stride == 1) {
      if (p->R - 1 - p->pad < 0 || p->S - 1 - p->pad < 0) { *why = "pad larger than filter-1"; return false; }
      return true;
    }
    if (p->R == 1 && p->S == 1 && p->pad == 0) return true;  // strided 1x1: GEMM + scatter
    if (p->R == 3 && p->S == 3 && p->pad == 1 && p->stride == 2) return true;  // four parity-class sub-convolutions
    *why = "strided dgrad with R,S>1 exists for 3x3 stride 2 pad 1 only";
    return false;
  }
  if (pass == DV_CONV_WGRAD) {
    if (p->C % 8 || p->C < 16) { *why = "C must be a multiple of 8 (and >= 16)"; return false; }
    if (!(plain && p->stride == 1) && p->C % 64) { *why = "C must be a multiple of 64 for RxS>1 or strided convolutions"; return false; }
    if (p->K % 8) { *why = "K must be a multiple of 8"; return false; }
    return true;
  }
  *why = "bad pass";
  return false;
}

size_t conv_tc_workspace(const dv_conv_params* p, int pass) {
  if (pass != DV_CONV_WGRAD || p->accumulate) return 0;
  WgradPlan w = plan_wgrad(p);
  return (size_t)w.splits * p->K * p->R * p->S * p->C * sizeof(float);
}

static int pick_bn(int nout) {
  static const int forced = env_int("DV_TC_BN", 0);  // experiment knob: cap the N tile
  int bn = nout <= 64 ? 64 : nout <= 128 ? 128 : 256;
  if (forced && bn > forced) bn = forced;
  return bn;
}

// the CTA-pair kernel takes the wide (BN = 256), dense-output launches with at least one full pair of M tiles per cluster
static bool use_pair(const GemmArgs& g, int bn) {
  static const int enabled = env_int("DV_TC_PAIR", 1);
  return enabled && bn == 256 && !g.scatter && !g.out_f32 && g.M >= 2 * kBlockM * (num_sms() / 2);
}

// STAGED (TMA store / TMA reduce-add epilogue, optional BN sums) whenever the output is a dense bf16 [M][Nout] matrix
static int gemm_dispatch(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmArgs& g, int bn, bool b_mn, cudaStream_t st,
                         const void* bnb_x0 = nullptr) {
  const bool staged = !g.scatter && !g.out_f32;
  alignas(64) CUtensorMap tmO;
  if (staged) {
    int rc = make_map_2d(&tmO, g.out, (uint64_t)g.M, (uint64_t)g.Nout, 32);
    if (rc) return rc;
  } else {
    tmO = tmA;  // unused by the kernel
  }
  alignas(64) CUtensorMap tmR = tmO;  // identity tensor of a folded-BN epilogue / x0 of the BN-backward sums: same geometry as the output
  if (g.ep_res != nullptr) {
    DV_REQUIRE(staged, DV_ERR_UNSUPPORTED, "conv_tc: a residual epilogue needs the TMA-store output path");
    int rc = make_map_2d(&tmR, g.ep_res, (uint64_t)g.M, (uint64_t)g.Nout, 32);
    if (rc) return rc;
  }
  const bool fold = g.ep_scale != nullptr;  // fprop only (conv_tc checks): staged, K-major B
  if (fold) {
    DV_REQUIRE(staged && !b_mn, DV_ERR_UNSUPPORTED, "conv_tc: the folded epilogue exists on the fprop TMA-store path only");
    if (g.ep_res == nullptr) {   // no identity tile to fetch (EPI 3): no tile buffers, the plain kernels' ring depth
      if (use_pair(g, bn)) return launch_gemm_pair<6, false, 3>(tmA, tmB, tmO, tmR, g, st);
      switch (bn) {
        case 64: return launch_gemm_persistent<64, 4, true, false, 3>(tmA, tmB, tmO, tmR, g, st);
        case 128: return launch_gemm_persistent<128, 3, true, false, 3>(tmA, tmB, tmO, tmR, g, st);
        default: return launch_gemm_persistent<256, 4, true, false, 3>(tmA, tmB, tmO, tmR, g, st);
      }
    }
    if (use_pair(g, bn)) return launch_gemm_pair<6, false, 1>(tmA, tmB, tmO, tmR, g, st);
    switch (bn) {
      case 64: return launch_gemm_persistent<64, 3, true, false, 1>(tmA, tmB, tmO, tmR, g, st);    // one stage fewer: two CTAs per SM still fit
      case 128: return launch_gemm_persistent<128, 2, true, false, 1>(tmA, tmB, tmO, tmR, g, st);
      default: return launch_gemm_persistent<256, 4, true, false, 1>(tmA, tmB, tmO, tmR, g, st);
    }
  }
  if (g.ep_res != nullptr) {   // dgrad + masked shortcut gradient (conv_tc checked the preconditions): the same epilogue without scale
    DV_REQUIRE(staged && b_mn && !g.accumulate && g.bnb_dgamma == nullptr, DV_ERR_UNSUPPORTED,
               "conv_tc: the residual epilogue of dgrad needs the plain TMA-store path (stride 1, no accumulate, KRSC filters, no bnb_*)");
    if (use_pair(g, bn)) return launch_gemm_pair<6, true, 1>(tmA, tmB, tmO, tmR, g, st);
    switch (bn) {
      case 64: return launch_gemm_persistent<64, 3, true, true, 1>(tmA, tmB, tmO, tmR, g, st);
      case 128: return launch_gemm_persistent<128, 2, true, true, 1>(tmA, tmB, tmO, tmR, g, st);
      default: return launch_gemm_persistent<256, 4, true, true, 1>(tmA, tmB, tmO, tmR, g, st);
    }
  }
  if (g.bnb_dgamma != nullptr) {   // dgrad with the producing layer's BatchNorm-backward sums (conv_tc checked the preconditions)
    DV_REQUIRE(staged && b_mn && !g.accumulate && bnb_x0 != nullptr, DV_ERR_UNSUPPORTED,
               "conv_tc: BatchNorm-backward sums need the plain TMA-store dgrad path (stride 1, no accumulate, KRSC filters)");
    int rc = make_map_2d(&tmR, bnb_x0, (uint64_t)g.M, (uint64_t)g.Nout, 32);
    if (rc) return rc;
    if (use_pair(g, bn)) return launch_gemm_pair<6, true, 2>(tmA, tmB, tmO, tmR, g, st);
    switch (bn) {
      case 64: return launch_gemm_persistent<64, 3, true, true, 2>(tmA, tmB, tmO, tmR, g, st);
      case 128: return launch_gemm_persistent<128, 2, true, true, 2>(tmA, tmB, tmO, tmR, g, st);
      default: return launch_gemm_persistent<256, 4, true, true, 2>(tmA, tmB, tmO, tmR, g, st);
    }
  }
  if (use_pair(g, bn)) return b_mn ? launch_gemm_pair<6, true, 0>(tmA, tmB, tmO, tmR, g, st) : launch_gemm_pair<6, false, 0>(tmA, tmB, tmO, tmR, g, st);
#define DV_GEMM_CASE(BN_, ST_, STG_, BMN_) return launch_gemm_persistent<BN_, ST_, STG_, BMN_, 0>(tmA, tmB, tmO, tmR, g, st)
  if (b_mn) {
    if (staged) { switch (bn) { case 64: DV_GEMM_CASE(64, 4, true, true); case 128: DV_GEMM_CASE(128, 3, true, true); default: DV_GEMM_CASE(256, 4, true, true); } }
    switch (bn) { case 64: DV_GEMM_CASE(64, 4, false, true); case 128: DV_GEMM_CASE(128, 3, false, true); default: DV_GEMM_CASE(256, 4, false, true); }
  }
  if (staged) { switch (bn) { case 64: DV_GEMM_CASE(64, 4, true, false); case 128: DV_GEMM_CASE(128, 3, true, false); default: DV_GEMM_CASE(256, 4, true, false); } }
  switch (bn) { case 64: DV_GEMM_CASE(64, 4, false, false); case 128: DV_GEMM_CASE(128, 3, false, false); default: DV_GEMM_CASE(256, 4, false, false); }
#undef DV_GEMM_CASE
}

bool conv3x3_halo_supported(const dv_conv_params* p, int pass);
int conv3x3_halo(const dv_conv_params* p, int pass, cudaStream_t st);
bool conv3x3_halo_wgrad_supported(const dv_conv_params* p);
int conv3x3_halo_wgrad(const dv_conv_params* p, cudaStream_t st);

int conv_tc(const dv_conv_params* p, int pass, cudaStream_t st) {
  const char* why;
  DV_REQUIRE(conv_tc_supported(p, pass, &why), DV_ERR_UNSUPPORTED, "conv_tc: %s", why);
  DV_REQUIRE(aligned16(p->x) && aligned16(p->y) && aligned16(p->w), DV_ERR_ARG, "conv_tc: pointers must be 16-byte aligned");
  if (conv3x3_halo_supported(p, pass)) return conv3x3_halo(p, pass, st);
  const int P = (p->H + 2 * p->pad - p->R) / p->stride + 1, Q = (p->W + 2 * p->pad - p->S) / p->stride + 1;
  alignas(64) CUtensorMap tmA, tmB;
  int rc;

  if (pass == DV_CONV_FPROP) {
    GemmArgs g{};
    g.out = (bf16*)p->y;
    g.M = p->N * P * Q;
    g.Nout = p->K;
    g.kb_per_tap = ceil_div(p->C, 64);
    g.Cin = p->C;
    g.taps = p->R * p->S;
    g.S = p->S;
    g.P = P; g.Q = Q; g.stride = p->stride; g.pad = p->pad;
    g.a_im2col = !(p->R == 1 && p->S == 1 && p->stride == 1 && p->pad == 0);
    g.out_f32 = p->y_f32 ? 1 : 0;
    g.stat_sums = p->y_f32 ? nullptr : p->stat_sums;
    if (p->ep_scale != nullptr || p->ep_shift != nullptr || p->ep_residual != nullptr || p->ep_relu) {
      DV_REQUIRE(p->ep_scale && p->ep_shift, DV_ERR_ARG, "conv_tc fprop: the folded epilogue needs both ep_scale and ep_shift");
      DV_REQUIRE(!p->y_f32 && p->stat_sums == nullptr, DV_ERR_UNSUPPORTED, "conv_tc fprop: folded epilogue with y_f32 or stat_sums");
      DV_REQUIRE(aligned16(p->ep_scale) && aligned16(p->ep_shift) && aligned16(p->ep_residual), DV_ERR_ARG,
                 "conv_tc fprop: ep_scale / ep_shift / ep_residual must be 16-byte aligned");
      g.ep_scale = p->ep_scale; g.ep_shift = p->ep_shift; g.ep_res = (const bf16*)p->ep_residual; g.ep_relu = p->ep_relu ? 1 : 0;
    }
    const int bn = pick_bn(p->K);
    if (g.a_im2col) rc = make_map_im2col(&tmA, p->x, p->N, p->H, p->W, p->C, p->R, p->S, p->stride, p->pad, kBlockM);
    else rc = make_map_2d(&tmA, p->x, (uint64_t)g.M, (uint64_t)p->C, kBlockM);
    if (rc) return rc;
    rc = make_map_2d(&tmB, p->w, (uint64_t)p->K, (uint64_t)p->R * p->S * p->C, use_pair(g, bn) ? 128u : (uint32_t)bn);
    if (rc) return rc;
    return gemm_dispatch(tmA, tmB, g, bn, false, st);
  }

  if (pass == DV_CONV_DGRAD) {
    DV_REQUIRE(p->w_t == nullptr || aligned16(p->w_t), DV_ERR_ARG, "conv_tc dgrad: w_t must be 16-byte aligned");
    const bool b_mn = (p->w_t == nullptr);  // no transposed copy given: read the KRSC filters MN-major, flip taps by index
    GemmArgs g{};
    g.out = (bf16*)p->x;
    g.accumulate = p->accumulate ? 1 : 0;
    g.Nout = p->C;
    g.kb_per_tap = ceil_div(p->K, 64);
    g.Cin = p->K;
    const int bn = pick_bn(p->C);
    const void* bnb_x0 = nullptr;
    if (p->bnb_x != nullptr) {
      DV_REQUIRE(p->stride == 1 && !p->accumulate && b_mn, DV_ERR_UNSUPPORTED,
                 "conv_tc dgrad: bnb_* (BatchNorm-backward sums) needs stride 1, accumulate == 0 and w_t == NULL");
      DV_REQUIRE(p->bnb_gamma && p->bnb_beta && p->bnb_mean && p->bnb_invstd && p->bnb_dgamma && p->bnb_dbeta, DV_ERR_ARG,
                 "conv_tc dgrad: bnb_x given but one of bnb_gamma/beta/mean/invstd/dgamma/dbeta is null");
      DV_REQUIRE(aligned16(p->bnb_x), DV_ERR_ARG, "conv_tc dgrad: bnb_x must be 16-byte aligned");
      g.bnb_gamma = p->bnb_gamma; g.bnb_beta = p->bnb_beta; g.bnb_mean = p->bnb_mean; g.bnb_invstd = p->bnb_invstd;
      g.bnb_dgamma = p->bnb_dgamma; g.bnb_dbeta = p->bnb_dbeta;
      bnb_x0 = p->bnb_x;
    }
    if (p->ep_residual != nullptr) {
      DV_REQUIRE(p->stride == 1 && !p->accumulate && b_mn && p->bnb_x == nullptr, DV_ERR_UNSUPPORTED,
                 "conv_tc dgrad: ep_residual needs stride 1, accumulate == 0, w_t == NULL and no bnb_*");
      DV_REQUIRE(!p->ep_scale && !p->ep_shift && !p->ep_relu, DV_ERR_ARG, "conv_tc dgrad: ep_scale / ep_shift / ep_relu are fprop fields");
      DV_REQUIRE(aligned16(p->ep_residual) && (p->ep_residual_mask == nullptr || ((uintptr_t)p->ep_residual_mask & 3) == 0) && p->C % 32 == 0,
                 DV_ERR_ARG, "conv_tc dgrad: ep_residual must be 16-byte and ep_residual_mask 4-byte aligned, C a multiple of 32");
      g.ep_res = (const bf16*)p->ep_residual;
      g.ep_mask = p->ep_residual_mask;
    }
    if (p->stride == 1) {
      // dx = conv(dy, w_t) with pad' = R-1-pad: an fprop whose "input" is dy [N,P,Q,K] and whose output is [N,H,W,C]
      const int padp = p->R - 1 - p->pad;
      g.M = p->N * p->H * p->W;
      g.taps = p->R * p->S;
      g.S = p->S;
      g.P = p->H; g.Q = p->W; g.stride = 1; g.pad = padp;
      g.a_im2col = !(p->R == 1 && p->S == 1);
      if (g.a_im2col) rc = make_map_im2col(&tmA, p->y, p->N, P, Q, p->K, p->R, p->S, 1, padp, kBlockM);
      else rc = make_map_2d(&tmA, p->y, (uint64_t)g.M, (uint64_t)p->K, kBlockM);
    } else if (p->R == 3) {
      // 3x3 stride 2 pad 1 (ResNet-34's down-sampling convs, resnet34.py:96-103): h = 2p + r - 1 ties the parity of h to the
      // filter row — even h only meets r = 1, odd h meets r = 2 (p = (h-1)/2) and r = 0 (p = (h+1)/2); the same along w.
      // So dx splits into four parity classes, each a STRIDE-1 convolution of dy with 1, 2, 2 or 4 of the nine taps: four
      // launches with exactly the useful FLOPs, the im2col map walking a 1x1 / 1x2 / 2x1 / 2x2 window over dy (one past the
      // edge reads as zero), rows scattered to (2i + a, 2j + b).  Every dx pixel belongs to one class: no memset.
      // (accumulate: the scatter epilogue reads-modifies-writes its own pixels, as for the strided 1x1 case)
      if (b_mn) rc = make_map_2d(&tmB, p->w, (uint64_t)p->K, (uint64_t)p->R * p->S * p->C, 64);
      else rc = make_map_2d(&tmB, p->w_t, (uint64_t)p->C, (uint64_t)p->R * p->S * p->K, (uint32_t)bn);
      if (rc) return rc;
      for (int a = 0; a < 2; ++a) {
        for (int b = 0; b < 2; ++b) {
          const int Hc = (p->H - a + 1) / 2, Wc = (p->W - b + 1) / 2;
          if (Hc <= 0 || Wc <= 0) continue;
          const int Ta = a ? 2 : 1, Tb = b ? 2 : 1;
          GemmArgs gc = g;
          gc.M = p->N * Hc * Wc;
          gc.taps = Ta * Tb;
          gc.S = Tb;
          gc.P = Hc; gc.Q = Wc; gc.stride = 1; gc.pad = 0;
          gc.a_im2col = 1;
          gc.scatter = 1; gc.Ho = p->H; gc.Wo = p->W; gc.os = 2; gc.oa = a; gc.ob = b;
          gc.use_btap = 1;
          for (int dp = 0; dp < Ta; ++dp) {
            for (int dq = 0; dq < Tb; ++dq) {
              const int r = a ? 2 - 2 * dp : 1, s2 = b ? 2 - 2 * dq : 1;
              const int fr = r * 3 + s2;
              gc.btap[dp * Tb + dq] = (signed char)(b_mn ? fr : 8 - fr);  // w_t holds the taps flipped
            }
          }
          const int corners[4] = {0, 0, Wc - Q, Hc - P};  // origins 0..Wc-1 / 0..Hc-1; the second tap may read column Q / row P = zero
          rc = make_map_im2col(&tmA, p->y, p->N, P, Q, p->K, Ta, Tb, 1, 0, kBlockM, corners);
          if (rc) return rc;
          rc = gemm_dispatch(tmA, tmB, gc, bn, b_mn, st);
          if (rc) return rc;
        }
      }
      return DV_OK;
    } else {
      // 1x1 stride s: dx[n, p*s, q*s, :] = dy[n,p,q,:] * w ; every other pixel of dx is zero
      if (!p->accumulate) DV_CUDA_OK(cudaMemsetAsync(p->x, 0, (size_t)p->N * p->H * p->W * p->C * sizeof(bf16), st));
      g.M = p->N * P * Q;
      g.taps = 1;
      g.S = 1;
      g.P = P; g.Q = Q; g.stride = 1; g.pad = 0;
      g.a_im2col = 0;
      g.scatter = 1; g.Ho = p->H; g.Wo = p->W; g.os = p->stride;
      rc = make_map_2d(&tmA, p->y, (uint64_t)g.M, (uint64_t)p->K, kBlockM);
    }
    if (rc) return rc;
    if (b_mn) rc = make_map_2d(&tmB, p->w, (uint64_t)p->K, (uint64_t)p->R * p->S * p->C, 64);
    else rc = make_map_2d(&tmB, p->w_t, (uint64_t)p->C, (uint64_t)p->R * p->S * p->K, use_pair(g, bn) ? 128u : (uint32_t)bn);
    if (rc) return rc;
    return gemm_dispatch(tmA, tmB, g, bn, b_mn, st, bnb_x0);
  }

  // wgrad
  if (conv3x3_halo_wgrad_supported(p)) return conv3x3_halo_wgrad(p, st);
  WgradPlan w = plan_wgrad(p);
  const size_t need = p->accumulate ? 0 : (size_t)w.splits * p->K * p->R * p->S * p->C * sizeof(float);
  DV_REQUIRE(need == 0 || (p->workspace != nullptr && p->workspace_bytes >= need), DV_ERR_WORKSPACE,
             "conv_tc wgrad: workspace of %zu bytes required, %zu given", need, p->workspace_bytes);
  DV_REQUIRE(need == 0 || aligned16(p->workspace), DV_ERR_ARG, "conv_tc wgrad: workspace must be 16-byte aligned");
  WgradArgs g{};
  g.ws = (float*)p->workspace;
  g.dw_atomic = p->accumulate ? (float*)p->w : nullptr;
  g.Mpix = p->N * P * Q;
  g.K = p->K; g.C = p->C; g.taps = p->R * p->S; g.S = p->S;
  g.P = P; g.Q = Q; g.stride = p->stride; g.pad = p->pad;
  g.x_im2col = !(p->R == 1 && p->S == 1 && p->stride == 1 && p->pad == 0);
  g.kb_per_split = w.kb_per_split;
  g.kb_total = w.kb_total;
  g.c_tiles = w.c_tiles;
  rc = make_map_2d(&tmA, p->y, (uint64_t)g.Mpix, (uint64_t)p->K, 64);
  if (rc) return rc;
  if (g.x_im2col) rc = make_map_im2col(&tmB, p->x, p->N, p->H, p->W, p->C, p->R, p->S, p->stride, p->pad, 64);
  else rc = make_map_2d(&tmB, p->x, (uint64_t)g.Mpix, (uint64_t)p->C, 64);
  if (rc) return rc;
  alignas(64) CUtensorMap tmW = tmA;   // only read by the kernel when g.tma_reduce is set
  static const int tma_reduce_enabled = env_int("DV_TC_WGRAD_TMA", 1);
  if (p->accumulate && tma_reduce_enabled && w.mh == 1 && aligned16(p->w) && (p->C % 4) == 0) {
    rc = make_map_dw(&tmW, (float*)p->w, p->K, p->R * p->S, p->C);
    if (rc) return rc;
    g.tma_reduce = 1;
  }
  if (w.mh == 1) {
    switch (w.bnc) {
      case 64: rc = launch_wgrad<64, 4, 1>(tmA, tmB, tmW, g, w.tiles, w.splits, st); break;
      case 128: rc = launch_wgrad<128, 3, 1>(tmA, tmB, tmW, g, w.tiles, w.splits, st); break;
      case 192: rc = launch_wgrad<192, 4, 1>(tmA, tmB, tmW, g, w.tiles, w.splits, st); break;
      default: rc = launch_wgrad<256, 4, 1>(tmA, tmB, tmW, g, w.tiles, w.splits, st); break;
    }
  } else {
    switch (w.bnc) {
      case 64: rc = launch_wgrad<64, 5, 2>(tmA, tmB, tmW, g, w.tiles, w.splits, st); break;
      case 128: rc = launch_wgrad<128, 4, 2>(tmA, tmB, tmW, g, w.tiles, w.splits, st); break;
      case 192: rc = launch_wgrad<192, 3, 2>(tmA, tmB, tmW, g, w.tiles, w.splits, st); break;
      default: rc = launch_wgrad<256, 3, 2>(tmA, tmB, tmW, g, w.tiles, w.splits, st); break;
    }
  }
  if (rc) return rc;
  if (p->accumulate) return DV_OK;  // the tiles were added into dw directly
  const int64_t n4 = (int64_t)p->K * p->R * p->S * p->C / 4;
  wgrad_reduce_kernel<<<(unsigned)ceil_div64(n4, 256), 256, 0, st>>>(g.ws, (float*)p->w, n4, w.splits);
  DV_LAUNCH_OK("wgrad_reduce_kernel");
  return DV_OK;
}

}  // namespace dv

// ============================================================================================================
// C <= 4 stems (resnet50.py:20-28 7x7/2, mobilenet_v1.py:30 3x3/2) without a patch matrix
// ============================================================================================================
// The image is kept as NHWC4 bf16 with its zero border written out ([N][Hp][Wp][4], 8 bytes per pixel, `pad` pixels on the
// left / top).  With stride 2 the window of output column q starts at padded pixel 2q = byte 16q of the row, so ONE tiled
// tensor map with OVERLAPPING strides describes every patch row:
//     dim0 = 32 elements (8 pixels x 4 channels: the S <= 7 taps of one filter row, the rest meets zero weights),
//     dim1 = Q window origins, 16 bytes apart,   dim2 = image rows,   dim3 = images.
// A box {32, Q, 1, 1} is the A operand of one filter row for one whole output row: [Q pixels][64 bytes], SWIZZLE_64B, K-major —
// no im2col buffer in HBM, no index arithmetic on the SM.  K = R blocks of 32.
//   fprop : D[128 px (Q used) x 64 k] += A_r [px x 32] * W_r [64 k x 32]      (filters resident in shared memory for the CTA's life)
//   wgrad : D[128 (64 k used) x R*32] += dY^T [k x 16 px] * [X_0 | ... | X_{R-1}] [16 px x R*32]    (one MMA, N = R*32, per 16 pixels)
namespace dv {
namespace {

constexpr int kStemWin = 32;                       // elements per filter row in the packed K layout (8 pixels x 4 channels)
constexpr int kStemABytes = kBlockM * kStemWin * 2;  // 8 KB: 128 rows x 64 B
constexpr int kStemMaxR = 7;

__device__ __forceinline__ void tma_load_4d(void* smem, const CUtensorMap* tm, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
          smem_u32(smem)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* tm, uint32_t smem_addr, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(reinterpret_cast<uint64_t>(tm)),
               "r"(smem_addr), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// SWIZZLE_64B descriptors: 8-row groups are 512 B apart (SBO); layout type 4
constexpr uint32_t kDescHi64 = (512u >> 4) | (1u << 14) | (4u << 29);
__device__ __forceinline__ uint64_t desc_pack_hi(uint32_t lo, uint32_t hi) {
  uint64_t d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
  return d;
}

struct StemArgs {
  int tiles;       // N * P output rows
  int P, Q, R;
  int Nout;        // output channels (<= 64)
  float* stat_sums;
  float* dw;       // wgrad, accumulate mode: packed fp32 [64][R*32], tiles are ADDED with red.global
  float* ws;       // wgrad, fixed-order mode: [grid][64][R*32] partials
};

template <int STAGES>
__global__ void __launch_bounds__(kThreads)
conv_stem_fprop_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmO,
                       const StemArgs g) {
  constexpr int BN = 64, ACC = 2;
  constexpr int kBRow = BN * kStemWin * 2;   // 4 KB: one filter row of all 64 output channels
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0) {
    if (threadIdx.x == 0) printf("dv conv_stem: dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }
  uint8_t* sB = smem + STAGES * kStemABytes;
  uint8_t* stage_out = sB + kStemMaxR * kBRow;           // 28 KB of filters, then 4 x [32 rows][128 B]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(stage_out + kEpiStageBytes);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + ACC;
  uint64_t* bfull_bar = tempty_bar + ACC;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bfull_bar + 1);
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;

  // rows Q..127 of every A stage are never written by TMA: zero them once, so that those accumulator rows (clipped by the store,
  // but summed into the BatchNorm statistics) are exactly zero
  for (int i = threadIdx.x * 16; i < STAGES * kStemABytes; i += kThreads * 16) {
    if ((i % kStemABytes) >= g.Q * 64) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmO);
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < ACC; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 4); }
    mbar_init(bfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_ptr, ACC * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    if (elect_one()) {   // the filters: R boxes of [64 k][32], once
      mbar_expect_tx(bfull_bar, (uint32_t)(g.R * kBRow));
      for (int r = 0; r < g.R; ++r) tma_load_2d(sB + r * kBRow, &tmB, bfull_bar, r * kStemWin, 0);
    }
    __syncwarp();
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < g.tiles; tile += gridDim.x) {
      const int n = tile / g.P, p = tile - n * g.P;
      for (int r = 0; r < g.R; ++r) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          mbar_expect_tx(&full_bar[stage], (uint32_t)(g.Q * 64));
          tma_load_4d(smem + stage * kStemABytes, &tmA, &full_bar[stage], 0, 0, 2 * p + r, n);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(kBlockM, BN, 0, 0);
    const uint32_t a_lo0 = desc_lo_kmajor(smem_u32(smem));
    const uint32_t b_lo0 = desc_lo_kmajor(smem_u32(sB));
    mbar_wait(bfull_bar, 0);
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < g.tiles; tile += gridDim.x) {
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
      for (int r = 0; r < g.R; ++r) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_lo = a_lo0 + (uint32_t)stage * (kStemABytes >> 4);
          const uint32_t b_lo = b_lo0 + (uint32_t)r * (kBRow >> 4);
          umma_f16(d_tmem, desc_pack_hi(a_lo, kDescHi64), desc_pack_hi(b_lo, kDescHi64), idesc, r ? 1u : 0u);
          umma_f16(d_tmem, desc_pack_hi(a_lo + 2u, kDescHi64), desc_pack_hi(b_lo + 2u, kDescHi64), idesc, 1u);   // +32 B: the second 16 of K
          umma_commit(&empty_bar[stage]);
          if (r == g.R - 1) umma_commit(&tfull_bar[acc]);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == ACC) { acc = 0; acc_phase ^= 1; }
    }
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    uint8_t* my_stage = stage_out + quarter * (32 * 128);
    const uint32_t my_stage_u32 = smem_u32(my_stage);
    int acc = 0;
    uint32_t acc_phase = 0;
    float s1[2] = {0.f, 0.f}, s2[2] = {0.f, 0.f};
    for (int tile = blockIdx.x; tile < g.tiles; tile += gridDim.x) {
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN);
      if (lane == 0) tma_store_wait_read();
      __syncwarp();
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t v[32];
        tmem_ld_32x32(t_addr + (uint32_t)(half * 32), v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[j * 8 + 2 * e]), __uint_as_float(v[j * 8 + 2 * e + 1]));
            w[e] = *reinterpret_cast<uint32_t*>(&h);
          }
          const int ch = (half * 4 + j) ^ (lane & 7);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_stage_u32 + lane * 128 + ch * 16), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                       "r"(w[3])
                       : "memory");
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&tempty_bar[acc]);                                   // every TMEM read of this warp is complete
        if (quarter * 32 < g.Q) tma_store_3d(&tmO, my_stage_u32, 0, quarter * 32, tile);   // rows past Q are clipped by the map
      }
      if (g.stat_sums != nullptr) {
        const int wch = lane >> 2, wof = (lane & 3) * 4;
        float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
#pragma unroll 8
        for (int r = 0; r < 32; ++r) {
          const uint32_t word = *reinterpret_cast<const uint32_t*>(my_stage + r * 128 + ((wch ^ (r & 7)) * 16) + wof);
          const float x0 = __uint_as_float(word << 16), x1 = __uint_as_float(word & 0xffff0000u);
          a0 += x0; a1 += x1;
          b0 = fmaf(x0, x0, b0); b1 = fmaf(x1, x1, b1);
        }
        s1[0] += a0; s1[1] += a1; s2[0] += b0; s2[1] += b1;
      }
      if (++acc == ACC) { acc = 0; acc_phase ^= 1; }
    }
    if (g.stat_sums != nullptr) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int col = 2 * lane + e;
        if (col < g.Nout) {
          atomicAdd(g.stat_sums + col, s1[e]);
          atomicAdd(g.stat_sums + g.Nout + col, s2[e]);
        }
      }
    }
    if (lane == 0) tma_store_wait_read();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, ACC * BN);
}


// Stem wgrad.  One stage = one output row (n, p): dY [Q px][64 k] (MN-major, 128 B per pixel) and the R patch rows [Q px][32]
// (64 B per pixel, SWIZZLE_64B).  Per 16 pixels ONE MMA: D[128 x R*32] += dY^T * [X_0 | ... | X_{R-1}]; the R blocks are 8 KB
// apart, which is the descriptor's stride between 32-column atoms.  Only 64 of the 128 accumulator rows carry filters: the
// second 64-channel chunk of the A operand is a zero buffer shared by all stages (reached through the descriptor's chunk
// stride), not a TMA fill.  A CTA accumulates all of its rows in TMEM and adds (or stores) its [64][R*32] tile once.
template <int STAGES>
__global__ void __launch_bounds__(kThreads)
conv_stem_wgrad_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX, const StemArgs g) {
  constexpr int kDyBytes = 16384;                         // [<=128 px][128 B]
  constexpr int kStageBytes = kDyBytes + kStemMaxR * kStemABytes;   // 72 KB
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0) {
    if (threadIdx.x == 0) printf("dv conv_stem: dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }
  uint8_t* zero_chunk = smem + STAGES * kStageBytes;      // 16 KB of zeros: channels 64..127 of dY
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(zero_chunk + kDyBytes);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int ncols = g.R * kStemWin;                        // 224 for the 7x7 stem
  for (int i = threadIdx.x * 16; i < kDyBytes; i += kThreads * 16) *reinterpret_cast<uint4*>(zero_chunk + i) = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmDY);
    prefetch_tmap(&tmX);
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_ptr, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const int my_tiles = (g.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // >= 1: the grid never exceeds the tile count

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < g.tiles; tile += gridDim.x) {
      const int n = tile / g.P, p = tile - n * g.P;
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (elect_one()) {
        uint8_t* sDY = smem + stage * kStageBytes;
        mbar_expect_tx(&full_bar[stage], (uint32_t)(g.Q * 128 + g.R * g.Q * 64));
        tma_load_2d(sDY, &tmDY, &full_bar[stage], 0, tile * g.Q);
        for (int r = 0; r < g.R; ++r) tma_load_4d(sDY + kDyBytes + r * kStemABytes, &tmX, &full_bar[stage], 0, 0, 2 * p + r, n);
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    const uint32_t idesc = make_idesc(128, ncols, 1, 1);
    constexpr uint32_t kHiA = (1024u >> 4) | (1u << 14) | (2u << 29);   // dY: SWIZZLE_128B, 8-pixel groups 1 KB apart
    constexpr uint32_t kHiB = (512u >> 4) | (1u << 14) | (4u << 29);    // X:  SWIZZLE_64B, 8-pixel groups 512 B apart
    const int ksteps = g.Q / 16;
    int stage = 0, it = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < g.tiles; tile += gridDim.x, ++it) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sDY = smem_u32(smem + stage * kStageBytes);
        const uint32_t lbo_a = (smem_u32(zero_chunk) - sDY) >> 4;       // "next 64 channels" = the zero chunk
        const uint32_t a_lo0 = ((sDY & 0x3FFFFu) >> 4) | (lbo_a << 16);
        const uint32_t b_lo0 = (((sDY + kDyBytes) & 0x3FFFFu) >> 4) | ((uint32_t)(kStemABytes >> 4) << 16);   // 32-column atoms 8 KB apart
        for (int k = 0; k < ksteps; ++k)
          umma_f16(tmem_base, desc_pack_hi(a_lo0 + (uint32_t)k * (2048u >> 4), kHiA), desc_pack_hi(b_lo0 + (uint32_t)k * (1024u >> 4), kHiB), idesc,
                   (it == 0 && k == 0) ? 0u : 1u);
        umma_commit(&empty_bar[stage]);
        if (it == my_tiles - 1) umma_commit(tmem_full_bar);
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else {
    const int quarter = warp & 3;
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    if (quarter < 2) {   // accumulator rows 0..63 = the 64 filters
      const int k = quarter * 32 + lane;
      float* orow = (g.dw != nullptr ? g.dw : g.ws + (size_t)blockIdx.x * 64 * ncols) + (size_t)k * ncols;
      for (int c = 0; c < ncols / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(c * 32), v);
        tmem_ld_wait();
        if (k < g.Nout) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (g.dw != nullptr)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(orow + c * 32 + j * 4), "f"(__uint_as_float(v[4 * j])),
                           "f"(__uint_as_float(v[4 * j + 1])), "f"(__uint_as_float(v[4 * j + 2])), "f"(__uint_as_float(v[4 * j + 3]))
                           : "memory");
            else
              *reinterpret_cast<uint4*>(orow + c * 32 + j * 4) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// x [N][H][W][C<=4] bf16 -> xp [N][Hp][Wp][4] bf16 with `pad` zero pixels on the left / top, zeros to the right / bottom and in
// the unused channels.  One thread per padded pixel (8-byte store).
__global__ void stem_pad_kernel(const bf16* __restrict__ x, bf16* __restrict__ xp, int N, int H, int W, int C, int Hp, int Wp, int pad) {
  const int64_t total = (int64_t)N * Hp * Wp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int wp = (int)(i % Wp);
    const int64_t t = i / Wp;
    const int hp = (int)(t % Hp), n = (int)(t / Hp);
    const int h = hp - pad, w = wp - pad;
    uint16_t v[4] = {0, 0, 0, 0};
    if ((unsigned)h < (unsigned)H && (unsigned)w < (unsigned)W) {
      const uint16_t* src = reinterpret_cast<const uint16_t*>(x) + (((size_t)n * H + h) * W + w) * C;
      for (int c = 0; c < C; ++c) v[c] = src[c];
    }
    reinterpret_cast<uint2*>(xp)[i] = make_uint2((uint32_t)v[0] | ((uint32_t)v[1] << 16), (uint32_t)v[2] | ((uint32_t)v[3] << 16));
  }
}
// The same from the batch as the reference's loader yields it (train.py:444-449: fp32 [N][C][H][W]): layout change, rounding to bf16
// and zero border in one pass — 12 bytes read and 8 written per pixel, where dv_nchw_to_nhwc + stem_pad_kernel move 12 + 6 + 6 + 8.
// Consecutive threads take consecutive padded columns, so each channel plane is read in whole lines.
__global__ void stem_pad_nchw_kernel(const float* __restrict__ x, bf16* __restrict__ xp, int N, int H, int W, int C, int Hp, int Wp, int pad) {
  const int64_t total = (int64_t)N * Hp * Wp;
  const size_t plane = (size_t)H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int wp = (int)(i % Wp);
    const int64_t t = i / Wp;
    const int hp = (int)(t % Hp), n = (int)(t / Hp);
    const int h = hp - pad, w = wp - pad;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if ((unsigned)h < (unsigned)H && (unsigned)w < (unsigned)W) {
      const float* src = x + (size_t)n * C * plane + (size_t)h * W + w;
      for (int c = 0; c < C; ++c) v[c] = __ldg(src + c * plane);
    }
    __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]), hi = __floats2bfloat162_rn(v[2], v[3]);
    reinterpret_cast<uint2*>(xp)[i] = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
  }
}
// fp32 KRSC filters -> bf16 [64][R][32]: wp[k][r][s*4 + c] = w[k][r][s][c], zero elsewhere (rows k >= K, taps s >= S, channels c >= C)
__global__ void stem_pack_w_kernel(const float* __restrict__ w, bf16* __restrict__ wp, int K, int R, int S, int C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * R * kStemWin) return;
  const int e = i % kStemWin, r = (i / kStemWin) % R, k = i / (kStemWin * R);
  const int s = e >> 2, c = e & 3;
  const float v = (k < K && s < S && c < C) ? w[(((size_t)k * R + r) * S + s) * C + c] : 0.f;
  wp[i] = __float2bfloat16_rn(v);
}
// packed fp32 [splits][64][R*32] -> dW [K][R][S][C] (KRSC): sum over splits in a fixed order, drop the padding columns
__global__ void stem_unpack_dw_kernel(const float* __restrict__ packed, float* __restrict__ dw, int splits, int K, int R, int S, int C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K * R * S * C) return;
  const int c = i % C, s = (i / C) % S, r = (i / (C * S)) % R, k = i / (C * S * R);
  const size_t col = (size_t)k * R * kStemWin + r * kStemWin + s * 4 + c, stride = (size_t)64 * R * kStemWin;
  float acc = 0.f;
  for (int j = 0; j < splits; ++j) acc += packed[(size_t)j * stride + col];
  dw[i] = acc;
}

struct StemGeom {
  int P, Q, Hp, Wp;
};
static bool stem_geometry(const dv_stem_params* p, StemGeom* gm, const char** why) {
  *why = "";
  if (p->C < 1 || p->C > 4) { *why = "C must be 1..4"; return false; }
  if (p->stride != 2) { *why = "stride must be 2 (the window origin advances by 16 bytes)"; return false; }
  if (p->R < 1 || p->R > kStemMaxR || p->S < 1 || p->S > 7) { *why = "filter larger than 7x7"; return false; }
  if (p->K < 8 || p->K > 64 || p->K % 8) { *why = "K must be a multiple of 8, at most 64"; return false; }
  if (p->pad < 0 || p->pad > 7) { *why = "bad pad"; return false; }
  gm->P = (p->H + 2 * p->pad - p->R) / 2 + 1;
  gm->Q = (p->W + 2 * p->pad - p->S) / 2 + 1;
  if (gm->Q < 16 || gm->Q > 128 || gm->Q % 16) { *why = "output width must be a multiple of 16, at most 128"; return false; }
  gm->Wp = p->pad + p->W;
  if (gm->Wp < 2 * (gm->Q - 1) + 8) gm->Wp = 2 * (gm->Q - 1) + 8;
  gm->Wp += gm->Wp & 1;                                      // row pitch a multiple of 16 bytes
  gm->Hp = p->pad + p->H;
  if (gm->Hp < 2 * (gm->P - 1) + p->R) gm->Hp = 2 * (gm->P - 1) + p->R;
  if ((int64_t)p->N * gm->P >= (1LL << 31) / 128) { *why = "too many output rows"; return false; }
  return true;
}
// the sliding-window map over the padded image (see the section header)
static int make_map_stem_x(CUtensorMap* tm, const void* xp, int N, const StemGeom& gm) {
  const MapKey key{{4, (uint64_t)(uintptr_t)xp, (uint64_t)N, (uint64_t)gm.Hp, (uint64_t)gm.Wp, (uint64_t)gm.Q}};
  if (map_cache_get(key, tm)) return DV_OK;
  DriverApi& api = driver_api();
  DV_REQUIRE(api.ok, DV_ERR_CUDA, "conv_stem: cuTensorMapEncode* entry points not available from the driver");
  cuuint64_t dims[4] = {(cuuint64_t)kStemWin, (cuuint64_t)gm.Q, (cuuint64_t)gm.Hp, (cuuint64_t)N};
  cuuint64_t strides[3] = {16, (cuuint64_t)gm.Wp * 8, (cuuint64_t)gm.Hp * gm.Wp * 8};
  cuuint32_t box[4] = {(cuuint32_t)kStemWin, (cuuint32_t)gm.Q, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = api.tiled(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(xp), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DV_REQUIRE(r == CUDA_SUCCESS, DV_ERR_CUDA, "cuTensorMapEncodeTiled(stem image) failed (%d) N=%d Hp=%d Wp=%d Q=%d", (int)r, N, gm.Hp, gm.Wp, gm.Q);
  map_cache_put(key, *tm);
  return DV_OK;
}

}  // namespace
}  // namespace dv

using namespace dv;

extern "C" int dv_stem_supported(const dv_stem_params* p, int* Hp, int* Wp) {
  if (!p) return 0;
  StemGeom gm;
  const char* why;
  if (!stem_geometry(p, &gm, &why)) return 0;
  if (Hp) *Hp = gm.Hp;
  if (Wp) *Wp = gm.Wp;
  return 1;
}

extern "C" int dv_stem_pad_input(const dv_stem_params* p, const void* x_nhwc, void* stream) {
  DV_REQUIRE(p && x_nhwc && p->x, DV_ERR_ARG, "dv_stem_pad_input: null pointer");
  StemGeom gm;
  const char* why;
  DV_REQUIRE(stem_geometry(p, &gm, &why), DV_ERR_UNSUPPORTED, "dv_stem_pad_input: %s", why);
  DV_REQUIRE(aligned16(p->x), DV_ERR_ARG, "dv_stem_pad_input: the padded image must be 16-byte aligned");
  const int64_t total = (int64_t)p->N * gm.Hp * gm.Wp;
  const unsigned grid = (unsigned)(ceil_div64(total, 256) < (int64_t)num_sms() * 16 ? ceil_div64(total, 256) : (int64_t)num_sms() * 16);
  stem_pad_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x_nhwc, (bf16*)p->x, p->N, p->H, p->W, p->C, gm.Hp, gm.Wp, p->pad);
  DV_LAUNCH_OK("stem_pad_kernel");
  return DV_OK;
}

extern "C" int dv_stem_pad_input_nchw(const dv_stem_params* p, const float* x_nchw, void* stream) {
  DV_REQUIRE(p && x_nchw && p->x, DV_ERR_ARG, "dv_stem_pad_input_nchw: null pointer");
  StemGeom gm;
  const char* why;
  DV_REQUIRE(stem_geometry(p, &gm, &why), DV_ERR_UNSUPPORTED, "dv_stem_pad_input_nchw: %s", why);
  DV_REQUIRE(aligned16(p->x), DV_ERR_ARG, "dv_stem_pad_input_nchw: the padded image must be 16-byte aligned");
  const int64_t total = (int64_t)p->N * gm.Hp * gm.Wp;
  const unsigned grid = (unsigned)(ceil_div64(total, 256) < (int64_t)num_sms() * 16 ? ceil_div64(total, 256) : (int64_t)num_sms() * 16);
  stem_pad_nchw_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x_nchw, (bf16*)p->x, p->N, p->H, p->W, p->C, gm.Hp, gm.Wp, p->pad);
  DV_LAUNCH_OK("stem_pad_nchw_kernel");
  return DV_OK;
}

extern "C" int dv_stem_pack_filters(const dv_stem_params* p, const float* w_krsc, void* stream) {
  DV_REQUIRE(p && w_krsc && p->w, DV_ERR_ARG, "dv_stem_pack_filters: null pointer");
  StemGeom gm;
  const char* why;
  DV_REQUIRE(stem_geometry(p, &gm, &why), DV_ERR_UNSUPPORTED, "dv_stem_pack_filters: %s", why);
  const int n = 64 * p->R * kStemWin;
  stem_pack_w_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(w_krsc, (bf16*)p->w, p->K, p->R, p->S, p->C);
  DV_LAUNCH_OK("stem_pack_w_kernel");
  return DV_OK;
}

extern "C" int dv_stem_fprop(const dv_stem_params* p, void* stream) {
  DV_REQUIRE(p && p->x && p->w && p->y, DV_ERR_ARG, "dv_stem_fprop: null pointer");
  StemGeom gm;
  const char* why;
  DV_REQUIRE(stem_geometry(p, &gm, &why), DV_ERR_UNSUPPORTED, "dv_stem_fprop: %s", why);
  DV_REQUIRE(aligned16(p->x) && aligned16(p->w) && aligned16(p->y), DV_ERR_ARG, "dv_stem_fprop: pointers must be 16-byte aligned");
  alignas(64) CUtensorMap tmA, tmB, tmO;
  int rc = make_map_stem_x(&tmA, p->x, p->N, gm);
  if (rc) return rc;
  DriverApi& api = driver_api();
  {  // filters [64][R*32] bf16, box {32, 64}, SWIZZLE_64B
    cuuint64_t dims[2] = {(cuuint64_t)p->R * kStemWin, 64};
    cuuint64_t strides[1] = {(cuuint64_t)p->R * kStemWin * 2};
    cuuint32_t box[2] = {(cuuint32_t)kStemWin, 64}, estr[2] = {1, 1};
    CUresult r = api.tiled(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p->w), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DV_REQUIRE(r == CUDA_SUCCESS, DV_ERR_CUDA, "cuTensorMapEncodeTiled(stem filters) failed (%d)", (int)r);
  }
  {  // output [N*P][Q][K] bf16, box {64, 32, 1}: a warp's 32 pixels of one output row; pixels past Q and channels past K are clipped
    cuuint64_t dims[3] = {(cuuint64_t)p->K, (cuuint64_t)gm.Q, (cuuint64_t)p->N * gm.P};
    cuuint64_t strides[2] = {(cuuint64_t)p->K * 2, (cuuint64_t)gm.Q * p->K * 2};
    cuuint32_t box[3] = {64, 32, 1}, estr[3] = {1, 1, 1};
    CUresult r = api.tiled(&tmO, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, p->y, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DV_REQUIRE(r == CUDA_SUCCESS, DV_ERR_CUDA, "cuTensorMapEncodeTiled(stem output) failed (%d) K=%d Q=%d", (int)r, p->K, gm.Q);
  }
  StemArgs g{};
  g.tiles = p->N * gm.P; g.P = gm.P; g.Q = gm.Q; g.R = p->R; g.Nout = p->K; g.stat_sums = p->stat_sums;
  constexpr int STAGES = 8;
  constexpr int smem = STAGES * kStemABytes + kStemMaxR * 64 * kStemWin * 2 + kEpiStageBytes + 256;
  DV_CUDA_OK(cudaFuncSetAttribute(conv_stem_fprop_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  DV_CUDA_OK(cudaFuncSetAttribute(conv_stem_fprop_kernel<STAGES>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  const int slots = 2 * num_sms();
  conv_stem_fprop_kernel<STAGES><<<g.tiles < slots ? g.tiles : slots, kThreads, smem, (cudaStream_t)stream>>>(tmA, tmB, tmO, g);
  DV_LAUNCH_OK("conv_stem_fprop_kernel");
  return DV_OK;
}

extern "C" size_t dv_stem_workspace_bytes(const dv_stem_params* p) {
  StemGeom gm;
  const char* why;
  if (!p || !stem_geometry(p, &gm, &why)) return 0;
  const int tiles = p->N * gm.P;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  return (size_t)(p->accumulate ? 1 : grid) * 64 * p->R * kStemWin * sizeof(float);
}

extern "C" int dv_stem_wgrad(const dv_stem_params* p, void* stream) {
  DV_REQUIRE(p && p->x && p->y && p->dw, DV_ERR_ARG, "dv_stem_wgrad: null pointer");
  StemGeom gm;
  const char* why;
  DV_REQUIRE(stem_geometry(p, &gm, &why), DV_ERR_UNSUPPORTED, "dv_stem_wgrad: %s", why);
  const size_t need = dv_stem_workspace_bytes(p);
  DV_REQUIRE(p->workspace && p->workspace_bytes >= need && aligned16(p->workspace), DV_ERR_WORKSPACE,
             "dv_stem_wgrad: workspace of %zu bytes required, %zu given", need, p->workspace_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  alignas(64) CUtensorMap tmDY, tmX;
  int rc = make_map_2d(&tmDY, p->y, (uint64_t)p->N * gm.P * gm.Q, (uint64_t)p->K, (uint32_t)gm.Q);
  if (rc) return rc;
  rc = make_map_stem_x(&tmX, p->x, p->N, gm);
  if (rc) return rc;
  StemArgs g{};
  g.tiles = p->N * gm.P; g.P = gm.P; g.Q = gm.Q; g.R = p->R; g.Nout = p->K;
  const int grid = g.tiles < num_sms() ? g.tiles : num_sms();
  const int splits = p->accumulate ? 1 : grid;
  if (p->accumulate) {   // one packed tile that every CTA adds into; the caller's dw is then updated by the unpack pass below
    DV_CUDA_OK(cudaMemsetAsync(p->workspace, 0, (size_t)64 * p->R * kStemWin * sizeof(float), st));
    g.dw = (float*)p->workspace;
  } else {
    g.ws = (float*)p->workspace;
  }
  constexpr int STAGES = 2;
  constexpr int smem = STAGES * (16384 + kStemMaxR * kStemABytes) + 16384 + 256;
  DV_CUDA_OK(cudaFuncSetAttribute(conv_stem_wgrad_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  conv_stem_wgrad_kernel<STAGES><<<grid, kThreads, smem, st>>>(tmDY, tmX, g);
  DV_LAUNCH_OK("conv_stem_wgrad_kernel");
  const int n = p->K * p->R * p->S * p->C;
  stem_unpack_dw_kernel<<<ceil_div(n, 256), 256, 0, st>>>((const float*)p->workspace, p->dw, splits, p->K, p->R, p->S, p->C);
  DV_LAUNCH_OK("stem_unpack_dw_kernel");
  return DV_OK;
}

// ------------------------------------------------------------------------------------------------------------
// 3x3 stride-1 pad-1 convolution, 64 -> 64 channels, over a halo tile resident in shared memory (fprop and dgrad).
//
// The im2col kernels above fetch a [128 pixels][64 channels] tile from L2 once per filter tap: nine fetches of (almost) the
// same pixels, 1.39 GB of L2->SM traffic for a 103 MB input at 64 channels @56x56 (profiles/r02_ncu_full_3x3_64_56.csv).
// Here ONE tiled TMA box brings rows h0-1 .. h0+2, columns -1 .. 62 of the image ([4][64 pixels][64 channels] = 32 KB, everything
// outside the image zero-filled by the TMA unit: that is the padding) and the nine taps are nine START ADDRESSES into it: with
// the rows of the box 64 pixels apart, output pixel m = 64*row + col of the tile reads input pixel m + 64*r + s for tap (r, s),
// i.e. the K-major SWIZZLE_128B descriptor of the A operand starts (64*r + s) * 128 bytes into the tile.  The swizzle is a
// function of the shared-memory address, so a start on any 128-byte row is legal (tools/probes/umma_row_offset.cu).  Output
// columns >= W of a row are computed from wrapped-around pixels and never stored: the TMA store clips them.  The nine 64x64
// filter taps (72 KB) stay resident for the CTA's life.  W <= 62 (W + 2 columns must fit the 64-pixel box row).
// ------------------------------------------------------------------------------------------------------------
namespace dv {
namespace {

__device__ __forceinline__ void tma_store_4d(const CUtensorMap* tm, uint32_t smem_addr, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(reinterpret_cast<uint64_t>(tm)),
               "r"(smem_addr), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

struct HaloArgs {
  int N, H, W;
  int tiles_h;        // ceil(H / 2): a tile is two output rows
  int total;          // N * tiles_h
  float* stat_sums;   // fprop, optional: [2][64] fp32 (as GemmArgs::stat_sums)
};

constexpr int kHaloWP = 64;                              // pixels per box row
constexpr int kHaloABytes = 4 * kHaloWP * 128;           // 32 KB
constexpr int kHaloBBytes = 9 * 64 * 128;                // 72 KB
constexpr int kHaloStages = 3;

// B_MN: dgrad — the filter tap is taken flipped and MN-major straight from the KRSC filters (rows = output channels of the
// forward convolution = reduction, columns = input channels).  BatchNorm statistics of the output when stat_sums != NULL.  The
// dgrad WITH BatchNorm-backward sums (bnb_*) stays on the im2col kernel: a version of this kernel that also fetched the x0 tile
// measured 115-146 us against 111 us there (64 -> 64 @56, batch 256): the extra shared-memory traffic of its epilogue competes
// with the operand reads of an already shared-memory-bound N = 64 MMA stream.
constexpr int kHaloThreads = 64 + 8 * 32;   // TMA warp, MMA warp, eight epilogue warps
template <bool B_MN>
__global__ void __launch_bounds__(kHaloThreads)
conv3x3_halo_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmO,
                    const HaloArgs g) {
  constexpr int STAGES = kHaloStages, ACC = 2, BN = 64;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0) {
    if (threadIdx.x == 0) printf("dv conv3x3_halo: dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }
  uint8_t* sB = smem + STAGES * kHaloABytes;     // also what the last stage's wrapped-around rows (2 pixels) read: never stored
  uint8_t* stage_out = sB + kHaloBBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(stage_out + kEpiStageBytes);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + ACC;
  uint64_t* bfull_bar = tempty_bar + ACC;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bfull_bar + 1);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmX);
    prefetch_tmap(&tmW);
    prefetch_tmap(&tmO);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < ACC; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], 8);
    }
    mbar_init(bfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_ptr, ACC * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    if (elect_one()) {   // the nine filter taps, once
      mbar_expect_tx(bfull_bar, kHaloBBytes);
#pragma unroll
      for (int t = 0; t < 9; ++t) tma_load_2d(sB + t * 8192, &tmW, bfull_bar, t * 64, 0);
    }
    __syncwarp();
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < g.total; tile += gridDim.x) {
      const int n = tile / g.tiles_h, h0 = (tile - n * g.tiles_h) * 2;
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (elect_one()) {
        mbar_expect_tx(&full_bar[stage], kHaloABytes);
        tma_load_4d(smem + stage * kHaloABytes, &tmX, &full_bar[stage], 0, -1, h0 - 1, n);
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(kBlockM, BN, 0, B_MN ? 1 : 0);
    constexpr uint32_t kBStep = B_MN ? (2048u >> 4) : 2u;
    const uint32_t a_lo0 = desc_lo_kmajor(smem_u32(smem));
    const uint32_t b_lo0 = B_MN ? desc_lo_mnmajor(smem_u32(sB)) : desc_lo_kmajor(smem_u32(sB));
    mbar_wait(bfull_bar, 0);
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < g.total; tile += gridDim.x) {
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        const uint32_t a_st = a_lo0 + (uint32_t)stage * (uint32_t)(kHaloABytes >> 4);
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          const int r = t / 3, s = t - 3 * r;
          const uint32_t a_lo = a_st + (uint32_t)((r * kHaloWP + s) * 128 >> 4);
          const uint32_t b_lo = b_lo0 + (uint32_t)((B_MN ? 8 - t : t) * (8192 >> 4));
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16(d_tmem, desc_pack(a_lo + k * 2u), desc_pack(b_lo + k * kBStep), idesc, (t | k) ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);
        umma_commit(&tfull_bar[acc]);
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
      if (++acc == ACC) { acc = 0; acc_phase ^= 1; }
    }
    __syncwarp();
  } else {
    // Eight epilogue warps.  Warp (quarter, half) owns accumulator rows 32*quarter .. +31 — output row h0 + (quarter >> 1), image
    // columns 32*(quarter & 1) .. +31 — and output channels 32*half .. +31: one tcgen05.ld, a [32 pixels][32 channels] staging tile
    // (64-byte rows, SWIZZLE_64B: 16-byte chunk index XOR (row >> 1) & 3), one TMA store, column sums from the staged bf16 values.
    const int quarter = warp & 3, ew = warp - 2, col0 = (ew >> 2) * 32;
    const int rho = quarter >> 1, kap0 = (quarter & 1) * 32;
    uint8_t* my_stage = stage_out + ew * (32 * 64);
    const uint32_t my_stage_u32 = smem_u32(my_stage);
    const int nvalid_w = g.W - kap0 < 0 ? 0 : (g.W - kap0 > 32 ? 32 : g.W - kap0);   // columns of this warp inside the image
    // column sums: lane (rh, cp) = (lane >> 4, lane & 15) walks the even (rh = 0) or odd rows of channel pair 2*cp, 2*cp+1 (one 32-bit word per row)
    const int rh = lane >> 4, cp = lane & 15;
    const int my_off = ((cp >> 2) * 16) + (cp & 3) * 4;   // byte offset of that channel pair in an unswizzled 64-byte row
    int acc = 0;
    uint32_t acc_phase = 0;
    float2 s1 = make_float2(0.f, 0.f), s2 = make_float2(0.f, 0.f);   // running sums of the lane's channel pair and row half
    for (int tile = blockIdx.x; tile < g.total; tile += gridDim.x) {
      const int n = tile / g.tiles_h, h = (tile - n * g.tiles_h) * 2 + rho;
      const int nvalid = h < g.H ? nvalid_w : 0;   // rows of this warp's 32 that are pixels of the image
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN + col0);
      uint32_t v[32];
      tmem_ld_32x32(t_addr, v);
      tmem_ld_wait();
      // the accumulator columns are in registers: hand them back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&tempty_bar[acc]);
        tma_store_wait_read();   // the previous tile's store has read the staging buffer
      }
      if (++acc == ACC) { acc = 0; acc_phase ^= 1; }
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          __nv_bfloat162 hh = __floats2bfloat162_rn(__uint_as_float(v[j * 8 + 2 * e]), __uint_as_float(v[j * 8 + 2 * e + 1]));
          w[e] = *reinterpret_cast<uint32_t*>(&hh);
        }
        const int ch = j ^ ((lane >> 1) & 3);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_stage_u32 + lane * 64 + ch * 16), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                     "r"(w[3])
                     : "memory");
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0 && nvalid > 0) tma_store_4d(&tmO, my_stage_u32, col0, kap0, h, n);   // columns >= W of the box are clipped
      if (nvalid > 0 && g.stat_sums != nullptr) {
        float2 a = make_float2(0.f, 0.f), b = make_float2(0.f, 0.f);
#pragma unroll
        for (int i = 0; i < 16; ++i) {   // rows outside the image are skipped
          const int r = 2 * i + rh;   // even / odd rows: the two half-warps read different banks
          if (r < nvalid) {
            const uint32_t word = *reinterpret_cast<const uint32_t*>(my_stage + r * 64 + (my_off ^ (((r >> 1) & 3) << 4)));
            const float2 d = make_float2(__uint_as_float(word << 16), __uint_as_float(word & 0xffff0000u));
            a = __fadd2_rn(a, d);
            b = __ffma2_rn(d, d, b);
          }
        }
        s1 = __fadd2_rn(s1, a);
        s2 = __fadd2_rn(s2, b);
      }
    }
    if (g.stat_sums != nullptr) {   // the two row halves of a channel pair meet, then one atomic per channel
      s1.x += __shfl_xor_sync(0xffffffffu, s1.x, 16); s1.y += __shfl_xor_sync(0xffffffffu, s1.y, 16);
      s2.x += __shfl_xor_sync(0xffffffffu, s2.x, 16); s2.y += __shfl_xor_sync(0xffffffffu, s2.y, 16);
      if (rh == 0) {
        const int col = col0 + 2 * cp;
        atomicAdd(g.stat_sums + col, s1.x);
        atomicAdd(g.stat_sums + col + 1, s1.y);
        atomicAdd(g.stat_sums + 64 + col, s2.x);
        atomicAdd(g.stat_sums + 64 + col + 1, s2.y);
      }
    }
    if (lane == 0) tma_store_wait_read();   // shared memory must outlive the last bulk store's reads
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, ACC * BN);
}

// Weight gradient of the same layer over the same halo tile: dW[k][r][s][c] = sum over pixels of dy[pix][k] * x[pix + (r, s)][c].
// Both operands are MN-major (a shared-memory row = one pixel = one K index, 128 bytes = 64 channels), so the shifted windows of x are
// again start addresses, now along K.  The accumulator is kept TRANSPOSED — D[c][k], A = x, B = dy — because then M can be filled:
// M = 128 is two 64-channel chunks of A whose distance is the descriptor's leading-dimension byte offset, and with that offset set to
// (offset of tap b - offset of tap a) * 128 bytes ONE MMA produces two taps (rows 0-63: tap a, rows 64-127: tap b; any distance and
// any start row are legal, tools/probes/umma_row_offset.cu).  The im2col kernel spends an M = 128 MMA per tap with half of its rows
// on zero-filled output channels and fetches the dy tile and a shifted x tile per tap: 2.77 GB of L2->SM traffic per launch.
// Here: 5 MMAs per 16 pixels (taps 0+1, 2+3, 4+5, 6+7, 8), one x box + one dy box per two output rows (columns >= W and rows >= H of
// dy are zero-filled by the TMA unit, so the wrapped-around pixels of x meet zeros), five accumulators (320 TMEM columns) that live
// for the CTA's whole share of the image, and at the end 18 fp32 [32 k][32 c] tiles per warp added into dW by TMA reduce-add.
constexpr int kHaloWgStages = 4;
constexpr int kHaloDyBytes = 2 * kHaloWP * 128;   // 16 KB: two output rows of dy

__global__ void __launch_bounds__(kThreads)
conv3x3_halo_wgrad_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmDW,
                          const HaloArgs g) {
  constexpr int STAGES = kHaloWgStages, kStage = kHaloABytes + kHaloDyBytes;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0) {
    if (threadIdx.x == 0) printf("dv conv3x3_halo_wgrad: dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * kStage);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tfull_bar + 1);
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmX);
    prefetch_tmap(&tmDY);
    prefetch_tmap(&tmDW);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(tfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_ptr, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < g.total; tile += gridDim.x) {
      const int n = tile / g.tiles_h, h0 = (tile - n * g.tiles_h) * 2;
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (elect_one()) {
        uint8_t* sX = smem + stage * kStage;
        mbar_expect_tx(&full_bar[stage], kStage);
        tma_load_4d(sX, &tmX, &full_bar[stage], 0, -1, h0 - 1, n);
        tma_load_4d(sX + kHaloABytes, &tmDY, &full_bar[stage], 0, 0, h0, n);
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(128, 64, 1, 1);
    int stage = 0;
    uint32_t phase = 0;
    bool first = true;
    for (int tile = blockIdx.x; tile < g.total; tile += gridDim.x) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t x0 = (smem_u32(smem + stage * kStage) & 0x3FFFFu) >> 4;
        const uint32_t b_lo0 = desc_lo_mnmajor(smem_u32(smem + stage * kStage + kHaloABytes));
#pragma unroll
        for (int j = 0; j < 5; ++j) {
          const int ta = 2 * j, tb = j < 4 ? 2 * j + 1 : 2 * j;                       // taps of rows 0-63 / 64-127 (the last one twice)
          const int offa = (ta / 3) * kHaloWP + ta % 3, offb = (tb / 3) * kHaloWP + tb % 3;
          const uint32_t lbo = (uint32_t)((offb - offa) * 128) >> 4;                 // chunk distance, 16-byte units
          const uint32_t a_lo0 = (x0 + (uint32_t)(offa * 128 >> 4)) | (lbo << 16);
#pragma unroll
          for (int k = 0; k < 8; ++k)   // 128 pixels of the tile, 16 per MMA (+2048 bytes)
            umma_f16(tmem_base + (uint32_t)(j * 64), desc_pack(a_lo0 + (uint32_t)(k * 128)), desc_pack(b_lo0 + (uint32_t)(k * 128)), idesc,
                     (first && k == 0) ? 0u : 1u);
        }
        umma_commit(&empty_bar[stage]);
      }
      __syncwarp();
      first = false;
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
    if (elect_one()) umma_commit(tfull_bar);
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    mbar_wait(tfull_bar, 0);
    tc_fence_after();
    // every MMA has completed: the operand ring is free for the fp32 tiles ([32 k][32 c], 128-byte rows, SWIZZLE_128B as tmDW wants)
    uint8_t* tiles = smem + (size_t)quarter * (10 * 4096);
    const int c0 = (quarter & 1) * 32;
#pragma unroll 1
    for (int j = 0; j < 5; ++j) {
      const int tap = quarter < 2 ? 2 * j : 2 * j + 1;
      if (tap > 8) break;   // the fifth accumulator holds tap 8 twice
#pragma unroll 1
      for (int kh = 0; kh < 2; ++kh) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(j * 64 + kh * 32), v);
        tmem_ld_wait();
        uint8_t* tile = tiles + (j * 2 + kh) * 4096;
        // lane = channel c0 + lane holds 32 output channels k: transposed into rows k
#pragma unroll
        for (int kk = 0; kk < 32; ++kk)
          *reinterpret_cast<uint32_t*>(tile + kk * 128 + (((lane >> 2) ^ (kk & 7)) * 16) + (lane & 3) * 4) = v[kk];
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) tma_reduce_add_3d(&tmDW, smem_u32(tile), c0, tap, kh * 32);
      }
    }
    if (lane == 0) tma_store_wait_read();
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace

// fprop or stride-1 dgrad of a 3x3 pad-1 convolution with 64 input and 64 output channels whose rows fit the 64-pixel box
bool conv3x3_halo_supported(const dv_conv_params* p, int pass) {
  static const int enabled = [] { const char* v = getenv("DV_TC_HALO"); return v ? atoi(v) : 1; }();
  if (!enabled || (pass != DV_CONV_FPROP && pass != DV_CONV_DGRAD)) return false;
  if (p->dtype != DV_BF16 || p->R != 3 || p->S != 3 || p->stride != 1 || p->pad != 1 || p->C != 64 || p->K != 64) return false;
  if (p->W > kHaloWP - 2 || p->W < 24 || p->H < 2) return false;   // narrower rows waste most of the 64-pixel box
  if (p->y_f32 || p->accumulate || p->ep_scale || p->ep_shift || p->ep_residual || p->ep_relu) return false;
  if (pass == DV_CONV_DGRAD && p->w_t != nullptr) return false;     // KRSC filters, read MN-major
  if (p->bnb_x != nullptr) return false;
  return true;
}

int conv3x3_halo(const dv_conv_params* p, int pass, cudaStream_t st) {
  const bool dgrad = pass == DV_CONV_DGRAD;
  const void* in = dgrad ? p->y : p->x;     // dy [N,H,W,64] (stride 1, pad 1: P = H, Q = W)  /  x
  void* out = dgrad ? p->x : p->y;          // dx / y
  alignas(64) CUtensorMap tmX, tmW, tmO;
  const uint64_t dims[4] = {64, (uint64_t)p->W, (uint64_t)p->H, (uint64_t)p->N};
  const uint64_t strides[3] = {128, (uint64_t)p->W * 128, (uint64_t)p->H * p->W * 128};
  const uint32_t box_in[4] = {64, (uint32_t)kHaloWP, 4, 1}, box_out[4] = {32, 32, 1, 1};   // out: one epilogue warp's [32 px][32 ch], SWIZZLE_64B
  int rc = make_tiled_map_bf16(&tmX, in, 4, dims, strides, box_in, 3);
  if (rc) return rc;
  rc = make_tiled_map_bf16(&tmO, out, 4, dims, strides, box_out, 2);
  if (rc) return rc;
  rc = make_map_2d(&tmW, p->w, 64, (uint64_t)9 * 64, 64);   // KRSC [64][9*64]: one [64][64] box per tap, either way round
  if (rc) return rc;
  HaloArgs g{};
  g.N = p->N; g.H = p->H; g.W = p->W;
  g.tiles_h = (p->H + 1) / 2;
  g.total = p->N * g.tiles_h;
  g.stat_sums = dgrad ? nullptr : p->stat_sums;
  const int grid = g.total < num_sms() ? g.total : num_sms();
  constexpr int smem = kHaloStages * kHaloABytes + kHaloBBytes + kEpiStageBytes + 256;
  static_assert(smem <= 227 * 1024, "halo kernel exceeds shared memory");
  if (dgrad) {
    DV_CUDA_OK(cudaFuncSetAttribute(conv3x3_halo_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    conv3x3_halo_kernel<true><<<grid, kHaloThreads, smem, st>>>(tmX, tmW, tmO, g);
  } else {
    DV_CUDA_OK(cudaFuncSetAttribute(conv3x3_halo_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    conv3x3_halo_kernel<false><<<grid, kHaloThreads, smem, st>>>(tmX, tmW, tmO, g);
  }
  DV_LAUNCH_OK("conv3x3_halo_kernel");
  return DV_OK;
}


// weight gradient of the same geometry, added into dW (fp32 KRSC) by TMA reduce-add: the accumulate mode of dv_conv_wgrad
bool conv3x3_halo_wgrad_supported(const dv_conv_params* p) {
  static const int enabled = [] { const char* v = getenv("DV_TC_HALO_WGRAD"); return v ? atoi(v) : 1; }();
  if (!enabled || p->dtype != DV_BF16 || p->R != 3 || p->S != 3 || p->stride != 1 || p->pad != 1 || p->C != 64 || p->K != 64) return false;
  if (p->W > kHaloWP - 2 || p->W < 24 || p->H < 2) return false;
  if (!p->accumulate || p->y_f32 || p->ep_scale || p->ep_shift || p->ep_residual || p->ep_relu || p->bnb_x) return false;
  return aligned16(p->w);
}

int conv3x3_halo_wgrad(const dv_conv_params* p, cudaStream_t st) {
  alignas(64) CUtensorMap tmX, tmDY, tmDW;
  const uint64_t dims[4] = {64, (uint64_t)p->W, (uint64_t)p->H, (uint64_t)p->N};
  const uint64_t strides[3] = {128, (uint64_t)p->W * 128, (uint64_t)p->H * p->W * 128};
  const uint32_t box_x[4] = {64, (uint32_t)kHaloWP, 4, 1}, box_dy[4] = {64, (uint32_t)kHaloWP, 2, 1};
  int rc = make_tiled_map_bf16(&tmX, p->x, 4, dims, strides, box_x, 3);
  if (rc) return rc;
  rc = make_tiled_map_bf16(&tmDY, p->y, 4, dims, strides, box_dy, 3);
  if (rc) return rc;
  rc = make_map_dw(&tmDW, (float*)p->w, 64, 9, 64);
  if (rc) return rc;
  HaloArgs g{};
  g.N = p->N; g.H = p->H; g.W = p->W;
  g.tiles_h = (p->H + 1) / 2;
  g.total = p->N * g.tiles_h;
  const int grid = g.total < num_sms() ? g.total : num_sms();
  constexpr int smem = kHaloWgStages * (kHaloABytes + kHaloDyBytes) + 256;
  static_assert(smem <= 227 * 1024, "halo wgrad kernel exceeds shared memory");
  DV_CUDA_OK(cudaFuncSetAttribute(conv3x3_halo_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  conv3x3_halo_wgrad_kernel<<<grid, kThreads, smem, st>>>(tmX, tmDY, tmDW, g);
  DV_LAUNCH_OK("conv3x3_halo_wgrad_kernel");
  return DV_OK;
}

}  // namespace dv
