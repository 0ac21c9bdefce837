// augment.cu — the reference's per-image training / validation transforms after decode + rescale, as one GPU pass over a
// uint8 batch: RandomHorizontalFlip -> RandomCrop (or CenterCrop) -> ColorJitter(brightness, contrast, saturation in a
// drawn order; hue 0) -> ToTensor (float, no /255) -> Normalize   (ResNet/pytorch/data_load.py:102-297 as composed at
// train.py:315-331).  The arithmetic is Pillow's (ImageEnhance = Image.blend against a degenerate image, convert("L"))
// and torchvision's normalize, restated so that the uint8 stages are bit-exact:
//   blend(deg, img, a) = trunc(float(deg) + a * float(img - deg)), clipped to [0,255] first when a is outside [0,1]
//   L = (19595 R + 38470 G + 7471 B + 0x8000) >> 16;  contrast's degenerate = int(mean(L) + 0.5) of the image as it is
//   when the contrast step runs — hence a first pass that sums L per image, then the pass that writes the tensor.
#include "common.cuh"

namespace dv {
namespace {

struct AugArgs {
  const uint8_t* src;
  const int32_t* geom;
  const float* factors;
  const int32_t* widths;
  uint32_t* gray_sums;
  int B, Hs, Ws, crop;
  float mean[3], stdv[3];
};

__device__ __forceinline__ int blend(int deg, int img, float a) {
  float v = __fadd_rn((float)deg, __fmul_rn(a, (float)(img - deg)));  // no FMA: Pillow's C code rounds the product
  if (!(a >= 0.f && a <= 1.f)) v = fminf(fmaxf(v, 0.f), 255.f);
  return (int)v;
}
__device__ __forceinline__ int gray_l(const int* c) { return (c[0] * 19595 + c[1] * 38470 + c[2] * 7471 + 0x8000) >> 16; }

// the three operations in the order `order` names (index into the 6 permutations of brightness, contrast, saturation);
// `upto_contrast`: stop in front of the contrast step (first pass).  Returns false if the contrast step was not reached.
__device__ __forceinline__ void jitter(int* c, int order, const float* f, int contrast_mean, bool upto_contrast) {
  constexpr int kOrders[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 0, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int op = kOrders[order][i];
    if (op == 0) {
#pragma unroll
      for (int k = 0; k < 3; ++k) c[k] = blend(0, c[k], f[0]);
    } else if (op == 1) {
      if (upto_contrast) return;
#pragma unroll
      for (int k = 0; k < 3; ++k) c[k] = blend(contrast_mean, c[k], f[1]);
    } else {
      const int g = gray_l(c);
#pragma unroll
      for (int k = 0; k < 3; ++k) c[k] = blend(g, c[k], f[2]);
    }
  }
}

__device__ __forceinline__ void load_pixel(const AugArgs& a, int b, int y, int x, int* c) {
  const int top = a.geom[b * 4 + 0], left = a.geom[b * 4 + 1], flip = a.geom[b * 4 + 2];
  const int w = a.widths ? a.widths[b] : a.Ws;
  const int sx = flip ? (w - 1 - (left + x)) : (left + x);
  const uint8_t* p = a.src + (((size_t)b * a.Hs + (top + y)) * a.Ws + sx) * 3;
  c[0] = p[0]; c[1] = p[1]; c[2] = p[2];
}

__global__ void __launch_bounds__(256) augment_gray_sum_kernel(AugArgs a) {
  const int b = blockIdx.y;
  const int order = a.geom[b * 4 + 3];
  if (order < 0) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t l = 0;
  if (i < a.crop * a.crop) {
    int c[3];
    load_pixel(a, b, i / a.crop, i % a.crop, c);
    jitter(c, order, a.factors + b * 3, 0, true);
    l = (uint32_t)gray_l(c);
  }
  __shared__ uint32_t red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = l;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t s = 0;
    for (int k = 0; k < 8; ++k) s += red[k];
    atomicAdd(a.gray_sums + b, s);   // integer sum: exact in any order
  }
}

template <typename T, bool NCHW>
__global__ void __launch_bounds__(256) augment_apply_kernel(AugArgs a, T* __restrict__ dst) {
  const int b = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = a.crop * a.crop;
  if (i >= n) return;
  int c[3];
  load_pixel(a, b, i / a.crop, i % a.crop, c);
  const int order = a.geom[b * 4 + 3];
  if (order >= 0) {
    const int mean = (int)((double)a.gray_sums[b] / (double)n + 0.5);
    jitter(c, order, a.factors + b * 3, mean, false);
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float v = __fdiv_rn(__fsub_rn((float)c[k], a.mean[k]), a.stdv[k]);   // (x - mean) / std on 0..255 values: no /255, as the reference
    if (NCHW) dst[((size_t)b * 3 + k) * n + i] = from_f32<T>(v);
    else dst[((size_t)b * n + i) * 3 + k] = from_f32<T>(v);
  }
}

// ---- Rescale: cv2.resize(image, (new_w, new_h)) with INTER_LINEAR on uint8 (data_load.py:72-99), bit for bit ----------------
// OpenCV's 8-bit path is fixed point: per axis two taps with weights cvRound(w * 2048) from a float32 fraction of the aligned
// pixel centre; a horizontal pass into 32-bit sums, then (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2.
// Along x the fraction is zeroed where the window leaves the image, along y the row indices are clamped instead.
struct RescaleArgs {
  const uint8_t* src;
  const int64_t* src_off;
  const int32_t* src_hw;
  const int32_t* dst_hw;
  uint8_t* dst;
  int B, Hs, Ws;
};

__device__ __forceinline__ void resize_tap(int d, int dsize, int ssize, bool vertical, int& i0, int& i1, int& w0, int& w1) {
  const double scale = 1.0 / ((double)dsize / (double)ssize);
  float f = (float)(((double)d + 0.5) * scale - 0.5);
  int s0 = (int)floorf(f);
  float fr = f - (float)s0;
  if (!vertical) {
    if (s0 < 0) { fr = 0.f; s0 = 0; }
    if (s0 >= ssize - 1) { fr = 0.f; s0 = ssize - 1; }
  }
  w1 = __float2int_rn(__fmul_rn(fr, 2048.f));
  w0 = __float2int_rn(__fmul_rn(__fsub_rn(1.f, fr), 2048.f));
  i0 = min(max(s0, 0), ssize - 1);
  i1 = min(max(s0 + 1, 0), ssize - 1);
}

__global__ void __launch_bounds__(256) rescale_kernel(RescaleArgs a) {
  const int b = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.Hs * a.Ws) return;
  const int y = i / a.Ws, x = i - y * a.Ws;
  const int sh = a.src_hw[2 * b], sw = a.src_hw[2 * b + 1], dh = a.dst_hw[2 * b], dw = a.dst_hw[2 * b + 1];
  uint8_t* out = a.dst + (((size_t)b * a.Hs + y) * a.Ws + x) * 3;
  if (y >= dh || x >= dw) {  // the rest of the slot: defined, never read by dv_augment
    out[0] = out[1] = out[2] = 0;
    return;
  }
  int x0, x1, ax0, ax1, y0, y1, ay0, ay1;
  resize_tap(x, dw, sw, false, x0, x1, ax0, ax1);
  resize_tap(y, dh, sh, true, y0, y1, ay0, ay1);
  const uint8_t* img = a.src + a.src_off[b];
  const uint8_t* r0 = img + (size_t)y0 * sw * 3;
  const uint8_t* r1 = img + (size_t)y1 * sw * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int h0 = r0[x0 * 3 + c] * ax0 + r0[x1 * 3 + c] * ax1;
    const int h1 = r1[x0 * 3 + c] * ax0 + r1[x1 * 3 + c] * ax1;
    out[c] = (uint8_t)((((ay0 * (h0 >> 4)) >> 16) + ((ay1 * (h1 >> 4)) >> 16) + 2) >> 2);
  }
}

}  // namespace
}  // namespace dv

using namespace dv;

extern "C" int dv_rescale(const dv_rescale_params* p, void* stream) {
  DV_REQUIRE(p && p->src && p->src_off && p->src_hw && p->dst && p->dst_hw, DV_ERR_ARG, "dv_rescale: null pointer");
  DV_REQUIRE(p->B > 0 && p->B <= 65535 && p->Hs > 0 && p->Ws > 0 && (int64_t)p->Hs * p->Ws < (1LL << 31), DV_ERR_ARG,
             "dv_rescale: bad sizes B=%d Hs=%d Ws=%d", p->B, p->Hs, p->Ws);
  RescaleArgs a;
  a.src = p->src; a.src_off = p->src_off; a.src_hw = p->src_hw; a.dst_hw = p->dst_hw; a.dst = p->dst;
  a.B = p->B; a.Hs = p->Hs; a.Ws = p->Ws;
  dim3 grid((unsigned)ceil_div(p->Hs * p->Ws, 256), (unsigned)p->B);
  rescale_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  DV_LAUNCH_OK("rescale_kernel");
  return DV_OK;
}

extern "C" int dv_augment(const dv_augment_params* p, void* stream) {
  DV_REQUIRE(p && p->src && p->dst && p->geom && p->factors && p->gray_sums, DV_ERR_ARG, "dv_augment: null pointer");
  DV_REQUIRE(p->B > 0 && p->Hs > 0 && p->Ws > 0 && p->crop > 0 && p->crop <= p->Hs && p->crop <= p->Ws, DV_ERR_ARG,
             "dv_augment: bad sizes B=%d Hs=%d Ws=%d crop=%d", p->B, p->Hs, p->Ws, p->crop);
  DV_REQUIRE(p->dtype == DV_F32 || p->dtype == DV_BF16, DV_ERR_ARG, "dv_augment: bad dtype %d", p->dtype);
  DV_REQUIRE(p->B <= 65535, DV_ERR_UNSUPPORTED, "dv_augment: batch %d exceeds grid.y", p->B);
  DV_REQUIRE((int64_t)p->crop * p->crop <= 0xFFFFFFFFLL / 255, DV_ERR_UNSUPPORTED, "dv_augment: crop too large for the 32-bit gray sum");
  for (int k = 0; k < 3; ++k) DV_REQUIRE(p->std[k] != 0.f, DV_ERR_ARG, "dv_augment: std[%d] is zero", k);
  cudaStream_t st = (cudaStream_t)stream;
  AugArgs a;
  a.src = p->src; a.geom = p->geom; a.factors = p->factors; a.widths = p->widths; a.gray_sums = p->gray_sums;
  a.B = p->B; a.Hs = p->Hs; a.Ws = p->Ws; a.crop = p->crop;
  for (int k = 0; k < 3; ++k) { a.mean[k] = p->mean[k]; a.stdv[k] = p->std[k]; }
  DV_CUDA_OK(cudaMemsetAsync(p->gray_sums, 0, (size_t)p->B * sizeof(uint32_t), st));
  dim3 grid((unsigned)ceil_div(p->crop * p->crop, 256), (unsigned)p->B);
  augment_gray_sum_kernel<<<grid, 256, 0, st>>>(a);
  DV_LAUNCH_OK("augment_gray_sum_kernel");
  if (p->dtype == DV_BF16) {
    if (p->nchw) augment_apply_kernel<bf16, true><<<grid, 256, 0, st>>>(a, (bf16*)p->dst);
    else augment_apply_kernel<bf16, false><<<grid, 256, 0, st>>>(a, (bf16*)p->dst);
  } else {
    if (p->nchw) augment_apply_kernel<float, true><<<grid, 256, 0, st>>>(a, (float*)p->dst);
    else augment_apply_kernel<float, false><<<grid, 256, 0, st>>>(a, (float*)p->dst);
  }
  DV_LAUNCH_OK("augment_apply_kernel");
  return DV_OK;
}
