/*
 * dv.h — C ABI of libdv_b200.so: the B200 (sm_100a) kernels behind the conv-BN-ReLU hot path
 * of zackdilan/deep-vision (PyTorch ResNet-34/50/152 and MobileNetV1 blocks, forward+backward).
 *
 * The reference has no FFI of its own: its hot path is a sequence of torch.nn primitives called
 * from four model files.  Each entry point below replaces ONE such primitive (the reference call
 * site is cited beside it) and is what a ctypes / cffi / pybind stub on the reference side binds.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller; the library never allocates device
 *     memory, never frees, never synchronises the device;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream); calls return after
 *     enqueue;
 *   - activations are NHWC ([N][H][W][C], C contiguous); dense filters are KRSC ([K][R][S][C]);
 *     depth-wise filters are [C][3][3] (== torch [C,1,3,3]);
 *   - dtype: DV_F32 or DV_BF16 is the storage type of activations / filters; accumulation and
 *     all per-channel statistics are fp32;
 *   - return value: 0 on success, a DV_ERR_* code otherwise; dv_last_error() gives the message of
 *     the last failure on the calling thread.  Nothing throws, nothing calls exit().
 *   - there is no CPU path: on a machine without an sm_100 GPU every compute entry point fails
 *     with DV_ERR_CUDA / DV_ERR_ARCH.
 */
#ifndef DV_B200_H
#define DV_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DV_VERSION 100 /* 0.1.0 */

enum { DV_F32 = 0, DV_BF16 = 1 };

enum {
  DV_OK = 0,
  DV_ERR_ARG = 1,       /* bad dims / null pointer / misalignment */
  DV_ERR_UNSUPPORTED = 2,
  DV_ERR_CUDA = 3,      /* a CUDA runtime/driver call failed */
  DV_ERR_ARCH = 4,      /* device is not compute capability 10.x */
  DV_ERR_WORKSPACE = 5  /* workspace too small */
};

/* conv algorithm selector */
enum {
  DV_ALGO_AUTO = 0, /* tensor-core path when the shape qualifies, else SIMT */
  DV_ALGO_SIMT = 1, /* CUDA-core implicit GEMM (any shape, fp32 or bf16 storage) */
  DV_ALGO_TC = 2    /* tcgen05/TMEM/TMA implicit GEMM (bf16 only); error if shape unsupported */
};

/* ------------------------------------------------------------------------------------------
 * Dense convolution, no bias, zero padding, cross-correlation (torch semantics).
 * Replaces nn.Conv2d(bias=False) at ResNet/pytorch/models/resnet50.py:20-28,101-107,111-118,
 * 122-128,137-143; resnet34.py:20-28,96-112,120-126; MobileNet/pytorch/models/mobilenet_v1.py:30,
 * 139-145 and (as a 1x1 conv on [B,1,1,F]) nn.Linear at resnet50.py:44, mobilenet_v1.py:87.
 *   fprop:  y[n,p,q,k]  = sum_{r,s,c} x[n, p*stride+r-pad, q*stride+s-pad, c] * w[k,r,s,c]
 *   dgrad:  dx[n,h,w,c] = sum_{r,s,k} dy[n,p,q,k] * w[k,r,s,c]   over (p,q) with p*stride+r-pad==h ...
 *   wgrad:  dw[k,r,s,c] = sum_{n,p,q} dy[n,p,q,k] * x[n, p*stride+r-pad, q*stride+s-pad, c]   (fp32 out)
 * P = (H + 2*pad - R)/stride + 1, Q likewise.
 * ------------------------------------------------------------------------------------------ */
typedef struct dv_conv_params {
  /* fprop: x in, w in, y out.  dgrad: y = dy in, w in, x = dx out.  wgrad: x in, y = dy in, w = dw out (fp32 KRSC). */
  void* x;            /* [N,H,W,C] dtype */
  void* w;            /* [K,R,S,C] dtype (fprop,dgrad: for DV_ALGO_TC dgrad see w_t) ; wgrad: fp32 out */
  void* y;            /* [N,P,Q,K] dtype */
  const void* w_t;    /* dgrad on the tensor-core path only, OPTIONAL: filters as [C][R][S][K] with taps flipped
                         (w_t[c,r,s,k] = w[k,R-1-r,S-1-s,c]), made by dv_filter_prep.  NULL (the normal case): the kernel reads
                         the KRSC filters `w` directly as an MN-major operand and flips the taps by index */
  void* workspace;    /* scratch, see dv_conv_workspace_bytes; may be NULL when that returns 0 */
  size_t workspace_bytes;
  float* stat_sums;   /* fprop on the tensor-core path only, optional: [2][K] fp32 that the epilogue ADDS the per-channel sum
                         and sum of squares of the (bf16-rounded) output to — the BatchNorm statistics pass that follows the
                         convolution (resnet50.py:152-153) without re-reading y; see dv_bn_stats_from_sums.  Ignored by the
                         CUDA-core path (ask dv_conv_algo) */
  int32_t N, H, W, C, K, R, S, stride, pad;
  int32_t dtype;      /* DV_F32 | DV_BF16 */
  int32_t algo;       /* DV_ALGO_* */
  int32_t y_f32;      /* fprop only: 1 = write y as fp32 even when dtype is bf16 (the logits of nn.Linear) */
  int32_t accumulate; /* dgrad: 1 = dx += result instead of dx = result (the second gradient of a tensor that feeds two
                         consumers, e.g. a residual block's input: resnet50.py:147-163).
                         wgrad: 1 = dw += result, added straight into dw (fp32 tiles through TMA reduce-add) — no workspace, no
                         second pass; the caller provides a zeroed (or to-be-accumulated) dw and accepts a summation order
                         that varies run to run.  0 = split partials in the workspace + fixed-order reduce (bit-reproducible) */
  /* fprop on the tensor-core path only, optional — inference with the BatchNorm that follows the convolution folded into
   * its epilogue (eval-mode nn.BatchNorm2d + nn.ReLU + `x + identity`, resnet50.py:152-164 under net.eval(), train.py:489):
   *   y = relu( acc * ep_scale[k] + ep_shift[k] + ep_residual[n,p,q,k] )
   * with ep_scale = gamma / sqrt(running_var + eps), ep_shift = beta - running_mean * ep_scale (fp32 [K], made by the caller).
   * ep_residual (dtype [N,P,Q,K]) may be NULL; ep_relu 0/1.  All NULL/0 = plain convolution.  Any other path rejects them
   * with DV_ERR_UNSUPPORTED rather than ignoring them (dgrad gives ep_residual its own meaning: see ep_residual_mask below). */
  const float* ep_scale;
  const float* ep_shift;
  const void* ep_residual;
  int32_t ep_relu;
  /* dgrad on the tensor-core path only, optional — the backward sums of the BatchNorm whose OUTPUT is this convolution's input.
   * In conv1 -> bn1 -> relu -> conv2 (resnet50.py:152-158) the dgrad of conv2 produces dy of bn1; nn.BatchNorm2d's backward then
   * needs  dbeta = sum_m dz,  dgamma = sum_m dz * xhat,  dz = dy * (bn1(x0) > 0),  over the M = N*H*W rows of x0 (bn1's input, the
   * output of conv1).  With bnb_x set, the dgrad epilogue fetches the x0 tile that matches each dx tile while it is still on chip
   * and ADDS those sums into bnb_dgamma / bnb_dbeta (fp32 [C], zeroed or to-be-accumulated by the caller; run-to-run summation
   * order), so dv_bn_bwd_reduce need not run for that layer (it would re-read dy and x0 from HBM: 4 bytes per element).
   *   bnb_x [N,H,W,C] dtype (same geometry as dx); bnb_gamma / bnb_beta / bnb_mean / bnb_invstd: fp32 [C] of that BatchNorm
   *   (save_mean / save_invstd of its forward).  The ReLU mask is recomputed from x0 with bn_apply's own expressions.
   * Requires stride == 1, accumulate == 0, w_t == NULL and the tensor-core path (ask dv_conv_algo); anything else is rejected
   * with DV_ERR_UNSUPPORTED rather than ignored.  All NULL = plain dgrad. */
  const void* bnb_x;
  const float* bnb_gamma;
  const float* bnb_beta;
  const float* bnb_mean;
  const float* bnb_invstd;
  float* bnb_dgamma;
  float* bnb_dbeta;
  /* dgrad on the tensor-core path only, optional — the shortcut's share of a residual block's input gradient added in the
   * epilogue.  In  out = relu(bn3(conv3(...conv1(x)...)) + x)  (resnet50.py:147-164) the gradient of x is
   *   dgrad(conv1) + dout * (out > 0);
   * with ep_residual = dout [N,H,W,C] (dtype) and ep_residual_mask = the one-bit-per-element ReLU mask dv_bn_apply wrote for that
   * block ([N*H*W][C/8] bytes, bit c%8 of byte c/8; NULL = add ep_residual unmasked), the dgrad of conv1 writes
   *   dx = acc + (mask ? ep_residual : 0)       (one rounding to bf16)
   * so the masked gradient is never stored by dv_bn_bwd_apply (dres == NULL there) nor read-modified-written here.
   * Requires stride == 1, accumulate == 0, w_t == NULL, no bnb_*, C % 32 == 0; rejected with DV_ERR_UNSUPPORTED otherwise. */
  const uint8_t* ep_residual_mask;
} dv_conv_params;

enum { DV_CONV_FPROP = 0, DV_CONV_DGRAD = 1, DV_CONV_WGRAD = 2 };

int dv_conv_fprop(const dv_conv_params* p, void* stream);
int dv_conv_dgrad(const dv_conv_params* p, void* stream);
int dv_conv_wgrad(const dv_conv_params* p, void* stream);
/* bytes of scratch the given pass (DV_CONV_*) needs for these params (0 = none) */
size_t dv_conv_workspace_bytes(const dv_conv_params* p, int pass);
/* which algorithm DV_ALGO_AUTO resolves to for this pass: DV_ALGO_SIMT or DV_ALGO_TC */
int dv_conv_algo(const dv_conv_params* p, int pass);

/* fp32 KRSC master filters -> compute-dtype copies: w_out[K,R,S,C] (cast) and, if w_t_out != NULL,
 * w_t_out[C,R,S,K] with taps flipped (the dgrad operand).  No counterpart call site in the reference
 * (torch keeps one fp32 OIHW tensor); this is the "NHWC plumbing" of the weight side. */
int dv_filter_prep(const float* w_f32, void* w_out, void* w_t_out, int K, int R, int S, int C, int dtype, void* stream);

/* Small-C convolutions (the C=3 stems: resnet50.py:20-28 7x7/2, mobilenet_v1.py:30 3x3/2) reach the tensor cores
 * through an explicit patch matrix: patches[m, (r*S+s)*C + c] = x[n, p*stride+r-pad, q*stride+s-pad, c] (zero outside
 * the image and for columns >= R*S*C), m = (n*P+p)*Q+q, row pitch Kp (a multiple of 64).  The conv is then the 1x1
 * convolution of `patches` ([M,1,1,Kp]) with the filters padded to [K][Kp] (dv_pad_rows), and wgrad is the 1x1 wgrad
 * followed by dv_unpad_rows.  No counterpart call in the reference: torch's conv does this internally. */
int dv_im2col(const void* x, void* patches, int N, int H, int W, int C, int R, int S, int stride, int pad, int Kp, int dtype,
              void* stream);
/* dst[r, 0:cols] = src[r, 0:cols] (cast fp32 -> dtype), dst[r, cols:dst_pitch] = 0 */
int dv_pad_rows(const float* src, void* dst, int rows, int cols, int dst_pitch, int dtype, void* stream);
/* dst[r, 0:cols] = src[r, 0:cols] with src row pitch src_pitch (fp32 -> fp32) */
int dv_unpad_rows(const float* src, float* dst, int rows, int cols, int src_pitch, void* stream);

/* The same stems WITHOUT a patch matrix (tensor-core path, bf16): the image is kept as NHWC4 with its zero border written out
 * (dv_stem_pad_input: [N][Hp][Wp][4] bf16, 8 bytes per pixel) and every patch row is one TMA box of a tensor map with
 * overlapping strides (window origins 16 bytes = one stride-2 step apart), so fprop and wgrad read the 100 MB image where the
 * patch-matrix path writes and re-reads 1.2 GB.  Requirements: C <= 4, stride 2, R,S <= 7, K a multiple of 8 and <= 64, output
 * width Q a multiple of 16 and <= 128 (dv_stem_supported answers 1 and the padded extents).  wgrad writes dw = the gradient
 * (never accumulates into dw itself); `accumulate` only picks how CTA tiles are combined: 1 = fp32 atomics (run-to-run order),
 * 0 = per-CTA partials in the workspace + fixed-order sum.  Replaces nn.Conv2d at resnet50.py:20-28, mobilenet_v1.py:30. */
typedef struct dv_stem_params {
  void* x;            /* padded image [N][Hp][Wp][4] bf16: out of dv_stem_pad_input, in of fprop / wgrad */
  void* w;            /* packed filters [64][R][32] bf16: out of dv_stem_pack_filters, in of fprop */
  void* y;            /* fprop: out [N,P,Q,K] bf16; wgrad: dy in */
  float* dw;          /* wgrad: out, fp32 KRSC [K][R][S][C] */
  float* stat_sums;   /* fprop, optional: as dv_conv_params.stat_sums */
  void* workspace;    /* wgrad: dv_stem_workspace_bytes() */
  size_t workspace_bytes;
  int32_t N, H, W, C, K, R, S, stride, pad;
  int32_t accumulate;
} dv_stem_params;
int dv_stem_supported(const dv_stem_params* p, int* Hp, int* Wp);
int dv_stem_pad_input(const dv_stem_params* p, const void* x_nhwc, void* stream);   /* x_nhwc: [N,H,W,C] bf16 */
/* the same straight from the fp32 [N,C,H,W] batch the reference's loader yields (train.py:444-449): dv_nchw_to_nhwc's rounding and
 * dv_stem_pad_input's border in one pass over the image, bit-identical to the two calls */
int dv_stem_pad_input_nchw(const dv_stem_params* p, const float* x_nchw, void* stream);
int dv_stem_pack_filters(const dv_stem_params* p, const float* w_krsc, void* stream);
int dv_stem_fprop(const dv_stem_params* p, void* stream);
int dv_stem_wgrad(const dv_stem_params* p, void* stream);
size_t dv_stem_workspace_bytes(const dv_stem_params* p);

/* ------------------------------------------------------------------------------------------
 * Depth-wise 3x3 convolution, pad 1, stride 1 or 2, groups == C.
 * Replaces nn.Conv2d(C, C, 3, padding=1, stride, groups=C, bias=False) at mobilenet_v1.py:112-122.
 * w is fp32 [C][3][3] for all three passes (tiny); dw out is fp32 [C][3][3].
 * wgrad scratch: dv_dwconv_workspace_bytes().
 * ------------------------------------------------------------------------------------------ */
typedef struct dv_dwconv_params {
  void* x;   /* [N,H,W,C] dtype ; dgrad: dx out */
  void* w;   /* fp32 [C,3,3]    ; wgrad: dw out */
  void* y;   /* [N,P,Q,C] dtype ; dgrad/wgrad: dy in */
  void* workspace;
  size_t workspace_bytes;
  int32_t N, H, W, C, stride;
  int32_t dtype;
  /* fprop only, optional — inference with the following eval-mode BatchNorm + ReLU folded in (mobilenet_v1.py:129-133
   * under net.eval()):  y = relu(conv(x, w) + ep_shift[c]).  The BN scale needs no field: the taps are fp32 and per
   * channel, so the caller passes w[c,:,:] * scale[c].  dgrad / wgrad reject a non-NULL ep_shift. */
  const float* ep_shift;
  int32_t ep_relu;
  float* stat_sums;   /* fprop only, optional: [2][C] fp32 that the kernel ADDS the per-channel sum and sum of squares of the (bf16-rounded)
                         output to — the statistics pass of the BatchNorm that follows (mobilenet_v1.py:126), see dv_bn_stats_from_sums.
                         Honoured by the TMA-staged stride-1 kernel only: dv_dwconv_fuses_stats() answers 1 for such parameters, any
                         other path rejects a non-NULL stat_sums */
  /* dgrad only, optional — the backward sums of the BatchNorm whose OUTPUT is this convolution's input (pw.bn -> relu -> dw of the
   * next block, mobilenet_v1.py:129-133,152-156), exactly as dv_conv_params.bnb_*: with bnb_x = that BatchNorm's input x0 [N,H,W,C]
   * the kernel ADDS dbeta = sum dz and dgamma = sum dz * xhat (dz = dx masked by bn(x0) > 0, from the bf16-rounded dx it stores)
   * into bnb_dbeta / bnb_dgamma, so dv_bn_bwd_reduce is skipped for that layer.  Honoured by the TMA-staged stride-1 kernel only:
   * dv_dwconv_fuses_bn_bwd() answers 1 for such parameters, any other path rejects a non-NULL bnb_x. */
  const void* bnb_x;
  const float* bnb_gamma;
  const float* bnb_beta;
  const float* bnb_mean;
  const float* bnb_invstd;
  float* bnb_dgamma;
  float* bnb_dbeta;
} dv_dwconv_params;
int dv_dwconv_fuses_stats(const dv_dwconv_params* p);
int dv_dwconv_fuses_bn_bwd(const dv_dwconv_params* p);

int dv_dwconv_fprop(const dv_dwconv_params* p, void* stream);
int dv_dwconv_dgrad(const dv_dwconv_params* p, void* stream);
int dv_dwconv_wgrad(const dv_dwconv_params* p, void* stream);
size_t dv_dwconv_workspace_bytes(const dv_dwconv_params* p);

/* ------------------------------------------------------------------------------------------
 * Batch normalisation over (N,H,W) per channel, torch defaults (eps 1e-5, momentum 0.1).
 * Replaces nn.BatchNorm2d at resnet50.py:30,108,119,129,144; resnet34.py:30,104,113,127;
 * mobilenet_v1.py:31,126,149 — together with the nn.ReLU(inplace=True) that follows it
 * (resnet50.py:31,131) and the residual add (resnet50.py:163, resnet34.py:141), which are fused
 * into the apply pass:   y = relu( (x-mean)*invstd*gamma + beta  [+ residual] ).
 * M = N*H*W rows of C channels.
 * ------------------------------------------------------------------------------------------ */
typedef struct dv_bn_params {
  const void* x;        /* [M,C] dtype: conv output */
  void* y;              /* [M,C] dtype: block output (apply) */
  const void* residual; /* [M,C] dtype or NULL */
  const float* gamma;   /* [C] */
  const float* beta;    /* [C] */
  float* running_mean;  /* [C] updated by dv_bn_stats when not NULL */
  float* running_var;   /* [C] (unbiased variance, as torch) */
  float* save_mean;     /* [C] batch mean     (stats: out; others: in). eval mode: unused */
  float* save_invstd;   /* [C] 1/sqrt(var+eps) (stats: out; others: in) */
  void* workspace;      /* dv_bn_workspace_bytes() */
  size_t workspace_bytes;
  int64_t M;
  int32_t C;
  int32_t dtype;
  int32_t relu;         /* 1 = fuse ReLU */
  int32_t eval_mode;    /* apply only: 1 = normalise with running_mean/var instead of save_* */
  float eps, momentum;
  void* relu_mask;      /* optional (bf16 only): uint8 [M][C/8] written by dv_bn_apply when relu == 1; bit j of byte
                           (m, c/8) = (y[m][c+j] > 0).  Lets the backward of a RESIDUAL block read 1 bit instead of 16 per
                           element for the ReLU mask (see dv_bn_bwd_params.relu_mask) */
  int32_t reverse;      /* dv_bn_apply: 1 = walk the rows from the last to the first (same result).  A pass that starts where
                           its producer stopped finds those bytes still in the 126 MB L2 */
} dv_bn_params;

/* batch statistics (mean, biased var -> invstd) + running-stat update */
int dv_bn_stats(const dv_bn_params* p, void* stream);
/* the same results from sums a convolution's epilogue accumulated (dv_conv_params.stat_sums): sums[0:C] = sum x,
 * sums[C:2C] = sum x^2 over the M rows.  Writes save_mean / save_invstd, updates the running statistics and CLEARS sums
 * to zero for the next step.  p->x and the workspace are not used. */
int dv_bn_stats_from_sums(const dv_bn_params* p, float* sums, void* stream);
/* normalise + affine (+residual) (+ReLU) */
int dv_bn_apply(const dv_bn_params* p, void* stream);
/* dv_bn_stats for callers that cannot allocate device memory.  (This entry point has no counterpart in the
 * reference, which leaves every allocation to torch.)
 *
 * Who needs it: a host that hands this library tensors but keeps the allocator to itself — an inference or
 * training runtime that reserves all of its device memory at start-up and forbids cudaMalloc afterwards (an allocation
 * is a device-wide synchronisation point and can fail late, deep inside a step), or a plug-in that is handed
 * input/output buffers and has no sanctioned way to ask for a workspace.  Such a caller cannot size and provide the
 * `workspace` that dv_bn_stats wants.
 *
 * Same arguments, same results as dv_bn_stats (the block partials are summed over a different, coarser partition of
 * the rows, so the statistics agree to rounding, not bit for bit).  `workspace` / `workspace_bytes` are ignored: the
 * block partials go to a fixed array compiled into the library, 148 x 2 x 2048 floats = 2.3 MiB, which every GPU that
 * loads the library holds until the process exits.  Neither this call nor anything it calls allocates or frees
 * (no cudaMalloc / cudaFree, no torch).
 *   - Limit: C <= 2048 channels (any M).  Larger C is rejected with DV_ERR_UNSUPPORTED and a message; nothing is
 *     launched in that case.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress, result checked and
 *     cached per device).  Make that first call before any stream capture begins.
 *   - All calls on a device share the one array.  It keeps whatever the previous call left there; that is harmless
 *     because each call overwrites every partial it later reads, so nothing needs clearing.  But two calls that can
 *     overlap on the GPU (different streams, or two host threads) race on it: serialise them on one stream. */
int dv_bn_stats_static(const dv_bn_params* p, void* stream);

typedef struct dv_bn_bwd_params {
  const void* x;        /* [M,C] conv output saved from forward */
  const void* y;        /* [M,C] block output saved from forward (ReLU mask).  May be NULL when relu == 0, and also when
                           relu == 1 and the forward had NO residual: the mask is then recomputed bit-exactly from x
                           (needs beta), which saves reading y in both backward passes */
  const void* dy;       /* [M,C] gradient w.r.t. block output */
  void* dx;             /* [M,C] gradient w.r.t. conv output */
  void* dres;           /* [M,C] gradient w.r.t. residual (== masked dy) or NULL */
  const float* gamma;
  const float* beta;    /* [C]; only read when the mask is recomputed (relu == 1, y == NULL) */
  const float* save_mean;
  const float* save_invstd;
  float* dgamma;        /* [C] out */
  float* dbeta;         /* [C] out */
  void* workspace;
  size_t workspace_bytes;
  int64_t M;
  int32_t C;
  int32_t dtype;
  int32_t relu;
  int32_t atomic_sums;  /* bwd_reduce: 1 = block sums are ADDED into dgamma/dbeta with fp32 atomics (the caller zeroed them):
                           no workspace, no finalize launch, run-to-run summation order.  0 = partials + fixed-order
                           finalize (bit-reproducible) */
  const void* relu_mask; /* optional (bf16 only): the bit mask dv_bn_apply wrote; when set (and relu == 1) it is used
                           instead of y, which is then not read by either backward pass */
  int32_t reverse;      /* dv_bn_bwd_apply: 1 = walk the rows from the last to the first (same result); see dv_bn_params.reverse */
} dv_bn_bwd_params;

/* dgamma, dbeta (with the ReLU mask recomputed from y) */
int dv_bn_bwd_reduce(const dv_bn_bwd_params* p, void* stream);
/* dx = gamma*invstd*(dz - dbeta/M - xhat*dgamma/M), dres = dz, with dz = dy*(y>0) */
int dv_bn_bwd_apply(const dv_bn_bwd_params* p, void* stream);
size_t dv_bn_workspace_bytes(int64_t M, int C);
/* Inference: eval-mode BatchNorm (train.py:489 `net.eval()`, notebooks cell 3) as a per-channel affine map,
 *   scale[c] = gamma[c] / sqrt(running_var[c] + eps),  shift[c] = beta[c] - running_mean[c] * scale[c]   (fp32 [C]),
 * for the folded epilogues dv_conv_params.ep_* / dv_dwconv_params.ep_shift.  With dw_w (fp32 [C][3][3], the depth-wise
 * taps that precede this BN) it also writes dw_w_scaled[c][:] = dw_w[c][:] * scale[c].  No counterpart call in the
 * reference: torch evaluates eval-mode BN as its own kernel. */
int dv_bn_fold(const float* gamma, const float* beta, const float* running_mean, const float* running_var, float eps,
               float* scale, float* shift, int C, const float* dw_w, float* dw_w_scaled, void* stream);

/* nn.BatchNorm2d + nn.ReLU + nn.MaxPool2d(3, 2, 1) of the ResNet stem (resnet50.py:30-34, 49-52; resnet34.py likewise) as ONE pass
 * forward and two passes backward over conv1's output, the largest activation of the network: the normalised tensor and the
 * pool's input gradient (411 MB each at batch 256) never exist.  Bit for bit what dv_bn_apply(relu) -> dv_maxpool_fwd and
 * dv_maxpool_bwd -> dv_bn_bwd_reduce -> dv_bn_bwd_apply give.  The statistics come from dv_bn_stats / dv_bn_stats_from_sums as
 * usual (save_mean / save_invstd in).  C must be V times a power of two (V = 8 for bf16, 4 for fp32), C/V <= 256. */
typedef struct dv_bn_pool_params {
  const void* x;        /* [N,H,W,C] conv output (BatchNorm input) */
  void* y;              /* fwd: out [N,P,Q,C], P = (H-1)/2+1, Q = (W-1)/2+1 */
  uint8_t* idx;         /* [N,P,Q,C] window position (0..8) of the maximum: fwd out, bwd in */
  const void* dy;       /* bwd: [N,P,Q,C] gradient w.r.t. y */
  void* dx;             /* bwd: [N,H,W,C] gradient w.r.t. x */
  const float* gamma;
  const float* beta;
  const float* save_mean;
  const float* save_invstd;
  float* dgamma;        /* bwd: [C] out */
  float* dbeta;         /* bwd: [C] out */
  void* workspace;      /* bwd: dv_bn_pool_workspace_bytes() */
  size_t workspace_bytes;
  int32_t N, H, W, C;
  int32_t dtype;
} dv_bn_pool_params;
int dv_bn_relu_maxpool_fwd(const dv_bn_pool_params* p, void* stream);
int dv_bn_relu_maxpool_bwd(const dv_bn_pool_params* p, void* stream);
size_t dv_bn_pool_workspace_bytes(const dv_bn_pool_params* p);

/* stand-alone ReLU (nn.ReLU at a site that no BN precedes; not used by the four models) */
int dv_relu_fwd(const void* x, void* y, int64_t n, int dtype, void* stream);
int dv_relu_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * Input pipeline after decode + rescale: the reference's per-image transforms
 *   RandomHorizontalFlip -> RandomCrop | CenterCrop -> ColorJitter(brightness, contrast, saturation; hue 0)
 *   -> ToTensor (float, NOT divided by 255) -> Normalize(mean, std)
 * ResNet/pytorch/data_load.py:102-297 as composed at train.py:315-331, for a whole uint8 batch in two launches,
 * writing the model's input directly (NHWC compute dtype, or NCHW fp32 as the reference's loader yields).
 * The uint8 stages are Pillow's arithmetic bit for bit (oracle/ref_augment.py); the random draws stay with the caller.
 * ------------------------------------------------------------------------------------------ */
typedef struct dv_augment_params {
  const uint8_t* src;     /* [B][Hs][Ws][3] uint8 RGB: decoded and Rescale'd images (data_load.py:49-99), each in the
                             top-left corner of its Hs x Ws slot */
  void* dst;              /* [B][crop][crop][3] dtype (NHWC), or [B][3][crop][crop] when nchw != 0 */
  const int32_t* geom;    /* device [B][4]: top, left, flip (0/1), order = index into the permutations
                             (b,c,s) (b,s,c) (c,b,s) (c,s,b) (s,b,c) (s,c,b) of brightness/contrast/saturation
                             (data_load.py:270 shuffles them); order < 0 = no colour jitter (validation) */
  const float* factors;   /* device [B][3]: brightness, contrast, saturation factors (data_load.py:248-262) */
  const int32_t* widths;  /* device [B] true image widths (the flip mirrors the image, not the slot) or NULL = Ws */
  uint32_t* gray_sums;    /* device [B] scratch */
  int32_t B, Hs, Ws, crop;
  int32_t dtype;          /* of dst */
  int32_t nchw;
  float mean[3], std[3];  /* train.py:324: 0.485,0.456,0.406 / 0.229,0.224,0.225 applied to 0..255 values */
} dv_augment_params;
int dv_augment(const dv_augment_params* p, void* stream);

/* Rescale(256) (data_load.py:72-99): cv2.resize with INTER_LINEAR on uint8, bit for bit (OpenCV's 11-bit fixed-point
 * two-tap filter), for a ragged batch of decoded images in one launch, written into the slots dv_augment reads.
 * The sizes come from the caller (Rescale's own int() arithmetic); a slot's unused remainder is zeroed. */
typedef struct dv_rescale_params {
  const uint8_t* src;      /* device: decoded RGB images, HWC uint8, stored back to back */
  const int64_t* src_off;  /* device [B]: byte offset of image b inside src */
  const int32_t* src_hw;   /* device [B][2]: height, width of image b */
  uint8_t* dst;            /* device [B][Hs][Ws][3] */
  const int32_t* dst_hw;   /* device [B][2]: rescaled height, width (<= Hs, Ws) */
  int32_t B, Hs, Ws;
} dv_rescale_params;
int dv_rescale(const dv_rescale_params* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * Stem / tail ops.
 * ------------------------------------------------------------------------------------------ */
/* nn.MaxPool2d(3, 2, 1) — resnet50.py:34, resnet34.py:34.  idx: uint8 [N,P,Q,C] window position (0..8) of the max. */
int dv_maxpool_fwd(const void* x, void* y, uint8_t* idx, int N, int H, int W, int C, int dtype, void* stream);
int dv_maxpool_bwd(const void* dy, const uint8_t* idx, void* dx, int N, int H, int W, int C, int dtype, void* stream);
/* nn.AdaptiveAvgPool2d((1,1)) — resnet50.py:43, mobilenet_v1.py:84.  y: [N,C] */
int dv_avgpool_fwd(const void* x, void* y, int N, int HW, int C, int dtype, void* stream);
int dv_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, int dtype, void* stream);
/* NCHW fp32 image batch -> NHWC compute dtype (the layout change at the model entrance; train.py:444-449) */
int dv_nchw_to_nhwc(const float* x_nchw, void* y_nhwc, int N, int C, int H, int W, int dtype, void* stream);
/* plain cast between fp32 and bf16 (n elements) */
int dv_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n, void* stream);
/* y[m, :] += bias (fp32 y, fp32 bias) — the bias of nn.Linear, resnet50.py:44 */
int dv_bias_add(float* y, const float* bias, int64_t M, int C, void* stream);
/* dbias[c] = sum_m dy[m,c] (fp32) */
int dv_bias_grad(const float* dy, float* dbias, int64_t M, int C, void* stream);

/* nn.CrossEntropyLoss() (mean reduction) — ResNet/pytorch/train.py:358,452, forward and backward in one launch:
 *   loss[0]      = mean_b ( logsumexp(logits[b,:]) - logits[b, labels[b]] )
 *   dlogits[b,c] = (softmax(logits[b,:])[c] - (c == labels[b])) / B          (the gradient for d(loss) = 1)
 * logits, dlogits: fp32 [B,C]; labels: int64 [B]; loss: fp32 scalar (zeroed by the call). */
int dv_cross_entropy(const float* logits, const int64_t* labels, float* loss, float* dlogits, int B, int C, void* stream);
/* x[i] *= s[0] with s a DEVICE scalar (chain rule for a loss whose upstream gradient is not 1) */
int dv_scale_by(float* x, const float* s, int64_t n, void* stream);

/* ------------------------------------------------------------------------------------------
 * Optimisers over a flat fp32 parameter buffer (one launch for the whole model).
 * optim.SGD(lr, momentum, weight_decay) — ResNet/pytorch/train.py:166-184,361-365,467:
 *     g' = g*grad_scale + wd*p ; buf = first ? g' : mom*buf + g' ; p -= lr*buf
 * optim.RMSprop(lr, alpha, eps) — train.py:199-205:
 *     g' = g*grad_scale + wd*p ; sq = alpha*sq + (1-alpha)*g'^2 ; p -= lr*g'/(sqrt(sq)+eps)
 * ------------------------------------------------------------------------------------------ */
int dv_sgd_step(float* p, const float* g, float* buf, int64_t n, float lr, float momentum, float weight_decay,
                float grad_scale, int first_step, void* stream);
int dv_rmsprop_step(float* p, const float* g, float* sq, int64_t n, float lr, float alpha, float eps,
                    float weight_decay, float grad_scale, void* stream);
/* the same two updates that ALSO write the new parameters as bf16 into p_bf16[n] (same element order): the
 * compute-dtype filter copies of every convolution refreshed by the optimiser launch itself */
int dv_sgd_step_bf16(float* p, const float* g, float* buf, void* p_bf16, int64_t n, float lr, float momentum, float weight_decay,
                     float grad_scale, int first_step, void* stream);
int dv_rmsprop_step_bf16(float* p, const float* g, float* sq, void* p_bf16, int64_t n, float lr, float alpha, float eps,
                         float weight_decay, float grad_scale, void* stream);

/* ------------------------------------------------------------------------------------------
 * Library
 * ------------------------------------------------------------------------------------------ */
int dv_version(void);
const char* dv_last_error(void);
/* 0 if the current device is compute capability 10.x, else DV_ERR_ARCH / DV_ERR_CUDA */
int dv_device_check(void);
/* number of kernels this library has launched since load (all threads); used by bench.py */
uint64_t dv_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* DV_B200_H */
