// common.cuh — shared helpers for libdv_b200 (error reporting, dtype load/store, launch accounting).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/dv.h"

namespace dv {

// ---- error reporting (thread-local message buffer, never throws) -------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define DV_REQUIRE(cond, code, ...)            \
  do {                                         \
    if (!(cond)) {                             \
      ::dv::set_error(__VA_ARGS__);            \
      return (code);                           \
    }                                          \
  } while (0)

#define DV_CUDA_OK(expr)                                                                    \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      ::dv::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return DV_ERR_CUDA;                                                                   \
    }                                                                                       \
  } while (0)

// checks the launch that was just enqueued
#define DV_LAUNCH_OK(name)                                                                  \
  do {                                                                                      \
    cudaError_t _e = cudaGetLastError();                                                    \
    if (_e != cudaSuccess) {                                                                \
      ::dv::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));             \
      return DV_ERR_CUDA;                                                                   \
    }                                                                                       \
    ::dv::count_launch();                                                                   \
  } while (0)

int num_sms();  // cached cudaDevAttrMultiProcessorCount of the current device (148 on B200)

// ---- dtype helpers ------------------------------------------------------------------------------
typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

// A vector of VEC elements of T that is 16 bytes wide: 4 floats or 8 bf16.
template <typename T> struct Vec16;
template <> struct Vec16<float> {
  static constexpr int N = 4;
  float4 raw;
  __device__ __forceinline__ void load(const float* p) { raw = *reinterpret_cast<const float4*>(p); }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = raw; }
  __device__ __forceinline__ void unpack(float* f) const { f[0] = raw.x; f[1] = raw.y; f[2] = raw.z; f[3] = raw.w; }
  __device__ __forceinline__ void pack(const float* f) { raw = make_float4(f[0], f[1], f[2], f[3]); }
};
template <> struct Vec16<bf16> {
  static constexpr int N = 8;
  uint4 raw;
  __device__ __forceinline__ void load(const bf16* p) { raw = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void store(bf16* p) const { *reinterpret_cast<uint4*>(p) = raw; }
  __device__ __forceinline__ void unpack(float* f) const {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      // bf16 -> fp32 is a 16-bit shift
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ __forceinline__ void pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    raw = make_uint4(w[0], w[1], w[2], w[3]);
  }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// ---- per-module launchers (defined in the .cu files, called from api.cu) --------------------------
int conv_simt(const dv_conv_params* p, int pass, cudaStream_t st);
int conv_tc(const dv_conv_params* p, int pass, cudaStream_t st);
bool conv_tc_supported(const dv_conv_params* p, int pass, const char** why);
size_t conv_tc_workspace(const dv_conv_params* p, int pass);
// bf16 tiled tensor map (encoded through the driver entry point conv_tc.cu resolves): `tm` points at a 64-byte aligned CUtensorMap,
// dims in elements (innermost first), strides in bytes for dims 1..rank-1, swizzle 0 = none, 1 = 32 B, 2 = 64 B, 3 = 128 B.
// Out-of-range box elements read as zero.
int make_tiled_map_bf16(void* tm, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                        int swizzle);

}  // namespace dv
