// Probe: does cuTensorMapEncodeTiled accept OVERLAPPING strides (stride of dim 1 smaller than the extent of dim 0), and does a
// SWIZZLE_64B box load through such a map deliver the sliding windows?  Image row of 230 pixels x 4 channels (bf16, 8 B per pixel);
// dim0 = 32 elements (8 pixels), dim1 = 112 window origins 16 B apart, dim2 = rows.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tma_overlap tma_overlap.cu && ./tma_overlap
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__global__ void probe(const __grid_constant__ CUtensorMap tm, uint16_t* out, int row) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  const uint32_t sb = (uint32_t)__cvta_generic_to_shared(&bar), ss = (uint32_t)__cvta_generic_to_shared(smem);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sb));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sb), "r"(112 * 64));
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(ss),
                 "l"(reinterpret_cast<uint64_t>(&tm)), "r"(sb), "r"(0), "r"(0), "r"(row)
                 : "memory");
  }
  uint32_t done = 0;
  while (!done)
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(sb));
  for (int i = threadIdx.x; i < 112 * 32; i += blockDim.x) out[i] = reinterpret_cast<uint16_t*>(smem)[i];
}

int main() {
  const int Wp = 230, Hp = 9;
  std::vector<uint16_t> img((size_t)Hp * Wp * 4);
  for (size_t i = 0; i < img.size(); ++i) img[i] = (uint16_t)(i & 0xffff);
  uint16_t *d_img, *d_out;
  cudaMalloc(&d_img, img.size() * 2);
  cudaMalloc(&d_out, 112 * 32 * 2);
  cudaMemcpy(d_img, img.data(), img.size() * 2, cudaMemcpyHostToDevice);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  alignas(64) CUtensorMap tm;
  cuuint64_t dims[3] = {32, 112, (cuuint64_t)Hp};
  cuuint64_t strides[2] = {16, (cuuint64_t)Wp * 8};
  cuuint32_t box[3] = {32, 112, 1}, estr[3] = {1, 1, 1};
  CUresult r = ((EncodeTiledFn)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, d_img, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode overlapping strides: CUresult %d\n", (int)r);
  if (r != CUDA_SUCCESS) return 1;
  const int row = 5;
  probe<<<1, 128, 112 * 64 + 1024>>>(tm, d_out, row);
  cudaError_t e = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(e));
  if (e != cudaSuccess) return 2;
  std::vector<uint16_t> out(112 * 32);
  cudaMemcpy(out.data(), d_out, out.size() * 2, cudaMemcpyDeviceToHost);
  // SWIZZLE_64B: 16-byte chunk index (0..3) XOR ((row >> 1) & 3) within each 64-byte row
  int bad = 0;
  for (int q2 = 0; q2 < 112; ++q2)
    for (int e2 = 0; e2 < 32; ++e2) {
      const int chunk = e2 / 8, within = e2 % 8;
      const int pc = chunk ^ ((q2 >> 1) & 3);
      const uint16_t got = out[q2 * 32 + pc * 8 + within];
      const uint16_t want = img[((size_t)row * Wp + 2 * q2) * 4 + e2];
      if (got != want && bad++ < 5) printf("mismatch q=%d e=%d got %u want %u\n", q2, e2, got, want);
    }
  printf("windows %s (%d mismatches)\n", bad ? "WRONG" : "OK", bad);
  return bad ? 3 : 0;
}
