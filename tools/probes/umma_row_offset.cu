// Probe: may the start address of a K-major SWIZZLE_128B tcgen05 shared-memory descriptor sit on ANY 128-byte row of a tile, not only
// on the 1024-byte boundary of the 8-row swizzle pattern?  A 3x3 convolution over a halo tile resident in shared memory would address
// its nine shifted windows that way.  A[row][k] lives in shared memory as rows of 64 bf16 (128 B), 16-byte chunk
// index XOR (row & 7) as TMA would have written it; B = identity, so D[m][n] must equal A[m + shift][n].  Each shift is tried with the
// descriptor's base-offset field (bits 49-51) left 0 and set to (start address >> 7) & 7.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o umma_row_offset umma_row_offset.cu && ./umma_row_offset
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int kRows = 272;   // rows of A kept in shared memory

// MN-major variant (the weight gradient's operands: rows of shared memory = pixels = the K dimension, 128 bytes = 64 channels).
// A = X with M = 128 taken as TWO 64-channel chunks `lbo_rows` pixel rows apart (the descriptor's leading-dimension byte offset):
// rows 0-63 of D then see X shifted by `shift` pixels, rows 64-127 X shifted by `shift + lbo_rows` — two filter taps in one MMA.
// B = DY, N = 64, unshifted.  D[m][n] = sum_i X[shift (+ lbo_rows) + i][m % 64] * DY[i][n], i < 64.
__global__ void __launch_bounds__(128) probe_mn(float* out, int shift, int lbo_rows) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* sX = smem;                       // kRows x 128 B
  uint8_t* sY = smem + kRows * 128;         // 64 x 128 B
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < kRows * 64; i += 128) {
    const int row = i >> 6, c = i & 63;
    reinterpret_cast<__nv_bfloat16*>(sX + row * 128 + (((c >> 3) ^ (row & 7)) * 16))[c & 7] = __float2bfloat16((float)((row * 5 + c * 3) % 7));
  }
  for (int i = tid; i < 64 * 64; i += 128) {
    const int row = i >> 6, c = i & 63;
    reinterpret_cast<__nv_bfloat16*>(sY + row * 128 + (((c >> 3) ^ (row & 7)) * 16))[c & 7] = __float2bfloat16((float)((row * 3 + c) % 5));
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(64) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  if (tid == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t a_addr = smem_u32(sX) + (uint32_t)shift * 128u, b_addr = smem_u32(sY);
    for (int k = 0; k < 4; ++k) {
      auto desc = [&](uint32_t addr, uint32_t lbo_bytes) {
        uint64_t d = (uint64_t)(((addr + k * 2048u) & 0x3FFFFu) >> 4);   // a K step = 16 pixel rows
        d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;               // distance between the 64-element chunks along M / N
        d |= (uint64_t)(1024u >> 4) << 32;                               // 8-row groups along K
        d |= (uint64_t)1 << 46;
        d |= (uint64_t)2 << 61;
        return d;
      };
      const uint64_t ad = desc(a_addr, (uint32_t)lbo_rows * 128u), bd = desc(b_addr, 8192u);
      const uint32_t acc = k ? 1u : 0u;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(acc)
          : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
  }
  uint32_t done = 0;
  while (!done)
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int half = 0; half < 2; ++half) {
    uint32_t v[32];
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(half * 32);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * 64 + half * 32 + j] = __uint_as_float(v[j]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64) : "memory");
}

__global__ void __launch_bounds__(128) probe(float* out, int shift, int use_base_offset) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* sA = smem;                       // kRows x 128 B
  uint8_t* sB = smem + kRows * 128;         // 64 x 128 B (kRows * 128 is a multiple of 1024)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < kRows * 64; i += 128) {
    const int row = i >> 6, k = i & 63;
    const int chunk = (k >> 3) ^ (row & 7);
    reinterpret_cast<__nv_bfloat16*>(sA + row * 128 + chunk * 16)[k & 7] = __float2bfloat16((float)((row * 3 + k) % 251));
  }
  for (int i = tid; i < 64 * 64; i += 128) {
    const int n = i >> 6, k = i & 63;
    const int chunk = (k >> 3) ^ (n & 7);
    reinterpret_cast<__nv_bfloat16*>(sB + n * 128 + chunk * 16)[k & 7] = __float2bfloat16(n == k ? 1.f : 0.f);
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(64) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  if (tid == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t a_addr = smem_u32(sA) + (uint32_t)shift * 128u, b_addr = smem_u32(sB);
    for (int k = 0; k < 4; ++k) {
      auto desc = [&](uint32_t addr, bool with_off) {
        uint64_t d = (uint64_t)(((addr + k * 32u) & 0x3FFFFu) >> 4);
        d |= (uint64_t)1 << 16;                    // LBO (unused for swizzled K-major)
        d |= (uint64_t)(1024u >> 4) << 32;         // SBO: 8-row groups 1024 B apart
        d |= (uint64_t)1 << 46;                    // version
        if (with_off) d |= (uint64_t)((addr >> 7) & 7u) << 49;
        d |= (uint64_t)2 << 61;                    // SWIZZLE_128B
        return d;
      };
      const uint64_t ad = desc(a_addr, use_base_offset != 0), bd = desc(b_addr, false);
      const uint32_t acc = k ? 1u : 0u;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(acc)
          : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
  }
  uint32_t done = 0;
  while (!done)
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int half = 0; half < 2; ++half) {
    uint32_t v[32];
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(half * 32);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * 64 + half * 32 + j] = __uint_as_float(v[j]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64) : "memory");
}

int main() {
  float* d_out;
  cudaMalloc(&d_out, 128 * 64 * sizeof(float));
  const int smem = kRows * 128 + 64 * 128;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<float> h(128 * 64);
  const int shifts[] = {0, 1, 2, 3, 5, 7, 8, 9, 64, 65, 66, 129, 130};
  int all_ok[2] = {1, 1};
  for (int mode = 0; mode < 2; ++mode) {
    for (int shift : shifts) {
      cudaMemset(d_out, 0xFF, 128 * 64 * sizeof(float));
      probe<<<1, 128, smem>>>(d_out, shift, mode);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("mode %d shift %d: %s\n", mode, shift, cudaGetErrorString(e)); return 1; }
      cudaMemcpy(h.data(), d_out, h.size() * sizeof(float), cudaMemcpyDeviceToHost);
      int bad = 0, first_m = -1, first_n = -1;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 64; ++n) {
          const float want = (float)(((m + shift) * 3 + n) % 251);
          if (h[m * 64 + n] != want) { if (!bad) { first_m = m; first_n = n; } ++bad; }
        }
      if (bad) all_ok[mode] = 0;
      printf("base_offset %s  shift %3d rows: %s (%d of 8192 wrong", mode ? "(addr>>7)&7" : "0          ", shift, bad ? "MISMATCH" : "ok", bad);
      if (bad) printf("; first at m=%d n=%d got %.0f want %.0f", first_m, first_n, h[first_m * 64 + first_n], (float)(((first_m + shift) * 3 + first_n) % 251));
      printf(")\n");
    }
  }
  printf("RESULT: base_offset=0 %s, base_offset=(addr>>7)&7 %s\n", all_ok[0] ? "works for every shift" : "fails", all_ok[1] ? "works for every shift" : "fails");
  // MN-major: two chunks of M at an arbitrary row distance, start on any row
  cudaFuncSetAttribute(probe_mn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int cases[][2] = {{0, 64}, {0, 1}, {1, 1}, {2, 62}, {64, 1}, {65, 63}, {128, 1}, {130, 0}, {3, 8}};
  int mn_ok = 1;
  for (auto& cs : cases) {
    const int shift = cs[0], lbo = cs[1];
    cudaMemset(d_out, 0xFF, 128 * 64 * sizeof(float));
    probe_mn<<<1, 128, smem>>>(d_out, shift, lbo);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("mn shift %d lbo %d: %s\n", shift, lbo, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(h.data(), d_out, h.size() * sizeof(float), cudaMemcpyDeviceToHost);
    int bad = 0, fm = -1, fn = -1;
    float fw = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < 64; ++n) {
        float want = 0.f;
        const int r0 = shift + (m >= 64 ? lbo : 0), c = m & 63;
        for (int i = 0; i < 64; ++i) want += (float)(((r0 + i) * 5 + c * 3) % 7) * (float)((i * 3 + n) % 5);
        if (h[m * 64 + n] != want) { if (!bad) { fm = m; fn = n; fw = want; } ++bad; }
      }
    if (bad) mn_ok = 0;
    printf("MN-major  start row %3d, second chunk %2d rows on: %s (%d of 8192 wrong", shift, lbo, bad ? "MISMATCH" : "ok", bad);
    if (bad) printf("; first at m=%d n=%d got %.0f want %.0f", fm, fn, h[fm * 64 + fn], fw);
    printf(")\n");
  }
  printf("RESULT MN-major: %s\n", mn_ok ? "any start row and any chunk distance work" : "fails");
  return 0;
}
