// Probe: may a tcgen05 SWIZZLE_128B K-major operand descriptor start at an arbitrary ROW of a tile (start address
// base + r * 128 B, r not a multiple of 8)?  The DAC k=7 convolutions would then load their input window ONCE (with a
// halo) and issue the 7 taps as 7 row-shifted descriptors instead of 7 TMA loads of almost the same tile.
// Two variants per offset: descriptor base_offset field = 0, and = (addr >> 7) & 7.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I zonos_vibes_b200/csrc -o /tmp/desc_test scripts/desc_offset_test.cu -lcuda
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include "tc_common.cuh"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)
#define A_ROWS 160
#define N_T 64

// A: [A_ROWS][64] fp16 row-major in global; B: [N_T][64]; out: [128][N_T] fp32
__global__ void __launch_bounds__(128) k_probe(const __half* A, const __half* B, float* out, int off, int use_base_offset) {
  extern __shared__ unsigned char raw_smem[];
  const uint32_t raw = smem_u32(raw_smem);
  const uint32_t base = (raw + 1023u) & ~1023u;
  unsigned char* sm = raw_smem + (base - raw);
  unsigned char* sa = sm;                          // A_ROWS * 128 B
  unsigned char* sb = sm + A_ROWS * 128;           // 20480 is a multiple of 1024
  uint64_t* bar = reinterpret_cast<uint64_t*>(sb + N_T * 128);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // manual 128B swizzle: 16-byte chunk c of row r goes to chunk c ^ (r & 7)
  for (int i = tid; i < A_ROWS * 8; i += 128) {
    const int r = i >> 3, c = i & 7;
    *reinterpret_cast<uint4*>(sa + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const uint4*>(A + r * 64 + c * 8);
  }
  for (int i = tid; i < N_T * 8; i += 128) {
    const int r = i >> 3, c = i & 7;
    *reinterpret_cast<uint4*>(sb + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const uint4*>(B + r * 64 + c * 8);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (tid == 0) { mbar_init(smem_u32(bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(64u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *reinterpret_cast<volatile uint32_t*>(slot);
  if (tid == 0) {
    const uint32_t a_addr = base + (uint32_t)off * 128u;
    uint64_t adesc = umma_desc_sw128(a_addr);
    if (use_base_offset) adesc |= (uint64_t)((a_addr >> 7) & 7u) << 49;
    const uint64_t bdesc = umma_desc_sw128(base + A_ROWS * 128);
    const uint32_t idesc = umma_idesc_f16(N_T);
    for (int k = 0; k < 4; ++k) tc_mma_f16(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, k != 0);
    tc_commit(smem_u32(bar));
  }
  __syncwarp();
  mbar_wait(smem_u32(bar), 0);
  tc_fence_after();
  for (int c0 = 0; c0 < N_T; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
    for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * N_T + c0 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(64u) : "memory");
}

int main() {
  __half *hA = (__half*)malloc(A_ROWS * 64 * 2), *hB = (__half*)malloc(N_T * 64 * 2);
  float* fA = (float*)malloc(A_ROWS * 64 * 4); float* fB = (float*)malloc(N_T * 64 * 4);
  srand(1);
  for (int i = 0; i < A_ROWS * 64; ++i) { hA[i] = __float2half((rand() % 17 - 8) / 8.0f); fA[i] = __half2float(hA[i]); }
  for (int i = 0; i < N_T * 64; ++i) { hB[i] = __float2half((rand() % 13 - 6) / 4.0f); fB[i] = __half2float(hB[i]); }
  __half *dA, *dB; float* dO;
  CK(cudaMalloc(&dA, A_ROWS * 64 * 2)); CK(cudaMalloc(&dB, N_T * 64 * 2)); CK(cudaMalloc(&dO, 128 * N_T * 4));
  CK(cudaMemcpy(dA, hA, A_ROWS * 64 * 2, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, hB, N_T * 64 * 2, cudaMemcpyHostToDevice));
  const int smem = 1024 + A_ROWS * 128 + N_T * 128 + 64;
  CK(cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  float* hO = (float*)malloc(128 * N_T * 4);
  for (int variant = 0; variant < 2; ++variant)
    for (int off = 0; off <= 27; ++off) {
      CK(cudaMemset(dO, 0, 128 * N_T * 4));
      k_probe<<<1, 128, smem>>>(dA, dB, dO, off, variant);
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(hO, dO, 128 * N_T * 4, cudaMemcpyDeviceToHost));
      double maxerr = 0;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N_T; ++n) {
          float ref = 0;
          for (int k = 0; k < 64; ++k) ref += fA[(m + off) * 64 + k] * fB[n * 64 + k];
          maxerr = fmax(maxerr, fabs(ref - hO[m * N_T + n]));
        }
      printf("base_offset_field=%s off=%2d max_abs_err=%g %s\n", variant ? "(addr>>7)&7" : "0", off, maxerr, maxerr < 1e-3 ? "OK" : "WRONG");
    }
  return 0;
}
