// Probe of tcgen05.cp (shared memory -> tensor memory) semantics on sm_100a: which shared-memory word lands in which
// (TMEM lane, column) for the .128x256b / .128x128b shapes and a no-swizzle matrix descriptor with given LBO / SBO.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tmem_cp_test tmem_cp_test.cu && ./tmem_cp_test
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void k_probe(uint32_t* out, int shape, uint32_t lbo, uint32_t sbo, int version) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) unsigned long long bar;
  uint32_t* w = reinterpret_cast<uint32_t*>(smem);
  for (int i = threadIdx.x; i < 16384; i += blockDim.x) w[i] = i;           // 64 KB: word k holds k
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(32u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1u) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");             // generic-proxy smem writes -> async proxy reader
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = tmem_slot;
  if (threadIdx.x == 0) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_u32(smem) >> 4) & 0x3fff);
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)version << 46;
    if (shape == 0) asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(tm), "l"(d) : "memory");
    else asm volatile("tcgen05.cp.cta_group::1.128x128b [%0], %1;" ::"r"(tm), "l"(d) : "memory");
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
  }
  // everyone waits for the copy
  {
    uint32_t ok = 0;
    while (!ok) {
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    }
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t v[8];
  const uint32_t taddr = tm + ((uint32_t)(warp * 32) << 16);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int j = 0; j < 8; ++j) out[(warp * 32 + lane) * 8 + j] = v[j];
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(32u) : "memory");
}

int main() {
  uint32_t* d_out;
  cudaMalloc(&d_out, 128 * 8 * 4);
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  struct Cfg { int shape; uint32_t lbo, sbo; int version; } cfgs[] = {
      {0, 128, 256, 1}, {0, 256, 128, 1}, {0, 128, 256, 0}, {0, 16, 1024, 1}, {1, 128, 128, 1}, {1, 16, 256, 1}};
  for (const Cfg& c : cfgs) {
    cudaMemset(d_out, 0xff, 128 * 8 * 4);
    k_probe<<<1, 128, 65536>>>(d_out, c.shape, c.lbo, c.sbo, c.version);
    cudaError_t e = cudaDeviceSynchronize();
    uint32_t h[128 * 8];
    cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
    printf("shape %s lbo %u sbo %u version %d: %s\n", c.shape == 0 ? "128x256b" : "128x128b", c.lbo, c.sbo, c.version, cudaGetErrorString(e));
    const int lanes[] = {0, 1, 2, 7, 8, 9, 15, 16, 31, 32, 33, 64, 127};
    for (int l : lanes) {
      printf("  lane %3d: byte offsets", l);
      for (int j = 0; j < 8; ++j) printf(" %6d", (int)h[l * 8 + j] == -1 ? -1 : (int)(h[l * 8 + j] * 4));
      printf("\n");
    }
  }
  return 0;
}
