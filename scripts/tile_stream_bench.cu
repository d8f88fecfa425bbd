// Micro-benchmark: how fast does a weight matrix [N][K] bf16 stream from HBM into shared memory as a function of HOW the
// tiles are requested?  Question behind it (profiles/r02_summary.md section 3): k_bt_gemm's fc1 runs at 2.9 TB/s of weights and
// halving the activation bytes per SM (k_bt_gemm2) did not speed it up, so the limit is not the bytes delivered per SM.
//   mode 0: one TMA tensor box {64 k, 128 rows} per stage (what k_bt_gemm does: 128 segments of 128 B, 2*K bytes apart)
//   mode 1: two boxes (k, k + 64) per stage: 256 contiguous bytes per row
//   mode 2: four boxes per stage: 512 contiguous bytes per row
//   mode 3: the same bytes as one contiguous 16 KB cp.async.bulk per stage (a tile-blocked weight layout)
//   mode 4: contiguous 32 KB per stage
// One producer thread per CTA, one consumer thread that only waits and releases.  grid = N / 128 CTAs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tile_stream_bench scripts/tile_stream_bench.cu -lcuda && /tmp/tile_stream_bench
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint32_t b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(b) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t b, uint32_t par) {
  uint32_t ok = 0;
  while (!ok) asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(b), "r"(par) : "memory");
}
__device__ __forceinline__ void bulk(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst), "l"(map), "r"(c0),
               "r"(c1), "r"(bar) : "memory");
}

struct P {
  const unsigned char* w;     // matrix base (mode 3 / 4: read as a flat byte stream, 128 * K * 2 bytes per CTA)
  int K, mode, n_stages, rows;   // rows per box (128 or 256)
  int iters;                  // passes over the CTA's tile row (each pass uses the next matrix copy)
  size_t copy_stride;         // bytes between matrix copies
  int n_copies;
};

__global__ void __launch_bounds__(64, 1) k_stream(const __grid_constant__ CUtensorMap m0, const __grid_constant__ CUtensorMap m1,
                                                  const __grid_constant__ CUtensorMap m2, const __grid_constant__ CUtensorMap m3, const P p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t base = (s32(smem) + 1023u) & ~1023u;
  const int boxes = p.mode == 0 ? 1 : p.mode == 1 ? 2 : p.mode == 2 ? 4 : 0;
  const int box_bytes = p.rows * 128;
  const int stage_bytes = p.mode <= 2 ? boxes * box_bytes : (p.mode == 3 ? 16384 : 32768);
  const uint32_t bars = base + (uint32_t)p.n_stages * stage_bytes;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.n_stages; ++s) { mbar_init(bars + 8u * s, 1); mbar_init(bars + 8u * (p.n_stages + s), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int per_pass = (int)((long long)p.rows * p.K * 2 / stage_bytes);     // stages per pass over this CTA's tile row
  const int n = per_pass * p.iters;
  const CUtensorMap* maps[4] = {&m0, &m1, &m2, &m3};
  // no divisions in the loops: the producer of a real kernel spends a few dozen cycles per stage, and so must this one
  if (threadIdx.x == 0) {
    int s = 0, round = 0, j = 0, pass = 0;
    for (int i = 0; i < n; ++i) {
      if (round > 0) mbar_wait(bars + 8u * (p.n_stages + s), (uint32_t)(round - 1) & 1);
      const uint32_t dst = base + (uint32_t)s * stage_bytes, bar = bars + 8u * s;
      mbar_expect(bar, (uint32_t)stage_bytes);
      if (p.mode <= 2) {
        for (int b = 0; b < boxes; ++b) tma2d(dst + b * box_bytes, maps[pass], (j * boxes + b) * 64, blockIdx.x * p.rows, bar);
      } else {
        const unsigned char* src = p.w + (size_t)pass * p.copy_stride + (size_t)blockIdx.x * p.rows * p.K * 2 + (size_t)j * stage_bytes;
        bulk(dst, src, (uint32_t)stage_bytes, bar);
      }
      if (++s == p.n_stages) { s = 0; ++round; }
      if (++j == per_pass) { j = 0; if (++pass == p.n_copies) pass = 0; }
    }
  } else if (threadIdx.x == 32) {
    int s = 0, round = 0;
    for (int i = 0; i < n; ++i) {
      mbar_wait(bars + 8u * s, (uint32_t)round & 1);
      mbar_arrive(bars + 8u * (p.n_stages + s));
      if (++s == p.n_stages) { s = 0; ++round; }
    }
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                             CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int N = 16384, n_copies = 4;
  EncodeFn encode = nullptr;
  cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &qr));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  CK(cudaFuncSetAttribute(k_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  for (int K : {2048, 8192}) {
    const int Nn = K == 2048 ? N : N / 4;                     // same bytes: fc1-like [16384][2048], fc2-like [4096][8192]
    const size_t bytes = (size_t)Nn * K * 2;
    unsigned char* w;
    CK(cudaMalloc(&w, bytes * n_copies));
    CK(cudaMemset(w, 1, bytes * n_copies));
    for (int rows : {128, 256}) {
      CUtensorMap maps[4];
      for (int c = 0; c < n_copies; ++c) {
        cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)Nn};
        cuuint64_t strides[1] = {(cuuint64_t)K * 2};
        cuuint32_t box[2] = {64, (cuuint32_t)rows};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = encode(&maps[c], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w + c * bytes, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
      }
      for (int mode = 0; mode <= 4; ++mode) {
        if (rows == 256 && mode >= 2) continue;
        for (int depth_kb : {96, 192}) {
          P p;
          p.w = w; p.K = K; p.mode = mode; p.rows = rows; p.iters = 8; p.copy_stride = bytes; p.n_copies = n_copies;
          const int stage_bytes = mode <= 2 ? (mode == 0 ? 1 : mode == 1 ? 2 : 4) * rows * 128 : (mode == 3 ? 16384 : 32768);
          p.n_stages = depth_kb * 1024 / stage_bytes;
          if (p.n_stages < 2) continue;
          const int grid = Nn / rows;
          const size_t smem = (size_t)p.n_stages * stage_bytes + 2048;
          k_stream<<<grid, 64, smem>>>(maps[0], maps[1], maps[2], maps[3], p);      // warm-up
          CK(cudaDeviceSynchronize());
          CK(cudaEventRecord(e0));
          k_stream<<<grid, 64, smem>>>(maps[0], maps[1], maps[2], maps[3], p);
          CK(cudaEventRecord(e1));
          CK(cudaDeviceSynchronize());
          float ms;
          CK(cudaEventElapsedTime(&ms, e0, e1));
          const double gb = (double)bytes * p.iters / 1e9;
          printf("K=%5d rows/box=%3d mode=%d stage=%6d B x %2d stages grid=%3d : %7.1f us per pass, %7.1f GB/s (%5.1f GB/s per CTA)\n", K, rows, mode, stage_bytes,
                 p.n_stages, grid, ms * 1e3 / p.iters, gb / (ms * 1e-3), gb / (ms * 1e-3) / grid);
        }
      }
    }
    CK(cudaFree(w));
  }
  return 0;
}
