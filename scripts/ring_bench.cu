// Micro-benchmark for the weight-streaming ring of the persistent decode kernel (mega_kernel.cuh): how fast can one
// producer warp per SM pull HBM through cp.async.bulk + mbarriers, as a function of copy granularity, ring depth,
// consumer behaviour and polling style?  Build + run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ring_bench scripts/ring_bench.cu && /tmp/ring_bench
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint32_t b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(b) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint32_t b, uint32_t par) {
  uint32_t ok;
  asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(b), "r"(par) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void bulk(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
  const uint32_t z = 0u;
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(z), "r"(a2), "r"(z), "r"(b0), "r"(b1));
}

struct P {
  const unsigned char* src;   // total bytes = grid * per_cta
  size_t per_cta;             // bytes per CTA (multiple of stage payload)
  int n_stages;
  int rows;                   // copies per stage (1 or 8); payload per stage = rows * row_bytes
  int row_bytes;              // bytes per copy
  int row_stride;             // smem stride between rows
  size_t src_row_stride;      // global stride between the rows of one stage (row-major matrix with K*2 bytes per row)
  int mode;                   // 0: wait+release, 1: + LDS everything, 2: LDS + MMA (decode-like)
  int poll_all;               // 1: all lanes poll the full barrier, 0: lane 0 polls + __syncwarp
  int direct;                 // 1: no ring — consumers ld.global.nc straight to registers (8 x 16 B in flight per thread)
  long long wrap_iters;        // > 0: the producer re-reads the first wrap_iters stages over and over (L2-resident source)
  long long n_iter;            // > 0: stages per CTA (with wrap_iters)
  float* sink;
};

__global__ void __launch_bounds__(288, 1) k_ring(const P p) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int stage_bytes = p.rows * p.row_stride;
  unsigned char* ring = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)p.n_stages * stage_bytes);
  const uint32_t full0 = s32(bars), empty0 = full0 + 8u * p.n_stages;
  if (tid == 0) {
    for (int s = 0; s < p.n_stages; ++s) { mbar_init(full0 + 8u * s, 1); mbar_init(empty0 + 8u * s, 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int payload = p.rows * p.row_bytes;
  const long long n_iter = p.n_iter > 0 ? p.n_iter : (long long)(p.per_cta / payload);
  const unsigned char* base = p.src + (size_t)blockIdx.x * p.per_cta;
  if (p.direct) {
    if (warp == 8) return;
    // each thread: 16 B per load; CTA covers 256*16 = 4 KB per "row"; 8 loads in flight
    float acc = 0.f;
    const size_t n16 = p.per_cta / 16;
    for (size_t i = tid; i < n16; i += 256 * 8) {
      uint4 v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const size_t k = i + (size_t)j * 256;
        if (k < n16) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[j].x), "=r"(v[j].y), "=r"(v[j].z), "=r"(v[j].w) : "l"(base + k * 16));
        else v[j] = make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) acc += __uint_as_float(v[j].x ^ v[j].y ^ v[j].z ^ v[j].w);
    }
    if (acc == 123.456f) p.sink[0] = acc;
    return;
  }
  if (warp == 8) {
    int st = 0, par = 0;
    for (long long it0 = 0; it0 < n_iter; ++it0) {
      const long long it = p.wrap_iters > 0 ? it0 % p.wrap_iters : it0;
      if (lane == 0) {
        while (!mbar_try(empty0 + 8u * st, par ^ 1)) {}
        mbar_expect(full0 + 8u * st, payload);
      }
      __syncwarp();
      // stage `it` covers rows [it*rows, ...) of a [n][src_row_stride] matrix, column chunk handled by linearisation:
      const unsigned char* s = (p.rows == 1) ? base + (size_t)it * payload
                                             : base + ((size_t)(it / (p.src_row_stride / p.row_bytes)) * p.rows) * p.src_row_stride +
                                                   (size_t)(it % (p.src_row_stride / p.row_bytes)) * p.row_bytes;
      if (lane < p.rows) bulk(s32(ring + (size_t)st * stage_bytes + lane * p.row_stride), s + (size_t)lane * p.src_row_stride, p.row_bytes, full0 + 8u * st);
      if (++st == p.n_stages) { st = 0; par ^= 1; }
    }
    return;
  }
  int st = 0, par = 0;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  float fsum = 0.f;
  const int g = lane >> 2, t = lane & 3;
  for (long long it = 0; it < n_iter; ++it) {
    if (p.poll_all) { while (!mbar_try(full0 + 8u * st, par)) {} }
    else { if (lane == 0) { while (!mbar_try(full0 + 8u * st, par)) {} } __syncwarp(); }
    const unsigned char* sp = ring + (size_t)st * stage_bytes;
    if (p.mode == 1) {
      for (int o = tid * 16; o < payload; o += 256 * 16) {
        const int r = o / p.row_bytes, c = o % p.row_bytes;
        const uint4 v = *reinterpret_cast<const uint4*>(sp + r * p.row_stride + c);
        fsum += __uint_as_float(v.x ^ v.y ^ v.z ^ v.w);
      }
    } else if (p.mode == 2) {
      // decode-like: 8 rows x row_bytes; warp w takes row_bytes/8 bytes of each row; lane (g,t) reads row g
      const int slice = p.row_bytes / 8;            // bytes per warp per row
      const int rows = p.rows == 1 ? 8 : p.rows;
      const int rstride = p.rows == 1 ? p.row_bytes / 8 : p.row_stride;
      const int sl = p.rows == 1 ? slice / 8 : slice;
      const unsigned char* wp = sp + (g % rows) * rstride + warp * sl + t * 16;
      for (int j = 0; j < sl / 64; ++j) {
        const uint4 wv = *reinterpret_cast<const uint4*>(wp + j * 64);
        mma16816(acc, 0x3f803f80u, 0x3f803f80u, wv.x, wv.y);
        mma16816(acc, 0x3f803f80u, 0x3f803f80u, wv.z, wv.w);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty0 + 8u * st);
    if (++st == p.n_stages) { st = 0; par ^= 1; }
  }
  fsum += acc[0] + acc[1] + acc[2] + acc[3];
  if (fsum == 123.456f) p.sink[0] = fsum;
}

int main() {
  const size_t total = (size_t)148 * 16 * 1024 * 1024;       // 2.3 GB > L2
  unsigned char* buf;
  float* sink;
  CK(cudaMalloc(&buf, total));
  CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(buf, 1, total));
  CK(cudaFuncSetAttribute(k_ring, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  struct Cfg { const char* name; int rows, row_bytes, row_stride, n_stages, mode, poll_all, direct; size_t src_stride; };
  const Cfg cfgs[] = {
      {"direct ld.global.nc 8x16B/thread", 1, 16384, 16384, 2, 0, 0, 1, 0},
      {"1x16KB  st12 mode0", 1, 16384, 16384, 12, 0, 0, 0, 0},
      {"1x16KB  st6  mode0", 1, 16384, 16384, 6, 0, 0, 0, 0},
      {"1x16KB  st3  mode0", 1, 16384, 16384, 3, 0, 0, 0, 0},
      {"1x32KB  st6  mode0", 1, 32768, 32768, 6, 0, 0, 0, 0},
      {"1x8KB   st24 mode0", 1, 8192, 8192, 24, 0, 0, 0, 0},
      {"8x2KB(K=2048) st12 mode0", 8, 2048, 2112, 12, 0, 0, 0, 4096},
      {"8x2KB(K=8192) st12 mode0", 8, 2048, 2112, 12, 0, 0, 0, 16384},
      {"8x2KB(K=2048) st12 mode2 poll_all", 8, 2048, 2112, 12, 2, 1, 0, 4096},
      {"8x2KB(K=2048) st12 mode2 poll_lane0", 8, 2048, 2112, 12, 2, 0, 0, 4096},
      {"8x2KB(K=2048) st12 mode1 poll_lane0", 8, 2048, 2112, 12, 1, 0, 0, 4096},
      {"8x4KB(K=2048) st6 mode2 poll_lane0", 8, 4096, 4160, 6, 2, 0, 0, 4096},
      {"8x4KB(K=2048) st6 mode0", 8, 4096, 4160, 6, 0, 0, 0, 4096},
      {"1x16KB  st12 mode2 poll_lane0", 1, 16384, 16384, 12, 2, 0, 0, 0},
      {"8x4KB st5 mode2 poll_lane0", 8, 4096, 4160, 5, 2, 0, 0, 4096},
      {"8x4KB st5 mode2 poll_all", 8, 4096, 4160, 5, 2, 1, 0, 4096},
      {"8x4KB st4 mode2 poll_lane0", 8, 4096, 4160, 4, 2, 0, 0, 4096},
      {"8x4KB st3 mode2 poll_lane0", 8, 4096, 4160, 3, 2, 0, 0, 4096},
      {"8x4KB st2 mode2 poll_lane0", 8, 4096, 4160, 2, 2, 0, 0, 4096},
      {"8x4KB st5 mode0 poll_lane0", 8, 4096, 4160, 5, 0, 0, 0, 4096},
      {"16x4KB st3 mode2 poll_lane0", 16, 4096, 4160, 3, 2, 0, 0, 4096},
      {"16x4KB st2 mode2 poll_lane0", 16, 4096, 4160, 2, 2, 0, 0, 4096},
      {"8x8KB(K=8192) st3 mode2 poll_lane0", 8, 8192, 8256, 3, 2, 0, 0, 16384},
      {"8x8KB(K=8192) st2 mode2 poll_lane0", 8, 8192, 8256, 2, 2, 0, 0, 16384},
      {"8x4KB(K=8192) st5 mode2 poll_lane0", 8, 4096, 4160, 5, 2, 0, 0, 16384},
  };
  for (const Cfg& c : cfgs) {
    P p;
    p.src = buf; p.per_cta = total / 148; p.n_stages = c.n_stages; p.rows = c.rows; p.row_bytes = c.row_bytes; p.row_stride = c.row_stride;
    p.src_row_stride = c.src_stride; p.mode = c.mode; p.poll_all = c.poll_all; p.direct = c.direct; p.sink = sink; p.wrap_iters = 0; p.n_iter = 0;
    const size_t smem = (size_t)c.n_stages * c.rows * c.row_stride + 16 * c.n_stages + 64;
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
      CK(cudaEventRecord(e0));
      k_ring<<<148, 288, smem>>>(p);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      CK(cudaGetLastError());
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best) best = ms;
    }
    printf("%-42s smem %6zu  %8.3f ms  %8.1f GB/s\n", c.name, smem, best, total / (best * 1e-3) / 1e9);
    fflush(stdout);
  }
  // ---- L2-resident source: every CTA loops over its own first `wrap` stages (148 x wrap x 32 KB in total), so after the
  // first pass the ring is fed from L2: the ceiling of an L2-prefetched weight stream
  for (int wrap : {4, 8, 16, 32}) {
    for (int mode : {0, 2}) {
      P p;
      p.src = buf; p.per_cta = total / 148; p.n_stages = 5; p.rows = 8; p.row_bytes = 4096; p.row_stride = 4160;
      p.src_row_stride = 4096; p.mode = mode; p.poll_all = 0; p.direct = 0; p.sink = sink; p.wrap_iters = wrap; p.n_iter = 4096;
      const size_t smem = (size_t)5 * 8 * 4160 + 16 * 5 + 64;
      float best = 1e30f;
      for (int rep = 0; rep < 3; ++rep) {
        CK(cudaEventRecord(e0));
        k_ring<<<148, 288, smem>>>(p);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
      }
      const double bytes = 148.0 * 4096 * 32768;
      printf("L2-resident 8x4KB st5 mode%d wrap %2d (%5.1f MB set)   %8.3f ms  %8.1f GB/s\n", mode, wrap, 148.0 * wrap * 32768 / 1e6, best, bytes / (best * 1e-3) / 1e9);
      fflush(stdout);
    }
  }
  // ---- how much can ONE SM take in?  The same ring on fewer CTAs: if the per-SM rate stays at total / 148 the limit is the SM's
  // own ingest path, if it rises the limit is the fabric / memory in aggregate (decides whether fatter tiles on fewer CTAs can
  // pay for the batched GEMMs, whose per-CTA traffic is weights + activation tile).
  for (int grid : {148, 74, 37, 16, 4, 1}) {
    for (int l2 = 0; l2 < 2; ++l2) {
      P p;
      p.src = buf; p.per_cta = total / 148; p.n_stages = 5; p.rows = 8; p.row_bytes = 4096; p.row_stride = 4160;
      p.src_row_stride = 4096; p.mode = 0; p.poll_all = 0; p.direct = 0; p.sink = sink; p.wrap_iters = l2 ? 8 : 0; p.n_iter = l2 ? 4096 : 0;
      const size_t smem = (size_t)5 * 8 * 4160 + 16 * 5 + 64;
      float best = 1e30f;
      for (int rep = 0; rep < 3; ++rep) {
        CK(cudaEventRecord(e0));
        k_ring<<<grid, 288, smem>>>(p);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
      }
      const double bytes = l2 ? (double)grid * 4096 * 32768 : (double)grid * (double)(total / 148);
      printf("grid %3d CTAs, source %s: %8.3f ms  %8.1f GB/s total  %7.1f GB/s per SM\n", grid, l2 ? "L2 (8 stages per CTA)" : "HBM", best,
             bytes / (best * 1e-3) / 1e9, bytes / (best * 1e-3) / 1e9 / grid);
      fflush(stdout);
    }
  }
  return 0;
}
