// Micro-benchmark: how fast can the 8 consumer warps of one SM eat a weight stage that is ALREADY in shared memory?
// (the persistent decode kernel's GEMV phases turned out consumer-bound, not HBM-bound: profiles/r01_mega_trace_v3.md)
// Variants: mma.sync m16n8k16 with one / four accumulator chains, the "two K-halves in one MMA" packing for R <= 4,
// and plain FFMA for R = 2.  Prints the weight bytes per second the whole GPU could consume (148 SMs).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/consume_bench scripts/consume_bench.cu && /tmp/consume_bench
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)
#define ROW_BYTES 4160          // 4096 + 64 pad
#define KC 2048
#define XLD (KC + 32)

__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a));
  return r;
}
__device__ __forceinline__ void mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// one "stage" = 8 weight rows x 2048 bf16 (row stride 4160 B) + activations xs[R][2048+32] bf16
__global__ void __launch_bounds__(256, 1) k_consume(int mode, int iters, float* sink) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < (8 * ROW_BYTES + 8 * XLD * 2) / 4; i += 256) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
  __syncthreads();
  const uint32_t wbase = (uint32_t)__cvta_generic_to_shared(smem);
  const uint32_t xbase = wbase + 8 * ROW_BYTES;
  const int g = lane >> 2, t = lane & 3;
  float out = 0.f;
  if (mode == 0 || mode == 1) {
    // activations = A rows (R=2 real), weights = B (8 rows); warp handles a 256-column slice: 8 k-blocks of 32
    const uint32_t wa = wbase + g * ROW_BYTES + (warp * 256 + t * 8) * 2;
    const uint32_t xa = xbase + (g * XLD + warp * 256 + t * 8) * 2;
    float acc[4][4] = {};
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int j = 0; j < 8; j += 2) {
        const uint4 w0 = lds128(wa + j * 64), w1 = lds128(wa + j * 64 + 64);
        uint4 x0 = make_uint4(0, 0, 0, 0), x1 = x0;
        if (g < 2) { x0 = lds128(xa + j * 64); x1 = lds128(xa + j * 64 + 64); }
        if (mode == 0) {
          mma(acc[0], x0.x, 0, x0.y, 0, w0.x, w0.y); mma(acc[0], x0.z, 0, x0.w, 0, w0.z, w0.w);
          mma(acc[0], x1.x, 0, x1.y, 0, w1.x, w1.y); mma(acc[0], x1.z, 0, x1.w, 0, w1.z, w1.w);
        } else {
          mma(acc[0], x0.x, 0, x0.y, 0, w0.x, w0.y); mma(acc[1], x0.z, 0, x0.w, 0, w0.z, w0.w);
          mma(acc[2], x1.x, 0, x1.y, 0, w1.x, w1.y); mma(acc[3], x1.z, 0, x1.w, 0, w1.z, w1.w);
        }
      }
    }
    for (int a = 0; a < 4; ++a) out += acc[a][0] + acc[a][1];
  } else if (mode == 2) {
    // weights = A: rows 0-7 <- k-half 0, rows 8-15 <- k-half 1 of the SAME 8 weight rows; activations = B:
    // cols 0-3 <- x rows (k-half 0), cols 4-7 <- x rows (k-half 1).  Warp slice: 128 columns of each half.
    const uint32_t wa0 = wbase + g * ROW_BYTES + (warp * 128 + t * 8) * 2;
    const uint32_t wa1 = wa0 + 1024 * 2;
    const int n = g & 3, halfsel = g >> 2;                    // B column g -> activation row n of k-half halfsel
    const uint32_t xa = xbase + (n * XLD + halfsel * 1024 + warp * 128 + t * 8) * 2;
    float acc[2][4] = {};
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint4 w0 = lds128(wa0 + j * 64), w1 = lds128(wa1 + j * 64);
        uint4 x = make_uint4(0, 0, 0, 0);
        if (n < 2) x = lds128(xa + j * 64);
        mma(acc[0], w0.x, w1.x, w0.y, w1.y, x.x, x.y);
        mma(acc[1], w0.z, w1.z, w0.w, w1.w, x.z, x.w);
      }
    }
    out = acc[0][0] + acc[0][1] + acc[1][2] + acc[1][3];
  } else {
    // FFMA, R = 2: lane (g, t) owns weight row g, 8 consecutive columns per step (same smem pattern as mode 0)
    const uint32_t wa = wbase + g * ROW_BYTES + (warp * 256 + t * 8) * 2;
    const uint32_t xa = xbase + (warp * 256 + t * 8) * 2;
    float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint4 w = lds128(wa + j * 64);
        const uint4 x0 = lds128(xa + j * 64), x1 = lds128(xa + XLD * 2 + j * 64);
        a0 = fmaf(lo(w.x), lo(x0.x), a0); a1 = fmaf(hi(w.x), hi(x0.x), a1);
        a0 = fmaf(lo(w.y), lo(x0.y), a0); a1 = fmaf(hi(w.y), hi(x0.y), a1);
        a0 = fmaf(lo(w.z), lo(x0.z), a0); a1 = fmaf(hi(w.z), hi(x0.z), a1);
        a0 = fmaf(lo(w.w), lo(x0.w), a0); a1 = fmaf(hi(w.w), hi(x0.w), a1);
        b0 = fmaf(lo(w.x), lo(x1.x), b0); b1 = fmaf(hi(w.x), hi(x1.x), b1);
        b0 = fmaf(lo(w.y), lo(x1.y), b0); b1 = fmaf(hi(w.y), hi(x1.y), b1);
        b0 = fmaf(lo(w.z), lo(x1.z), b0); b1 = fmaf(hi(w.z), hi(x1.z), b1);
        b0 = fmaf(lo(w.w), lo(x1.w), b0); b1 = fmaf(hi(w.w), hi(x1.w), b1);
      }
    }
    out = a0 + a1 + b0 + b1;
  }
  if (out == 123.456f) sink[0] = out;
}

int main() {
  float* sink;
  CK(cudaMalloc(&sink, 4));
  const size_t smem = 8 * ROW_BYTES + 8 * XLD * 2;
  CK(cudaFuncSetAttribute(k_consume, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const char* names[] = {"HMMA act=A, 1 accumulator chain", "HMMA act=A, 4 accumulator chains", "HMMA weights=A, two K-halves per MMA (R<=4)", "FFMA R=2"};
  const int iters = 20000;
  for (int mode = 0; mode < 4; ++mode) {
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
      CK(cudaEventRecord(e0));
      k_consume<<<148, 256, smem>>>(mode, iters, sink);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best) best = ms;
    }
    const double bytes = 148.0 * iters * 8 * 4096;      // weight bytes consumed by the whole GPU
    printf("%-48s %8.3f ms   %9.1f GB/s of weights  (%.0f ns per 32 KB stage)\n", names[mode], best, bytes / (best * 1e-3) / 1e9, best * 1e6 / iters);
  }
  return 0;
}
