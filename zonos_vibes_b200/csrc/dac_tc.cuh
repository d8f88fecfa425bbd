// DAC 44 kHz decoder — tcgen05 implicit-GEMM convolutions (impl=0), fp16 operands, fp32 accumulation in TMEM.
//
// Every Conv1d / ConvTranspose1d of the decoder (modeling_dac.py:173-207,234-262,405-439) is the same multi-tap GEMM:
//     D[t, co] = sum_tap sum_ci  X[t + off(tap), ci] * W[tap][co, ci]
// with activations stored TIME-MAJOR [T][C] fp16.  A tap is a row shift of the activation matrix, so the A operand
// of every (tap, 64-channel chunk) k-block is one TMA box {64 ch, 128 t} at coordinate (c0, t0 + off) — TMA's
// out-of-bounds zero fill implements the conv's zero padding and the ragged last channel chunk (C=96) for free.
// The B operand is a box {64 ch, N} of the repacked weights [tap][Cout][Cin].  Both land in shared memory in the
// canonical K-major SWIZZLE_128B layout and feed `tcgen05.mma.cta_group::1.kind::f16` (M=128 time rows, N<=256
// output channels, K=16 per instruction); the accumulator lives in TMEM and is drained by 4 epilogue warps
// (`tcgen05.ld.32x32b.x32`) which fuse bias, residual add, the NEXT layer's Snake activation and the fp16 store.
// A ConvTranspose1d (k=2s, stride s) is s independent 2-tap GEMMs, one per output phase (grid.z).
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2..5 = epilogue.
// One output tile per CTA; two CTAs co-reside per SM so one tile's epilogue overlaps the other's main loop.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

#include "zb_common.cuh"

#define TC_THREADS 192
#define TC_BM 128
#define TC_BK 64
#define TC_A_BYTES (TC_BM * TC_BK * 2)     // 16 KB
#define TC_MAX_TAPS 8

struct TcConvParams {
  int ntaps;                 // conv: K taps; transposed conv: 2 taps per phase
  int dil, pad, stride;      // stride > 1 => transposed conv with K = 2*stride, grid.z = phase
  int kchunks;               // ceil(Cin / 64)
  int m_rows;                // GEMM rows (conv: T_out; convT: T_in + 1)
  int T_out;                 // output rows
  int Cout;                  // total output channels (row stride of the outputs)
  int n_tile;                // output channels per CTA (multiple of 32, <= 256)
  int tmem_cols;             // power of two >= n_tile
  int stages;
  const float* bias;         // [Cout]
  const __half* residual;    // [T_out][Cout] or NULL
  __half* out_raw;           // [T_out][Cout] or NULL   (acc + bias + residual)
  __half* out_act;           // [T_out][Cout] or NULL   snake(out_raw, alpha_next)
  const float* alpha_next;   // [Cout] (needed when out_act)
};

// ---- PTX wrappers -------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
               "l"(map), "r"(c0), "r"(c1), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start>>4 [0,14), LBO>>4 [16,30),
// SBO>>4 [32,46), version=1 [46,48), layout_type=2 (SWIZZLE_128B) [61,64).  8-row groups are 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::f16 instruction descriptor (cute::UMMA::InstrDescriptor): D=F32, A=B=F16, both K-major, M=128, N=n.
__device__ __forceinline__ uint32_t umma_idesc_f16(int n) {
  uint32_t d = 0;
  d |= 1u << 4;                       // c_format = F32
  d |= 0u << 7;                       // a_format = F16
  d |= 0u << 10;                      // b_format = F16
  d |= (uint32_t)(n >> 3) << 17;      // n_dim
  d |= (uint32_t)(TC_BM >> 4) << 24;  // m_dim
  return d;
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
        "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ float snake_fast(float x, float alpha, float inv_alpha) {
  const float s = __sinf(alpha * x);
  return fmaf(inv_alpha, s * s, x);
}

// grid = (ceil(m_rows/128), Cout/n_tile, phases); block = 192; dyn smem = 1024 (align) + stages*(16K + n_tile*128) + 256
__global__ void __launch_bounds__(TC_THREADS) k_dac_conv_tc(const __grid_constant__ CUtensorMap map_x,
                                                            const __grid_constant__ CUtensorMap map_w, const TcConvParams p) {
  extern __shared__ unsigned char tc_smem_raw[];
  const uint32_t raw = smem_u32(tc_smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;                 // SWIZZLE_128B tiles need 1024 B alignment
  const uint32_t b_bytes = (uint32_t)p.n_tile * 128u;
  const uint32_t stage_bytes = TC_A_BYTES + b_bytes;
  const uint32_t bar_base = base + (uint32_t)p.stages * stage_bytes;      // 8-byte aligned (stage_bytes % 1024 == 0)
  auto full_bar = [&](int s) { return bar_base + 8u * (uint32_t)s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (uint32_t)(p.stages + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (uint32_t)(2 * p.stages);
  const uint32_t tmem_slot = tmem_full_bar + 8u;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(tc_smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * TC_BM;
  const int n0 = blockIdx.y * p.n_tile;
  const int phase = blockIdx.z;
  const int num_kb = p.ntaps * p.kchunks;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)p.tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot_ptr;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer =====
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % p.stages;
        if (kb >= p.stages) mbar_wait(empty_bar(s), ((kb / p.stages) - 1) & 1);
        const int tap = kb / p.kchunks, kc = kb % p.kchunks;
        int row_off, w_row;
        if (p.stride > 1) { row_off = -tap; w_row = (phase + tap * p.stride) * p.Cout; }
        else { row_off = tap * p.dil - p.pad; w_row = tap * p.Cout; }
        const uint32_t a_dst = base + (uint32_t)s * stage_bytes;
        const uint32_t b_dst = a_dst + TC_A_BYTES;
        mbar_expect_tx(full_bar(s), stage_bytes);
        tma_load_2d(a_dst, &map_x, kc * TC_BK, m0 + row_off, full_bar(s));
        tma_load_2d(b_dst, &map_w, kc * TC_BK, w_row + n0, full_bar(s));
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer (single thread) =====
      const uint32_t idesc = umma_idesc_f16(p.n_tile);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % p.stages;
        mbar_wait(full_bar(s), (kb / p.stages) & 1);
        tc_fence_after();
        const uint32_t a_addr = base + (uint32_t)s * stage_bytes;
        const uint64_t adesc = umma_desc_sw128(a_addr);
        const uint64_t bdesc = umma_desc_sw128(a_addr + TC_A_BYTES);
#pragma unroll
        for (int k = 0; k < TC_BK / 16; ++k)     // +32 B per K=16 slice inside the 128 B swizzle atom
          tc_mma_f16(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0);
        tc_commit(empty_bar(s));                  // frees the smem stage when these MMAs retire
      }
      tc_commit(tmem_full_bar);                   // accumulator complete
    }
    __syncwarp();
  } else {
    // ===== epilogue: warp w owns TMEM lanes [32*(w%4), +32) =====
    const int lg = warp & 3;
    const int m = m0 + lg * 32 + lane;
    int out_row;
    if (p.stride > 1) out_row = m * p.stride + phase - p.pad; else out_row = m;
    const bool valid = (m < p.m_rows) && (out_row >= 0) && (out_row < p.T_out);
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const size_t row_base = (size_t)(valid ? out_row : 0) * p.Cout + n0;
    for (int c0 = 0; c0 < p.n_tile; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, v);
      if (valid) {
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) + __ldg(p.bias + n0 + c0 + j);
        if (p.residual) {
          const uint4* rp = reinterpret_cast<const uint4*>(p.residual + row_base + c0);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint4 r = rp[q];
            const __half2* h = reinterpret_cast<const __half2*>(&r);
#pragma unroll
            for (int j = 0; j < 4; ++j) { const float2 t = __half22float2(h[j]); f[q * 8 + 2 * j] += t.x; f[q * 8 + 2 * j + 1] += t.y; }
          }
        }
        if (p.out_raw) {
          uint4* op = reinterpret_cast<uint4*>(p.out_raw + row_base + c0);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            __half2 h[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) h[j] = __floats2half2_rn(f[q * 8 + 2 * j], f[q * 8 + 2 * j + 1]);
            op[q] = *reinterpret_cast<uint4*>(h);
          }
        }
        if (p.out_act) {
          uint4* op = reinterpret_cast<uint4*>(p.out_act + row_base + c0);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            __half2 h[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float a0 = __ldg(p.alpha_next + n0 + c0 + q * 8 + 2 * j), a1 = __ldg(p.alpha_next + n0 + c0 + q * 8 + 2 * j + 1);
              h[j] = __floats2half2_rn(snake_fast(f[q * 8 + 2 * j], a0, 1.0f / (a0 + 1e-9f)),
                                       snake_fast(f[q * 8 + 2 * j + 1], a1, 1.0f / (a1 + 1e-9f)));
            }
            op[q] = *reinterpret_cast<uint4*>(h);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)p.tmem_cols) : "memory");
  }
}

// Final conv (Cin -> 1 channel, k taps) + tanh: pure bandwidth, SIMT.  x fp16 [T][C] (already snaked), w fp32 [K][C].
__global__ void __launch_bounds__(256) k_dac_final_conv(const __half* __restrict__ x, const float* __restrict__ w, float bias,
                                                        float* __restrict__ out, int T, int C, int K, int pad) {
  // one warp per output sample; lanes over (tap, channel)
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nw = (gridDim.x * blockDim.x) >> 5;
  for (int t = warp; t < T; t += nw) {
    float acc = 0.f;
    for (int i = lane; i < K * C; i += 32) {
      const int j = i / C, c = i % C;
      const int tt = t + j - pad;
      if (tt >= 0 && tt < T) acc = fmaf(__half2float(x[(size_t)tt * C + c]), w[i], acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) out[t] = tanhf(acc + bias);
  }
}

// fp32 [T][C] -> fp16 [T][C] with optional snake (used to feed SIMT-produced activations into the TC path in tests)
__global__ void k_dac_f32_to_f16(const float* __restrict__ x, __half* __restrict__ out, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) out[i] = __float2half_rn(x[i]);
}
__global__ void k_dac_f16_to_channel_major(const __half* __restrict__ x, float* __restrict__ out, int T, int C) {
  const size_t n = (size_t)T * C;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const size_t t = i / C;
    out[(size_t)c * T + t] = __half2float(x[i]);
  }
}
