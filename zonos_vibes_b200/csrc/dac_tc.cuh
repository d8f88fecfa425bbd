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
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2..9 = epilogue (two per TMEM lane quarter).
// One output tile per CTA; two CTAs co-reside per SM so one tile's epilogue overlaps the other's main loop.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

#include "zb_common.cuh"

#define TC_THREADS 320                      // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue (two warps per TMEM lane quarter)
#define TC_BM 128
#define TC_BK 64
#define TC_A_BYTES (TC_BM * TC_BK * 2)     // 16 KB
#define TC_MAX_TAPS 8
#define TC_TAB_BYTES (3 * 256 * 4)          // bias / alpha / 1/(alpha + 1e-9) of the CTA's <= 256 output channels

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
  int m_tiles;               // 128-row tiles per utterance: grid.x = nb * m_tiles (utterance b = blockIdx.x / m_tiles)
  const float* bias;         // [Cout]
  const __half* residual;    // [T_out][Cout] or NULL
  __half* out_raw;           // [T_out][Cout] or NULL   (acc + bias + residual)
  __half* out_act;           // [T_out][Cout] or NULL   snake(out_raw, alpha_next)
  const float* alpha_next;   // [Cout] (needed when out_act)
};

#include "tc_common.cuh"

__device__ __forceinline__ float snake_fast(float x, float alpha, float inv_alpha) {
  const float s = __sinf(alpha * x);
  return fmaf(inv_alpha, s * s, x);
}

// ---- shared epilogue: one 32-column chunk of one accumulator row per thread ------------------------------------
// Thread `lane` holds row `lane` of the chunk (tcgen05.ld 32x32b).  A row-per-thread 16-byte access touches 32
// different sectors per warp instruction and uses half of each (ncu: "16 of 32 bytes per sector", 50 % excessive
// sectors on the k=1 layers).  Lanes 2i / 2i+1 therefore swap halves so that every LDG/STG covers whole 32-byte
// sectors: instruction j of the pair handles row 2i (j < 2) or 2i+1 (j >= 2), the even lane the even 16-byte quad.
__device__ __forceinline__ uint4 tc_shfl_xor1(const uint4 v) {
  uint4 r;
  r.x = __shfl_xor_sync(0xffffffffu, v.x, 1); r.y = __shfl_xor_sync(0xffffffffu, v.y, 1);
  r.z = __shfl_xor_sync(0xffffffffu, v.z, 1); r.w = __shfl_xor_sync(0xffffffffu, v.w, 1);
  return r;
}
__device__ __forceinline__ uint4 tc_sel(bool c, const uint4 a, const uint4 b) { return c ? a : b; }

// rowA/rowB: element offsets (row * Cout + first column of the chunk) of the pair's even/odd row, valid only when
// okA/okB.  All 32 lanes must call this (shuffles).
// the pair's residual quads as they sit in memory (issued early so that their latency hides behind the MMA wait)
__device__ __forceinline__ void tc_res_load(const __half* __restrict__ residual, size_t rowA, size_t rowB, bool okA, bool okB, bool odd,
                                            uint4 (&l)[4]) {
  const int qlo = odd ? 1 : 0, qhi = odd ? 3 : 2;
  const uint4 z = make_uint4(0, 0, 0, 0);
  l[0] = okA ? *reinterpret_cast<const uint4*>(residual + rowA + 8 * qlo) : z;
  l[1] = okA ? *reinterpret_cast<const uint4*>(residual + rowA + 8 * qhi) : z;
  l[2] = okB ? *reinterpret_cast<const uint4*>(residual + rowB + 8 * qlo) : z;
  l[3] = okB ? *reinterpret_cast<const uint4*>(residual + rowB + 8 * qhi) : z;
}

__device__ __forceinline__ void tc_epilogue_chunk(const uint32_t (&v)[32], const float* __restrict__ sb, const float* __restrict__ sa,
                                                  const float* __restrict__ si, bool has_res, const uint4 (&l)[4],
                                                  __half* __restrict__ out_raw, __half* __restrict__ out_act, size_t rowA, size_t rowB,
                                                  bool okA, bool okB, bool odd) {
  float f[32];
#pragma unroll
  for (int j4 = 0; j4 < 8; ++j4) {
    const float4 b = reinterpret_cast<const float4*>(sb)[j4];
    f[4 * j4 + 0] = __uint_as_float(v[4 * j4 + 0]) + b.x; f[4 * j4 + 1] = __uint_as_float(v[4 * j4 + 1]) + b.y;
    f[4 * j4 + 2] = __uint_as_float(v[4 * j4 + 2]) + b.z; f[4 * j4 + 3] = __uint_as_float(v[4 * j4 + 3]) + b.w;
  }
  const int qlo = odd ? 1 : 0, qhi = odd ? 3 : 2;             // the 16-byte quads this lane touches in memory
  if (has_res) {
    const uint4 l0 = l[0], l1 = l[1], l2 = l[2], l3 = l[3];
    const uint4 r0 = tc_shfl_xor1(odd ? l0 : l2), r1 = tc_shfl_xor1(odd ? l1 : l3);
    const uint4 q[4] = {tc_sel(odd, r0, l0), tc_sel(odd, l2, r0), tc_sel(odd, r1, l1), tc_sel(odd, l3, r1)};
#pragma unroll
    for (int qq = 0; qq < 4; ++qq) {
      const __half2* h = reinterpret_cast<const __half2*>(&q[qq]);
#pragma unroll
      for (int j = 0; j < 4; ++j) { const float2 t = __half22float2(h[j]); f[qq * 8 + 2 * j] += t.x; f[qq * 8 + 2 * j + 1] += t.y; }
    }
  }
  auto store_rows = [&](__half* __restrict__ out, const uint4 (&q)[4]) {
    const uint4 r0 = tc_shfl_xor1(odd ? q[0] : q[1]), r1 = tc_shfl_xor1(odd ? q[2] : q[3]);
    if (okA) {
      *reinterpret_cast<uint4*>(out + rowA + 8 * qlo) = tc_sel(odd, r0, q[0]);
      *reinterpret_cast<uint4*>(out + rowA + 8 * qhi) = tc_sel(odd, r1, q[2]);
    }
    if (okB) {
      *reinterpret_cast<uint4*>(out + rowB + 8 * qlo) = tc_sel(odd, q[1], r0);
      *reinterpret_cast<uint4*>(out + rowB + 8 * qhi) = tc_sel(odd, q[3], r1);
    }
  };
  if (out_raw) {
    uint4 q[4];
#pragma unroll
    for (int qq = 0; qq < 4; ++qq) {
      __half2* h = reinterpret_cast<__half2*>(&q[qq]);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2half2_rn(f[qq * 8 + 2 * j], f[qq * 8 + 2 * j + 1]);
    }
    store_rows(out_raw, q);
  }
  if (out_act) {
    uint4 q[4];
#pragma unroll
    for (int qq = 0; qq < 4; ++qq) {
      __half2* h = reinterpret_cast<__half2*>(&q[qq]);
      const float4 a0 = reinterpret_cast<const float4*>(sa)[2 * qq], a1 = reinterpret_cast<const float4*>(sa)[2 * qq + 1];
      const float4 i0 = reinterpret_cast<const float4*>(si)[2 * qq], i1 = reinterpret_cast<const float4*>(si)[2 * qq + 1];
      h[0] = __floats2half2_rn(snake_fast(f[qq * 8 + 0], a0.x, i0.x), snake_fast(f[qq * 8 + 1], a0.y, i0.y));
      h[1] = __floats2half2_rn(snake_fast(f[qq * 8 + 2], a0.z, i0.z), snake_fast(f[qq * 8 + 3], a0.w, i0.w));
      h[2] = __floats2half2_rn(snake_fast(f[qq * 8 + 4], a1.x, i1.x), snake_fast(f[qq * 8 + 5], a1.y, i1.y));
      h[3] = __floats2half2_rn(snake_fast(f[qq * 8 + 6], a1.z, i1.z), snake_fast(f[qq * 8 + 7], a1.w, i1.w));
    }
    store_rows(out_act, q);
  }
}

// grid = (ceil(m_rows/128), Cout/n_tile, phases); block = 320; dyn smem = 1024 (align) + stages*(16K + n_tile*128) + 256
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
  // per-channel epilogue constants: one shared-memory broadcast read per element instead of two global loads and a
  // reciprocal per element and thread (the epilogue, not the MMA, bounds the small-channel layers)
  float* s_bias = reinterpret_cast<float*>(tc_smem_raw + (((tmem_slot + 16u + 15u) & ~15u) - raw));
  float* s_alpha = s_bias + 256;
  float* s_inv = s_alpha + 256;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(tc_smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ub = blockIdx.x / p.m_tiles;                        // utterance: the activation map is 3-D {C, T, nb}
  const int m0 = (blockIdx.x - ub * p.m_tiles) * TC_BM;
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
        tma_load_3d(a_dst, &map_x, kc * TC_BK, m0 + row_off, ub, full_bar(s));
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
    for (int i = threadIdx.x - 64; i < p.n_tile; i += 256) {
      s_bias[i] = __ldg(p.bias + n0 + i);
      const float a = p.out_act ? __ldg(p.alpha_next + n0 + i) : 1.f;
      s_alpha[i] = a;
      s_inv[i] = 1.0f / (a + 1e-9f);
    }
    asm volatile("bar.sync 2, 256;" ::: "memory");          // the 8 epilogue warps
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const bool odd = lane & 1;
    const int vA = __shfl_sync(0xffffffffu, (int)valid, lane & ~1), vB = __shfl_sync(0xffffffffu, (int)valid, lane | 1);
    const int orA = __shfl_sync(0xffffffffu, out_row, lane & ~1), orB = __shfl_sync(0xffffffffu, out_row, lane | 1);
    const size_t ubase = (size_t)ub * p.T_out;
    const size_t rowA = (ubase + (size_t)(vA ? orA : 0)) * p.Cout + n0, rowB = (ubase + (size_t)(vB ? orB : 0)) * p.Cout + n0;
    const int chalf = (warp - 2) >> 2;                      // the two warps of a lane quarter alternate 32-column chunks
    for (int c0 = chalf * 32; c0 < p.n_tile; c0 += 64) {
      uint32_t v[32];
      tmem_ld32(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, v);
      uint4 res[4];
      if (p.residual) tc_res_load(p.residual, rowA + c0, rowB + c0, vA != 0, vB != 0, odd, res);
      tc_epilogue_chunk(v, s_bias + c0, s_alpha + c0, s_inv + c0, p.residual != nullptr, res, p.out_raw, p.out_act, rowA + c0, rowB + c0,
                        vA != 0, vB != 0, odd);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)p.tmem_cols) : "memory");
  }
}

// Final conv (Cin -> 1 channel, k taps) + tanh: pure bandwidth, SIMT.  x fp16 [T][C] (already snaked), w fp32 [K][C].
// One block = FC_TILE output samples: the [FC_TILE + K - 1][C] input window is staged in shared memory with coalesced
// 16 B loads (the first version used one warp per output sample and ran at 0.15 TB/s, 13 % of the whole decode), rows
// padded to an odd number of 32-bit words so that 32 consecutive samples read 32 different banks.
#define FC_TILE 256
__global__ void __launch_bounds__(FC_TILE) k_dac_final_conv(const __half* __restrict__ x, const float* __restrict__ w, float bias,
                                                            float* __restrict__ out, int T, int C, int K, int pad) {
  extern __shared__ __align__(16) unsigned char fc_smem[];
  const int ldw = (C / 2) | 1;                                   // row stride in 32-bit words (odd)
  uint32_t* xs = reinterpret_cast<uint32_t*>(fc_smem);           // [FC_TILE + K - 1][ldw]
  float* ws = reinterpret_cast<float*>(xs + (size_t)(FC_TILE + K - 1) * ldw);   // [K][C]
  const int t0 = blockIdx.x * FC_TILE;
  x += (size_t)blockIdx.y * T * C;                               // grid.y = utterance
  out += (size_t)blockIdx.y * T;
  const int rows = FC_TILE + K - 1;
  const int wpr = C / 2;                                         // words per row
  for (int i = threadIdx.x; i < rows * wpr; i += FC_TILE) {
    const int r = i / wpr, c2 = i % wpr;
    const int tt = t0 + r - pad;
    uint32_t v = 0u;
    if (tt >= 0 && tt < T) v = *reinterpret_cast<const uint32_t*>(x + (size_t)tt * C + 2 * c2);
    xs[(size_t)r * ldw + c2] = v;
  }
  for (int i = threadIdx.x; i < K * C; i += FC_TILE) ws[i] = w[i];
  __syncthreads();
  const int t = t0 + threadIdx.x;
  if (t >= T) return;
  float acc0 = 0.f, acc1 = 0.f;
  for (int j = 0; j < K; ++j) {
    const uint32_t* xr = xs + (size_t)(threadIdx.x + j) * ldw;
    const float* wr = ws + j * C;
#pragma unroll 4
    for (int c2 = 0; c2 < wpr; ++c2) {
      const float2 v = __half22float2(*reinterpret_cast<const __half2*>(xr + c2));
      acc0 = fmaf(v.x, wr[2 * c2], acc0);
      acc1 = fmaf(v.y, wr[2 * c2 + 1], acc1);
    }
  }
  out[t] = tanhf(acc0 + acc1 + bias);
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
