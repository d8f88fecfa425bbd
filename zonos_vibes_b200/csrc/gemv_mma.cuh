// Weight-streaming skinny GEMM on the legacy tensor path (mma.sync m16n8k16 bf16 -> fp32) for decode with up to 8
// activation rows (B <= 4 utterances with CFG).  Why not tcgen05 here: with M = 2..8 rows the contraction is a GEMV —
// HBM-bound by ~100x — and tcgen05's minimum tile (M=64/128 from shared memory) would only add a smem round trip for
// the 3.2 GB weight stream.  mma.sync takes the weights straight from registers (the ld.global destination), does the
// bf16->fp32 conversion and the FMAs in one instruction, and leaves the SM's issue slots free: the SIMT version of
// this kernel (k_gemv, kept as the cross-check) spent 9.3 lane-instructions per weight element and reached only 36 %
// of DRAM peak (profiles/r01_fc1_simt.md); this one needs ~0.4.
//
// Mapping.  D[16 x 8] = A[16 x 16] * B[16 x 8]:  A = activations (rows 0..R-1 real, the rest zero), B = 8 weight rows
// (column n = weight row n), so one CTA task = 8 weight rows x the whole K, K split over the 8 warps of the CTA and
// reduced through shared memory.  The k index inside a 32-wide block is permuted so that every lane's B fragment is
// ONE 128-bit load of 8 consecutive bf16 of its weight row (8 rows x 64 B per warp instruction, sector-exact), and the
// A fragment is one LDS.128 at the same (row, k) pattern from the staged activations.
#pragma once
#include "decode_kernels.cuh"

#define GM_THREADS 256
#define GM_WARPS 8
#define GM_ROWS 8            // weight rows per task (MMA N)
#define GM_XPAD 32           // bf16 elements of padding per staged activation row (64 B: conflict-free LDS.128)

__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
  const uint32_t z = 0u;       // A rows 8..15 are zero (R <= 8)
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(z), "r"(a2), "r"(z), "r"(b0), "r"(b1));
}

// all four A registers (16 rows)
__device__ __forceinline__ void mma_bf16_16816_full(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// ---- single-row epilogues (same arithmetic as gemv_epilogue in decode_kernels.cuh) ---------------------------
template <int EPI>
__device__ __forceinline__ void gm_epilogue_pair(const GemvParams& p, int a /*activation row*/, int n0, float s0, float s1) {
  if constexpr (EPI == EPI_RESID) {
    __nv_bfloat162* px = reinterpret_cast<__nv_bfloat162*>(p.out + (size_t)a * p.out_stride + n0);
    const float2 old = __bfloat1622float2(*px);
    *px = __floats2bfloat162_rn(old.x + bf16_round(s0), old.y + bf16_round(s1));
  } else if constexpr (EPI == EPI_PLAIN) {
    *reinterpret_cast<__nv_bfloat162*>(p.out + (size_t)a * p.out_stride + n0) = __floats2bfloat162_rn(s0, s1);
  } else if constexpr (EPI == EPI_QKV) {
    const int pos = p.st->pos;
    const float v0 = bf16_round(s0), v1 = bf16_round(s1);
    if (n0 < p.q_size + p.kv_size) {
      const int i = (n0 % ZB_DH) >> 1;
      const float2 cs = *reinterpret_cast<const float2*>(p.rope + ((size_t)p.rope_pos[a] * (ZB_DH / 2) + i) * 2);
      const float o0 = __fsub_rn(__fmul_rn(v0, cs.x), __fmul_rn(v1, cs.y));
      const float o1 = __fadd_rn(__fmul_rn(v1, cs.x), __fmul_rn(v0, cs.y));
      const __nv_bfloat162 o = __floats2bfloat162_rn(o0, o1);
      if (n0 < p.q_size) {
        *reinterpret_cast<__nv_bfloat162*>(p.out + (size_t)a * p.out_stride + n0) = o;
      } else {
        const int m = n0 - p.q_size;
        const int page = p.page_table[a * p.max_pages + pos / ZB_PAGE];
        bf16* dst = p.kv_layer + ((((size_t)page * 2 + 0) * p.Hkv + m / ZB_DH) * ZB_PAGE + pos % ZB_PAGE) * ZB_DH + m % ZB_DH;
        *reinterpret_cast<__nv_bfloat162*>(dst) = o;
      }
    } else {
      const int m = n0 - p.q_size - p.kv_size;
      const int page = p.page_table[a * p.max_pages + pos / ZB_PAGE];
      bf16* dst = p.kv_layer + ((((size_t)page * 2 + 1) * p.Hkv + m / ZB_DH) * ZB_PAGE + pos % ZB_PAGE) * ZB_DH + m % ZB_DH;
      *reinterpret_cast<__nv_bfloat162*>(dst) = __floats2bfloat162_rn(v0, v1);
    }
  }
}

__device__ __forceinline__ void gm_epilogue_heads(const GemvParams& p, int b /*utterance*/, int n0, float c0, float c1, float u0, float u1) {
  const int k = n0 / p.V, n = n0 % p.V;
  const float scale = p.st->cfg_scale;
  const int eos_bias = p.st->apply_eos_bias;
  float out[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const float c = bf16_round(j == 0 ? c0 : c1);
    const float u = bf16_round(j == 0 ? u0 : u1);
    float l = __fadd_rn(u, __fmul_rn(__fsub_rn(c, u), scale));
    const int nn = n + j;
    if (nn >= 1025) l = -INFINITY;
    if (eos_bias && k >= 1 && nn == p.eos) l = -INFINITY;
    out[j] = l;
  }
  *reinterpret_cast<float2*>(p.logits + ((size_t)b * p.n_q + k) * p.V + n) = make_float2(out[0], out[1]);
}

// grid = 2 x #SMs, block = 256 (8 warps), dyn smem = R*(K+32)*2 + 2*NT*8*64*4 + 8*R*4
template <int R, int EPI>
__global__ void __launch_bounds__(GM_THREADS, 2) k_gemv_mma(const GemvParams p) {
  constexpr int NT = (EPI == EPI_SWIGLU) ? 2 : 1;      // weight tiles per task (value + gate rows for SwiGLU)
  constexpr int U = 8 / NT;                            // 32-wide k-blocks per pipeline unit (8 x 128-bit loads in flight)
  extern __shared__ __align__(16) unsigned char zb_smem[];
  const int K = p.K;
  const int ld = K + GM_XPAD;
  bf16* xs = reinterpret_cast<bf16*>(zb_smem);
  float* part = reinterpret_cast<float*>(zb_smem + (size_t)R * ld * sizeof(bf16));      // [2][NT][8 warps][8 r][8 n]
  float* red = part + 2 * NT * GM_WARPS * 64;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int kslice = K / GM_WARPS;                     // this warp's share of K (multiple of 32)
  const int nkb = kslice >> 5;                         // k-blocks per task for this warp
  const int upt = (nkb + U - 1) / U;                   // pipeline units per task
  const int my_tasks = (int)blockIdx.x < p.n_tasks ? (p.n_tasks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int nunits = my_tasks * upt;
  const int kbase = warp * kslice + t * 8;

  pdl_launch_dependents();

  uint4 wa[NT][U], wb[NT][U];
  auto load_unit = [&](uint4(&d)[NT][U], int u) {
    const int task = blockIdx.x + (u / upt) * gridDim.x;
    const int kb0 = (u % upt) * U;
    int row = task * GM_ROWS + g;
    if (row >= p.n_rows) row = p.n_rows - 1;           // ragged last task (9 x 1026 head rows): clamp, never stored
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const bf16* wp = p.W + (size_t)(row + nt * p.row1_off) * K + kbase;
#pragma unroll
      for (int j = 0; j < U; ++j)
        if (kb0 + j < nkb) d[nt][j] = ldg_stream(wp + (kb0 + j) * 32);
    }
  };

  if (nunits > 0) load_unit(wa, 0);                    // weights do not depend on the previous kernel
  if (p.pf_bytes && tid < 2) {
    const char* base = tid == 0 ? p.pf0 : p.pf1;
    if (base) {
      const unsigned per = p.pf_bytes / gridDim.x;
      const char* q = base + (size_t)blockIdx.x * per;
      for (unsigned off = 0; off < per; off += 32768u) l2_prefetch_bulk(q + off, min(32768u, per - off));
    }
  }
  pdl_wait_prior_grid();
  if (p.st->done) return;

  gemv_stage_activations<R>(p, xs, red, ld);
  __syncthreads();

  // four independent accumulator sets per weight tile: HMMA chains through its accumulator, and a single chain of
  // dependent HMMAs leaves the tensor pipe idle between issues
  float acc[NT][4][4];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int a = 0; a < 4; ++a) { acc[nt][a][0] = acc[nt][a][1] = acc[nt][a][2] = acc[nt][a][3] = 0.f; }
  const bf16* xrow = xs + (size_t)g * ld + kbase;      // only dereferenced when g < R

  auto compute_unit = [&](const uint4(&d)[NT][U], int u) {
    const int kb0 = (u % upt) * U;
#pragma unroll
    for (int j = 0; j < U; ++j) {
      if (kb0 + j < nkb) {
        uint4 xv = make_uint4(0, 0, 0, 0);
        if (g < R) xv = *reinterpret_cast<const uint4*>(xrow + (kb0 + j) * 32);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          mma_bf16_16816(acc[nt][2 * (j & 1)], xv.x, xv.y, d[nt][j].x, d[nt][j].y);
          mma_bf16_16816(acc[nt][2 * (j & 1) + 1], xv.z, xv.w, d[nt][j].z, d[nt][j].w);
        }
      }
    }
    if ((u % upt) == upt - 1) {
      // ---- task finished for this warp: publish the K-slice partial, reduce over the 8 warps, run the epilogue
      const int ti = u / upt;
      const int task = blockIdx.x + ti * gridDim.x;
      float* pb = part + (size_t)(ti & 1) * NT * GM_WARPS * 64;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        // c0,c1 = D[row g][n = 2t, 2t+1]
        *reinterpret_cast<float2*>(pb + ((nt * GM_WARPS + warp) * 8 + g) * 8 + 2 * t) =
            make_float2((acc[nt][0][0] + acc[nt][1][0]) + (acc[nt][2][0] + acc[nt][3][0]),
                        (acc[nt][0][1] + acc[nt][1][1]) + (acc[nt][2][1] + acc[nt][3][1]));
#pragma unroll
        for (int a = 0; a < 4; ++a) { acc[nt][a][0] = acc[nt][a][1] = acc[nt][a][2] = acc[nt][a][3] = 0.f; }
      }
      __syncthreads();
      const int n_base = task * GM_ROWS;
      if constexpr (EPI == EPI_SWIGLU) {
        if (tid < R * 8) {
          const int r = tid >> 3, n = tid & 7;
          float y = 0.f, gt = 0.f;
#pragma unroll
          for (int w = 0; w < GM_WARPS; ++w) { y += pb[((0 * GM_WARPS + w) * 8 + r) * 8 + n]; gt += pb[((1 * GM_WARPS + w) * 8 + r) * 8 + n]; }
          bool ok;
          int a = r < R / 2 ? p.base0 + r : p.base1 + (r - R / 2);
          ok = a < (r < R / 2 ? p.lim0 : p.lim1);
          if (ok && n_base + n < p.n_rows) {
            y = bf16_round(y);
            gt = bf16_round(gt);
            const float sg = bf16_round(gt / (1.0f + expf(-gt)));
            p.out[(size_t)a * p.out_stride + n_base + n] = __float2bfloat16_rn(y * sg);
          }
        }
      } else if constexpr (EPI == EPI_HEADS) {
        if (tid < (R / 2) * 4) {
          const int b = tid >> 2, j = tid & 3;
          float c0 = 0.f, c1 = 0.f, u0 = 0.f, u1 = 0.f;
#pragma unroll
          for (int w = 0; w < GM_WARPS; ++w) {
            c0 += pb[(w * 8 + b) * 8 + 2 * j]; c1 += pb[(w * 8 + b) * 8 + 2 * j + 1];
            u0 += pb[(w * 8 + b + R / 2) * 8 + 2 * j]; u1 += pb[(w * 8 + b + R / 2) * 8 + 2 * j + 1];
          }
          if (p.base0 + b < p.lim0 && n_base + 2 * j < p.n_rows) gm_epilogue_heads(p, p.base0 + b, n_base + 2 * j, c0, c1, u0, u1);
        }
      } else {
        if (tid < R * 4) {
          const int r = tid >> 2, j = tid & 3;
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int w = 0; w < GM_WARPS; ++w) { s0 += pb[(w * 8 + r) * 8 + 2 * j]; s1 += pb[(w * 8 + r) * 8 + 2 * j + 1]; }
          const int a = r < R / 2 ? p.base0 + r : p.base1 + (r - R / 2);
          const bool ok = a < (r < R / 2 ? p.lim0 : p.lim1);
          if (ok && n_base + 2 * j < p.n_rows) gm_epilogue_pair<EPI>(p, a, n_base + 2 * j, s0, s1);
        }
      }
      // no second barrier: the next task writes the other half of `part`; this half is rewritten only after the
      // NEXT task's __syncthreads, which every epilogue thread reaches after finishing its reads.
    }
  };

  for (int u = 0; u < nunits; u += 2) {
    if (u + 1 < nunits) load_unit(wb, u + 1);
    compute_unit(wa, u);
    if (u + 1 < nunits) {
      if (u + 2 < nunits) load_unit(wa, u + 2);
      compute_unit(wb, u + 1);
    }
  }
}
