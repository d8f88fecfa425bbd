// Batched decode step (more than 8 activation rows, i.e. B > 4 utterances with CFG): once the batch makes the
// projections dense contractions they move to the 5th-gen tensor cores.  Reference arithmetic: the same
// zonos/backbone/_torch.py:83-152 and zonos/model.py:100-116 as the batch-1 kernels; batch semantics = B independent
// batch-1 runs.
//
//   D[m, n] = sum_k X[m, k] * W[n, k]        X = activations [rows][K] bf16 (L2 resident), W = weights [N][K] bf16 (HBM)
//
// k_bt_gemm: one CTA = one 128-row activation tile x one 128-row weight tile x one K split.  Warp 0 = TMA producer
// (A box {64 k, 128 rows} of the activations, B box {64 k, 128 rows} of the weights, both SWIZZLE_128B K-major),
// warp 1 = TMEM allocator + single-thread tcgen05.mma issuer (kind::f16, bf16 operands, fp32 accumulators in TMEM),
// warps 2..5 = epilogue (tcgen05.ld 32x32b).  The weights are streamed ONCE per step for all rows (the per-op GEMV
// path would re-stream them per group of 8 rows), so the step stays HBM-bound: 3.2 GB of weights + the KV cache.
// The grid is filled by split-K (out_proj / fc2 have only 16 weight tiles): every split stores its fp32 partial tile to
// a workspace and a small fused "finish" kernel sums the splits in a FIXED order (reproducible tokens) and applies the
// epilogue that needs whole rows anyway (residual + next LayerNorm, RoPE + KV append, CFG mix).  fc1 needs no split
// (128 tiles) and runs its SwiGLU epilogue straight out of TMEM: value rows and gate rows are two B boxes of one tile.
#pragma once
#include <cuda.h>

#include "decode_kernels.cuh"
#include "tc_common.cuh"

#define BT_THREADS 192
#define BT_BM 128
#define BT_BN 128
#define BT_BK 64
#define BT_A_BYTES (BT_BM * BT_BK * 2)      // 16 KB
#define BT_B_BYTES (BT_BN * BT_BK * 2)      // 16 KB
#define BT_STAGE_BYTES (BT_A_BYTES + BT_B_BYTES)
#define BT_STAGES 6
#define BT_SMEM_BYTES (1024 + BT_STAGES * BT_STAGE_BYTES + 256)

enum { BT_MODE_PARTIAL = 0, BT_MODE_SWIGLU = 1 };

struct BtGemmParams {
  int cs;                // cluster size along the weight-tile dimension (1, 2 or 4): the CTAs of a cluster share the activation
                         // tile, every CTA loads 1 / cs of it and multicasts it to all of them (the launch carries the cluster attribute)
  int M;                 // real activation rows
  int N;                 // weight rows (PARTIAL) or ffn size F (SWIGLU: value rows [0,F), gate rows [F,2F))
  int num_kb;            // K / 64
  int kb_per_split;      // k-blocks per split (grid.z splits)
  int mode;
  float* partial;        // PARTIAL: [splits][m_pad][N] fp32
  int m_pad;
  bf16* h;               // SWIGLU: [rows][F] bf16
  const ZbDevState* st;  // done flag
  size_t partial_elems;  // floats allocated behind `partial` (checked build)
};

__device__ __forceinline__ void bt_tma_load_2d_mc(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar, uint16_t mask) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%2, %3}], [%4], %5;" ::"r"(dst),
               "l"(map), "r"(c0), "r"(c1), "r"(bar), "h"(mask)
               : "memory");
}
__device__ __forceinline__ void bt_commit_mc(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ void bt_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// grid = (n_tiles, m_tiles, splits); block = 192; dyn smem = BT_SMEM_BYTES.  map_xs: the activations again with a box of
// BT_BM / cs rows (the slice one CTA of a cluster loads and multicasts); unused when cs == 1.
__global__ void __launch_bounds__(BT_THREADS, 1) k_bt_gemm(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_xs,
                                                           const __grid_constant__ CUtensorMap map_w, const BtGemmParams p) {
  extern __shared__ unsigned char bt_smem_raw[];
  const uint32_t raw = smem_u32(bt_smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bar_base = base + BT_STAGES * BT_STAGE_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * (uint32_t)s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (uint32_t)(BT_STAGES + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (uint32_t)(2 * BT_STAGES);
  const uint32_t tmem_slot = tmem_full_bar + 8u;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(bt_smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x, m0 = blockIdx.y * BT_BM, split = blockIdx.z;
  const int kb0 = split * p.kb_per_split;
  const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
  const int nkb = kb1 - kb0;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
    // a stage is free again when the MMAs of ALL CTAs of the cluster have read it (every CTA multicasts into every CTA)
    for (int s = 0; s < BT_STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), (uint32_t)p.cs); }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const int cs = p.cs;
  uint32_t crank = 0;
  if (cs > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  const uint16_t cmask = (uint16_t)((1u << cs) - 1u);
  const int a_rows = BT_BM / cs;                                  // rows of the activation tile this CTA loads
  const uint32_t a_off = crank * (uint32_t)(a_rows * BT_BK * 2);  // ... and where they land in every CTA's stage
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (cs > 1) bt_cluster_sync();                                  // every CTA's barriers exist before a peer's multicast signals them
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot_ptr;

  pdl_launch_dependents();
  const bool swiglu = p.mode == BT_MODE_SWIGLU;
  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer.  The weight tiles do not depend on the previous kernel: they are requested BEFORE the
      // grid dependency resolves (programmatic dependent launch), the activation tiles after.
      int pre = nkb < BT_STAGES ? nkb : BT_STAGES;
      for (int i = 0; i < pre; ++i) {
        const uint32_t b_dst = base + (uint32_t)i * BT_STAGE_BYTES + BT_A_BYTES;
        mbar_expect_tx(full_bar(i), BT_STAGE_BYTES);
        if (swiglu) {
          tma_load_2d(b_dst, &map_w, (kb0 + i) * BT_BK, n_tile * 64, full_bar(i));
          tma_load_2d(b_dst + BT_B_BYTES / 2, &map_w, (kb0 + i) * BT_BK, p.N + n_tile * 64, full_bar(i));
        } else {
          tma_load_2d(b_dst, &map_w, (kb0 + i) * BT_BK, n_tile * BT_BN, full_bar(i));
        }
      }
      pdl_wait_prior_grid();
      for (int i = 0; i < pre; ++i) {
        if (cs > 1) bt_tma_load_2d_mc(base + (uint32_t)i * BT_STAGE_BYTES + a_off, &map_xs, (kb0 + i) * BT_BK, m0 + (int)crank * a_rows, full_bar(i), cmask);
        else tma_load_2d(base + (uint32_t)i * BT_STAGE_BYTES, &map_x, (kb0 + i) * BT_BK, m0, full_bar(i));
      }
      for (int i = pre; i < nkb; ++i) {
        const int s = i % BT_STAGES;
        mbar_wait(empty_bar(s), ((i / BT_STAGES) - 1) & 1);
        const uint32_t a_dst = base + (uint32_t)s * BT_STAGE_BYTES;
        const uint32_t b_dst = a_dst + BT_A_BYTES;
        mbar_expect_tx(full_bar(s), BT_STAGE_BYTES);
        if (cs > 1) bt_tma_load_2d_mc(a_dst + a_off, &map_xs, (kb0 + i) * BT_BK, m0 + (int)crank * a_rows, full_bar(s), cmask);
        else tma_load_2d(a_dst, &map_x, (kb0 + i) * BT_BK, m0, full_bar(s));
        if (swiglu) {
          tma_load_2d(b_dst, &map_w, (kb0 + i) * BT_BK, n_tile * 64, full_bar(s));
          tma_load_2d(b_dst + BT_B_BYTES / 2, &map_w, (kb0 + i) * BT_BK, p.N + n_tile * 64, full_bar(s));
        } else {
          tma_load_2d(b_dst, &map_w, (kb0 + i) * BT_BK, n_tile * BT_BN, full_bar(s));
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_f16(BT_BN, 1u);        // bf16 x bf16 -> fp32
      for (int i = 0; i < nkb; ++i) {
        const int s = i % BT_STAGES;
        mbar_wait(full_bar(s), (i / BT_STAGES) & 1);
        tc_fence_after();
        const uint32_t a_addr = base + (uint32_t)s * BT_STAGE_BYTES;
        const uint64_t adesc = umma_desc_sw128(a_addr);
        const uint64_t bdesc = umma_desc_sw128(a_addr + BT_A_BYTES);
#pragma unroll
        for (int k = 0; k < BT_BK / 16; ++k)
          tc_mma_f16(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (i | k) != 0);
        if (cs > 1) bt_commit_mc(empty_bar(s), cmask);
        else tc_commit(empty_bar(s));
      }
      tc_commit(tmem_full_bar);
    }
    __syncwarp();
  } else {
    // ===== epilogue: warp w owns TMEM lanes [32*(w%4), +32) = activation rows
    const int lg = warp & 3;
    const int m = m0 + lg * 32 + lane;
    const bool valid = m < p.M;
    if (nkb > 0) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    if (swiglu) {
      // fc1 -> y * silu(gate), all three rounded to bf16 like the reference's bf16 tensors (_torch.py:150-152)
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t vy[32], vg[32];
        tmem_ld32(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, vy);
        tmem_ld32(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)(64 + c0), vg);
        if (valid) {
          bf16* dst = p.h + (size_t)m * p.N + n_tile * 64 + c0;
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            __nv_bfloat162 o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float r[2];
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                const float y = bf16_round(__uint_as_float(vy[q4 * 8 + 2 * j + e]));
                const float gt = bf16_round(__uint_as_float(vg[q4 * 8 + 2 * j + e]));
                const float sg = bf16_round(gt / (1.0f + expf(-gt)));
                r[e] = y * sg;
              }
              o[j] = __floats2bfloat162_rn(r[0], r[1]);
            }
            if (n_tile * 64 + c0 + q4 * 8 < p.N) *reinterpret_cast<uint4*>(dst + q4 * 8) = *reinterpret_cast<uint4*>(o);
          }
        }
      }
    } else {
      float* dst_row = p.partial + ((size_t)split * p.m_pad + m) * p.N + (size_t)n_tile * BT_BN;
      ZB_ASSERT(!valid || ((size_t)split * p.m_pad + m) * p.N + min(p.N, (n_tile + 1) * BT_BN) <= p.partial_elems);
      ZB_ASSERT(m < p.m_pad || !valid);
      for (int c0 = 0; c0 < BT_BN; c0 += 32) {
        uint32_t v[32];
        if (nkb > 0) {
          tmem_ld32(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, v);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        if (valid) {
          const int nb = n_tile * BT_BN + c0;
          if (nb + 32 <= p.N && (p.N & 3) == 0) {
#pragma unroll
            for (int q4 = 0; q4 < 8; ++q4) *reinterpret_cast<uint4*>(dst_row + c0 + q4 * 4) = make_uint4(v[q4 * 4], v[q4 * 4 + 1], v[q4 * 4 + 2], v[q4 * 4 + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) if (nb + j < p.N) dst_row[c0 + j] = __uint_as_float(v[j]);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (cs > 1) bt_cluster_sync();                                  // no CTA leaves while a peer may still multicast into it / arrive on its barriers
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(128u) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------
// k_bt_gemm2: the same contraction on CTA PAIRS (tcgen05.mma.cta_group::2, M = 256).  What bounds k_bt_gemm at batch 64 is
// the bytes delivered into each SM (profiles/r02_summary.md section 3): every CTA pulls its 128-row weight tile AND the whole
// 128-row activation tile.  Here the roles are swapped - the WEIGHTS are the M operand (128 rows per CTA, 256 per pair, each
// CTA loads only its own), the ACTIVATIONS the N operand (N = 128 rows, split between the pair: each CTA loads 64) - so a
// CTA ingests 16 + 8 KB per k-block instead of 16 + 16.  The accumulator comes out transposed (TMEM lane = weight row,
// column = activation row); the epilogues store it into the same [rows][N] layouts as k_bt_gemm.
//   grid = (weight tiles padded to even, activation tiles, splits), cluster (2, 1, 1); rank 0 of a pair issues the MMAs, both
//   CTAs' TMA loads complete on rank 0's full barrier, the commits are multicast to both CTAs' empty / accumulator barriers.
// SWIGLU: the CTA's 128 weight rows are, per TMEM lane quarter, 16 value rows then the 16 gate rows of the same ffn columns
// (one 4-D TMA box over fc1 viewed as [16-column group][value | gate][16 rows][K]): a warp pairs value and gate with one
// shuffle per two activation rows, every lane evaluates one SwiGLU, and the tile leaves through shared memory as 16-byte
// row segments.
// ---------------------------------------------------------------------------------------------------------------
#define BT2_A_BYTES (128 * BT_BK * 2)       // 16 KB: this CTA's weight rows
#define BT2_B_BYTES (64 * BT_BK * 2)        // 8 KB: this CTA's half of the activation rows
#define BT2_STAGE_BYTES (BT2_A_BYTES + BT2_B_BYTES)
#define BT2_STAGES 8
#define BT2_SMEM_BYTES (1024 + BT2_STAGES * BT2_STAGE_BYTES + 256)

__device__ __forceinline__ void bt2_tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t leader_bar) {
  asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
               "l"(map), "r"(c0), "r"(c1), "r"(leader_bar)
               : "memory");
}
__device__ __forceinline__ void bt2_tma_load_4d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, int c3, uint32_t leader_bar) {
  asm volatile("cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(dst),
               "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(leader_bar)
               : "memory");
}
// Execution barrier of the pair without the release fence (barrier.cluster.arrive.release is a MEMBAR.ALL.GPU: at the end of
// the kernel it would wait for every epilogue store to become visible device-wide).  Barrier initialisation is published by
// fence.mbarrier_init.release.cluster; nothing else is communicated through memory across this barrier.
__device__ __forceinline__ void bt2_pair_sync() {
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void bt2_commit(uint32_t bar) {         // arrives on `bar` of both CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void bt2_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// map_xh: the activations with a box of {64 k, 64 rows}; map_w: the weights with a box of {64 k, 128 rows}, or (SWIGLU) fc1 as
// the 4-D tensor {K, 16 rows, value | gate, F / 16 groups} with a box of {64 k, 16, 2, 4}
__global__ void __launch_bounds__(BT_THREADS, 1) k_bt_gemm2(const __grid_constant__ CUtensorMap map_xh, const __grid_constant__ CUtensorMap map_w,
                                                            const BtGemmParams p) {
  extern __shared__ unsigned char bt_smem_raw[];
  const uint32_t raw = smem_u32(bt_smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bar_base = base + BT2_STAGES * BT2_STAGE_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * (uint32_t)s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (uint32_t)(BT2_STAGES + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (uint32_t)(2 * BT2_STAGES);
  const uint32_t tmem_slot = tmem_full_bar + 8u;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(bt_smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x, m0 = blockIdx.y * BT_BM, split = blockIdx.z;
  const int kb0 = split * p.kb_per_split;
  const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
  const int nkb = kb1 - kb0;
  uint32_t crank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  const bool leader = crank == 0;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_xh) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
    for (int s = 0; s < BT2_STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  bt2_pair_sync();                                              // the peer's barriers and accumulator exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot_ptr;

  pdl_launch_dependents();
  const bool swiglu = p.mode == BT_MODE_SWIGLU;
  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer (both CTAs).  Every load of the pair completes on the LEADER's full barrier: the address of the
      // same barrier in the even CTA of the pair is this CTA's address with the peer bit cleared.
      auto lfull = [&](int s) { return full_bar(s) & 0xFEFFFFFFu; };
      auto load_w = [&](int i, int s) {
        const uint32_t a_dst = base + (uint32_t)s * BT2_STAGE_BYTES;
        if (swiglu) {
          bt2_tma_load_4d(a_dst, &map_w, (kb0 + i) * BT_BK, 0, 0, n_tile * 4, lfull(s));
        } else {
          bt2_tma_load_2d(a_dst, &map_w, (kb0 + i) * BT_BK, n_tile * 128, lfull(s));
        }
      };
      auto load_x = [&](int i, int s) {
        bt2_tma_load_2d(base + (uint32_t)s * BT2_STAGE_BYTES + BT2_A_BYTES, &map_xh, (kb0 + i) * BT_BK, m0 + (int)crank * 64, lfull(s));
      };
      const int pre = nkb < BT2_STAGES ? nkb : BT2_STAGES;
      for (int i = 0; i < pre; ++i) {                              // weights do not depend on the previous kernel
        if (leader) mbar_expect_tx(full_bar(i), 2 * BT2_STAGE_BYTES);
        load_w(i, i);
      }
      pdl_wait_prior_grid();
      for (int i = 0; i < pre; ++i) load_x(i, i);
      for (int i = pre; i < nkb; ++i) {
        const int s = i % BT2_STAGES;
        mbar_wait(empty_bar(s), ((i / BT2_STAGES) - 1) & 1);
        if (leader) mbar_expect_tx(full_bar(s), 2 * BT2_STAGE_BYTES);
        load_x(i, s);
        load_w(i, s);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0 && leader) {
      uint32_t idesc = 0;                                         // kind::f16: D = F32, A = B = BF16, K-major, M = 256 (pair), N = 128
      idesc |= 1u << 4; idesc |= 1u << 7; idesc |= 1u << 10; idesc |= (uint32_t)(128 >> 3) << 17; idesc |= (uint32_t)(256 >> 4) << 24;
      for (int i = 0; i < nkb; ++i) {
        const int s = i % BT2_STAGES;
        mbar_wait(full_bar(s), (i / BT2_STAGES) & 1);
        tc_fence_after();
        const uint32_t a_addr = base + (uint32_t)s * BT2_STAGE_BYTES;
        const uint64_t adesc = umma_desc_sw128(a_addr);
        const uint64_t bdesc = umma_desc_sw128(a_addr + BT2_A_BYTES);
#pragma unroll
        for (int k = 0; k < BT_BK / 16; ++k)
          bt2_mma_f16(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (i | k) != 0);
        bt2_commit(empty_bar(s));
      }
      bt2_commit(tmem_full_bar);
    }
    __syncwarp();
  } else {
    // ===== epilogue: warp w owns TMEM lanes [32*(w%4), +32) = weight rows of this CTA; columns = activation rows
    const int lg = warp & 3;
    const int mcols = min(BT_BM, (p.M - m0 + 31) & ~31);          // accumulator columns that are real activation rows (in 32s)
    if (nkb > 0) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    if (swiglu) {
      // fc1 -> y * silu(gate), all three rounded to bf16 like the reference's bf16 tensors (_torch.py:150-152)
      // lanes 0..15 of the quarter hold value rows, lanes 16..31 the gate rows of the same 16 ffn columns.  For activation rows
      // (e, e + 1): the value lane keeps y[e] and receives g[e], the gate lane keeps g[e + 1] and receives y[e + 1].
      const int BT2_HROW = 160;                                                  // bytes per staged row: 64 bf16 + pad (bank spread, 16 B aligned)
      unsigned char* s_h = bt_smem_raw + (base - raw);                           // [128 activation rows][BT2_HROW]; the stages are drained
      const bool gate_lane = lane >= 16;
      const int jc = lg * 16 + (lane & 15);                                      // ffn column of the tile
      for (int c0 = 0; c0 < mcols; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          const uint32_t got = __shfl_xor_sync(0xffffffffu, gate_lane ? v[e] : v[e + 1], 16);
          const float y = bf16_round(__uint_as_float(gate_lane ? got : v[e]));
          const float gt = bf16_round(__uint_as_float(gate_lane ? v[e + 1] : got));
          const float sg = bf16_round(gt / (1.0f + expf(-gt)));
          ZB_ASSERT((size_t)(c0 + e + 1) * BT2_HROW + jc * 2 + 2 <= (size_t)BT2_STAGE_BYTES);
          *reinterpret_cast<bf16*>(s_h + (size_t)(c0 + e + (gate_lane ? 1 : 0)) * BT2_HROW + jc * 2) = __float2bfloat16_rn(y * sg);
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      {
        const int t = threadIdx.x - 64;                                          // 0..127 over the four epilogue warps
        const int seg = t & 7;                                                   // 16-byte segment of a 128-byte row
        if (n_tile * 64 + seg * 8 < p.N) {
#pragma unroll 4
          for (int r = t >> 3; r < mcols; r += 16) {
            const int m = m0 + r;
            if (m < p.M) *reinterpret_cast<uint4*>(p.h + (size_t)m * p.N + n_tile * 64 + seg * 8) = *reinterpret_cast<const uint4*>(s_h + (size_t)r * BT2_HROW + seg * 16);
          }
        }
      }
    } else {
      const int n = n_tile * 128 + lg * 32 + lane;
      for (int c0 = 0; c0 < mcols; c0 += 32) {
        uint32_t v[32];
        if (nkb > 0) {
          tmem_ld32(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, v);
        } else {
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = 0u;
        }
        float* dst = p.partial + ((size_t)split * p.m_pad + m0 + c0) * p.N + n;
        ZB_ASSERT(n >= p.N || m0 + c0 >= p.M || ((size_t)split * p.m_pad + min(p.M, m0 + c0 + 32) - 1) * p.N + n < p.partial_elems);
        ZB_ASSERT(p.M <= p.m_pad);
#pragma unroll
        for (int e = 0; e < 32; ++e)
          if (m0 + c0 + e < p.M && n < p.N) dst[(size_t)e * p.N] = __uint_as_float(v[e]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  bt2_pair_sync();                                              // the leader's MMAs read the peer's shared memory and write its accumulator
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(128u) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------
// finish kernels: fixed-order split sums + the epilogues that need whole rows
// ---------------------------------------------------------------------------------------------------------------
struct BtFinishParams {
  const float* partial;      // [splits][m_pad][N] fp32 (nullptr: no GEMM result to add, LayerNorm only)
  int splits, m_pad, N;
  bf16* x;                   // residual stream [rows][d]
  bf16* xn;                  // LayerNorm output [rows][d] (A operand of the next GEMM); nullptr = none
  const bf16 *ln_w, *ln_b;
  float eps;
  int d;
  // qkv
  bf16* q;                   // [rows][q_size]
  bf16* kv_layer;
  const int* page_table;
  int max_pages, Hkv, q_size, kv_size;
  const float* rope;
  const int* rope_pos;
  // heads
  float* logits;             // [B][n_q][V]
  int n_q, V, eos, B;
  const ZbDevState* st;
  int vrows;                 // > 0 (block prefill): activation row a = token a / vrows of cache row a % vrows, position pos + a / vrows
  int total_pages;           // pages of one layer's cache (checked build)
};

__device__ __forceinline__ float bt_split_sum(const float* partial, int splits, size_t split_stride, size_t off) {
  float s = 0.f;
  for (int i = 0; i < splits; ++i) s += __ldcg(partial + (size_t)i * split_stride + off);
  return s;
}

// 8 consecutive columns: all split partials are requested before the first add (one L2 round trip), fixed summation order
__device__ __forceinline__ void bt_split_sum8(const float* partial, int splits, size_t split_stride, size_t off, float (&a)[8]) {
  float4 lo[8], hi[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (i < splits) {
      lo[i] = __ldcg(reinterpret_cast<const float4*>(partial + (size_t)i * split_stride + off));
      hi[i] = __ldcg(reinterpret_cast<const float4*>(partial + (size_t)i * split_stride + off + 4));
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) a[j] = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (i < splits) {
      a[0] += lo[i].x; a[1] += lo[i].y; a[2] += lo[i].z; a[3] += lo[i].w;
      a[4] += hi[i].x; a[5] += hi[i].y; a[6] += hi[i].z; a[7] += hi[i].w;
    }
  }
}

// x[row] += bf16(sum of the split partials)  (bf16 residual stream, _torch.py:100-101), then xn[row] = LayerNorm(x[row])
// with the NEXT norm's parameters (F.layer_norm: biased variance, eps inside the sqrt, fp32 statistics).
// grid = rows, block = 256; d <= 2048 * k handled by a strided loop with the row kept in shared memory.
__global__ void __launch_bounds__(256) k_bt_resid_ln(const BtFinishParams p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  extern __shared__ float bt_row[];          // [d]
  __shared__ float red[16];
  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const size_t split_stride = (size_t)p.m_pad * p.N;
  float sum = 0.f;
  for (int e = tid * 8; e < p.d; e += 2048) {
    uint4* px = reinterpret_cast<uint4*>(p.x + (size_t)row * p.d + e);
    const uint4 xr = *px;
    float v[8] = {bf16lo(xr.x), bf16hi(xr.x), bf16lo(xr.y), bf16hi(xr.y), bf16lo(xr.z), bf16hi(xr.z), bf16lo(xr.w), bf16hi(xr.w)};
    if (p.partial) {
      float a[8];
      bt_split_sum8(p.partial, p.splits, split_stride, (size_t)row * p.N + e, a);
      __nv_bfloat162 o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) o[j] = __floats2bfloat162_rn(v[2 * j] + bf16_round(a[2 * j]), v[2 * j + 1] + bf16_round(a[2 * j + 1]));
      *px = *reinterpret_cast<uint4*>(o);
#pragma unroll
      for (int j = 0; j < 4; ++j) { const float2 f = __bfloat1622float2(o[j]); v[2 * j] = f.x; v[2 * j + 1] = f.y; }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { bt_row[e + j] = v[j]; sum += v[j]; }
  }
  if (!p.xn) return;
  sum = warp_sum(sum);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  float tot = 0.f;
  for (int w = 0; w < 8; ++w) tot += red[w];
  const float mean = tot / (float)p.d;
  float var = 0.f;
  for (int e = tid * 2; e < p.d; e += 512) { const float a = bt_row[e] - mean, b = bt_row[e + 1] - mean; var += a * a + b * b; }
  var = warp_sum(var);
  if (lane == 0) red[8 + warp] = var;
  __syncthreads();
  float vt = 0.f;
  for (int w = 0; w < 8; ++w) vt += red[8 + w];
  const float rstd = rsqrtf(vt / (float)p.d + p.eps);
  for (int e = tid * 2; e < p.d; e += 512) {
    const float2 g = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p.ln_w + e));
    const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p.ln_b + e));
    *reinterpret_cast<__nv_bfloat162*>(p.xn + (size_t)row * p.d + e) =
        __floats2bfloat162_rn((bt_row[e] - mean) * rstd * g.x + b.x, (bt_row[e + 1] - mean) * rstd * g.y + b.y);
  }
}

// in_proj finish: bf16 round, interleaved-pair RoPE in fp32 on q and k (_torch.py:18-30), q -> q buffer, k/v -> paged cache
// at the current position (_torch.py:33-49).  grid = (rows, ceil(N / 2048)), block = 256.
__global__ void __launch_bounds__(256) k_bt_qkv_finish(const BtFinishParams p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;
  const int row = blockIdx.x;
  const int crow = p.vrows ? row % p.vrows : row, tok = p.vrows ? row / p.vrows : 0;
  const int pos = p.st->pos + tok;
  const size_t split_stride = (size_t)p.m_pad * p.N;
  ZB_ASSERT(pos >= 0 && pos / ZB_PAGE < p.max_pages && row < p.m_pad);
  const int page = p.page_table[(size_t)crow * p.max_pages + pos / ZB_PAGE];
  ZB_ASSERT(page >= 0 && page < p.total_pages);
  const float* cs_row = p.rope + (size_t)(p.rope_pos[crow] + tok) * (ZB_DH / 2) * 2;
  for (int n8 = blockIdx.y * 2048 + threadIdx.x * 8; n8 < p.N; n8 += 2048 * gridDim.y) {     // grid.y = ceil(N / 2048): one trip per thread
    float a[8];
    bt_split_sum8(p.partial, p.splits, split_stride, (size_t)row * p.N + n8, a);
    __nv_bfloat162 o[4];
    const bool roped = n8 < p.q_size + p.kv_size;                 // 8 | 128: a group of 8 columns never straddles q / k / v
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float v0 = bf16_round(a[2 * j]), v1 = bf16_round(a[2 * j + 1]);
      if (roped) {
        const int i = ((n8 + 2 * j) % ZB_DH) >> 1;
        const float2 cs = *reinterpret_cast<const float2*>(cs_row + i * 2);
        o[j] = __floats2bfloat162_rn(__fsub_rn(__fmul_rn(v0, cs.x), __fmul_rn(v1, cs.y)), __fadd_rn(__fmul_rn(v1, cs.x), __fmul_rn(v0, cs.y)));
      } else {
        o[j] = __floats2bfloat162_rn(v0, v1);
      }
    }
    bf16* dst;
    if (n8 < p.q_size) {
      dst = p.q + (size_t)row * p.q_size + n8;
    } else {
      const int isv = n8 >= p.q_size + p.kv_size;
      const int m = n8 - p.q_size - (isv ? p.kv_size : 0);
      dst = p.kv_layer + ((((size_t)page * 2 + isv) * p.Hkv + m / ZB_DH) * ZB_PAGE + pos % ZB_PAGE) * ZB_DH + m % ZB_DH;
    }
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<uint4*>(o);
  }
}

// plain finish (hybrid backbone, Mamba2 in_proj): out[row][n] = bf16(sum of the split partials).  grid = (rows, ceil(N / 2048)), block = 256.
__global__ void __launch_bounds__(256) k_bt_plain_finish(const BtFinishParams p, bf16* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;
  const int row = blockIdx.x;
  const size_t split_stride = (size_t)p.m_pad * p.N;
  for (int n8 = blockIdx.y * 2048 + threadIdx.x * 8; n8 < p.N; n8 += 2048 * gridDim.y) {   // N % 8 == 0 (checked by the host)
    float a[8];
    bt_split_sum8(p.partial, p.splits, split_stride, (size_t)row * p.N + n8, a);
    __nv_bfloat162 o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = __floats2bfloat162_rn(a[2 * j], a[2 * j + 1]);
    *reinterpret_cast<uint4*>(out + (size_t)row * p.N + n8) = *reinterpret_cast<uint4*>(o);
  }
}

// heads finish: bf16 round, fp32 CFG mix of rows b (cond) and B+b (uncond), pad / EOS masks (model.py:111-115,266-267,280).
// grid = (n_q, B), block = 256.
__global__ void __launch_bounds__(256) k_bt_heads_finish(const BtFinishParams p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;
  const int k = blockIdx.x, b = blockIdx.y;
  const float scale = p.st->cfg_scale;
  const int eos_bias = p.st->apply_eos_bias;
  const size_t split_stride = (size_t)p.m_pad * p.N;
  for (int n = threadIdx.x; n < p.V; n += 256) {
    const size_t col = (size_t)k * p.V + n;
    const float c = bf16_round(bt_split_sum(p.partial, p.splits, split_stride, (size_t)b * p.N + col));
    const float u = bf16_round(bt_split_sum(p.partial, p.splits, split_stride, (size_t)(b + p.B) * p.N + col));
    float l = __fadd_rn(u, __fmul_rn(__fsub_rn(c, u), scale));
    if (n >= 1025) l = -INFINITY;
    if (eos_bias && k >= 1 && n == p.eos) l = -INFINITY;
    p.logits[((size_t)b * p.n_q + k) * p.V + n] = l;
  }
}

// plugin-level forward: post-norm_f hidden state of token t -> out[row, t, :]   (backbone/_torch.py:80)
__global__ void k_bt_store_hidden(const bf16* __restrict__ xn, bf16* __restrict__ out, int d, int S, int t) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  const int row = blockIdx.x;
  for (int e = threadIdx.x * 8; e < d; e += blockDim.x * 8)
    *reinterpret_cast<uint4*>(out + ((size_t)row * S + t) * d + e) = *reinterpret_cast<const uint4*>(xn + (size_t)row * d + e);
}
