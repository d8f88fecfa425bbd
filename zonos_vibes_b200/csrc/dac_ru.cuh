// DAC 44 kHz decoder — one FUSED kernel per ResidualUnit for the narrow blocks (96 and 192 channels):
//
//     out = x + conv_k1( snake2( conv_k7_dilated( snake1(x) ) + b1 ) ) + b2            (modeling_dac.py:173-207)
//
// Before, the unit was two launches: the k = 7 conv wrote Y = snake2(.) to HBM and the k = 1 conv read it back together with
// the residual.  The k = 1 layers are pure streams (ncu, profiles/r02_summary.md: 4.4-4.9 TB/s, tensor pipe 7-18 % active)
// and cost 35 % of the whole decode for 12 % of its FLOPs.  Here Y never leaves the SM:
//   P1  k = 7 implicit GEMM (row-shifted views of one activation window per 64-channel chunk, like dac_tc2.cuh) into a
//       TMEM accumulator of mt x C columns (double-buffered);
//   E1  epilogue warps drain it: + b1, snake2, fp16, and write it to shared memory in the K-major SWIZZLE_128B layout a
//       tcgen05 A operand needs (fence.proxy.async makes the generic-proxy stores visible to the tensor core);
//   P2  k = 1 GEMM straight from that shared-memory tile (weights through the same TMA ring) into a second, 96-column
//       TMEM region, one (row half, 96-channel) chunk at a time;
//   E2  + b2 + residual, raw output and the NEXT layer's Snake, coalesced stores (tc_epilogue_chunk of dac_tc.cuh).
// The MMA warp issues P1(t) then P2(t-1); the epilogue warps run E2(t-1) then E1(t): E1 / E2 of a tile overlap the P1 of
// the next one.  HBM traffic per unit: activations in + residual in + raw out + activated out (4 streams instead of 7).
//
// Shapes: C in {96, 192}; tile = mt x 128 time rows with mt * C == 192 (C = 96: 256-row tiles, C = 192: 128-row tiles), so
// TMEM = 2 x 192 (P1, double-buffered) + 96 (P2) = 480 columns.  Warp 0 = TMA producer, warp 1 = TMEM allocator + MMA
// issuer, warps 2..9 = epilogue (two per TMEM lane quarter).
#pragma once
#include "dac_tc2.cuh"

#define RU_EW 8
#define RU_THREADS (64 + 32 * RU_EW)
#define RU_MAX_A_SLOTS 4
#define RU_MAX_B_STAGES 8
#define RU_N2 96                       // output channels per P2 chunk
#define RU_TAB_BYTES (6 * 192 * 4)     // b1, alpha2, 1/alpha2, b2, alpha_next, 1/alpha_next

struct RuParams {
  int C, mt, kchunks;        // channels (96 / 192), 128-row halves per tile (2 / 1), ceil(C / 64)
  int dil, pad;              // k = 7 conv: dilation, padding (= 3 * dil)
  int T;                     // time rows per utterance
  int nb;                    // utterances (activation map is 3-D {C, T, nb})
  int m_tiles;               // tiles per utterance
  int box_rows, n_aload;     // activation window = n_aload TMA boxes of box_rows rows
  int a_bytes, a_slots;      // bytes per window slot, ring depth
  int b_stages;              // weight-tile ring depth (C * 128 bytes each)
  const float *bias1, *alpha2, *bias2, *alpha_next;
  const __half* residual;    // x (raw) [nb * T][C]
  __half* out_raw;           // optional
  __half* out_act;           // snake(out, alpha_next)
};

__device__ __forceinline__ void ru_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// dyn smem = 1024 + a_slots * a_bytes + b_stages * C * 128 + kchunks * mt * 128 * 128 + 512 + RU_TAB_BYTES
__global__ void __launch_bounds__(RU_THREADS, 1) k_dac_ru_fused(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w1,
                                                                const __grid_constant__ CUtensorMap map_w2, const RuParams p) {
  extern __shared__ unsigned char ru_smem_raw[];
  const uint32_t raw = smem_u32(ru_smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t b_bytes = (uint32_t)p.C * 128u;
  const uint32_t tile_rows = (uint32_t)p.mt * TC_BM;
  const uint32_t y_chunk = tile_rows * 128u;                     // one 64-channel K chunk of Y
  const uint32_t a_base = base;
  const uint32_t b_base = a_base + (uint32_t)p.a_slots * (uint32_t)p.a_bytes;
  const uint32_t y_base = b_base + (uint32_t)p.b_stages * b_bytes;
  const uint32_t bar_base = y_base + (uint32_t)p.kchunks * y_chunk;
  auto a_full = [&](int s) { return bar_base + 8u * (uint32_t)s; };
  auto a_empty = [&](int s) { return bar_base + 8u * (uint32_t)(RU_MAX_A_SLOTS + s); };
  auto b_full = [&](int s) { return bar_base + 8u * (uint32_t)(2 * RU_MAX_A_SLOTS + s); };
  auto b_empty = [&](int s) { return bar_base + 8u * (uint32_t)(2 * RU_MAX_A_SLOTS + RU_MAX_B_STAGES + s); };
  const uint32_t misc = bar_base + 8u * (uint32_t)(2 * RU_MAX_A_SLOTS + 2 * RU_MAX_B_STAGES);
  auto acc1_full = [&](int s) { return misc + 8u * (uint32_t)s; };
  auto acc1_free = [&](int s) { return misc + 8u * (uint32_t)(2 + s); };
  const uint32_t y_ready = misc + 32u, y_free = misc + 40u, acc2_full = misc + 48u, acc2_free = misc + 56u;
  const uint32_t tmem_slot = misc + 64u;
  float* s_tab = reinterpret_cast<float*>(ru_smem_raw + ((bar_base + 512u) - raw));
  float *s_b1 = s_tab, *s_a2 = s_tab + 192, *s_i2 = s_tab + 384, *s_b2 = s_tab + 576, *s_an = s_tab + 768, *s_in = s_tab + 960;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(ru_smem_raw + (tmem_slot - raw));
  unsigned char* y_ptr = ru_smem_raw + (y_base - raw);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = p.nb * p.m_tiles;
  const int n_mine = (int)blockIdx.x < total_tiles ? (total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w1) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w2) : "memory");
    for (int s = 0; s < p.a_slots; ++s) { mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1); }
    for (int s = 0; s < p.b_stages; ++s) { mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(acc1_full(s), 1); mbar_init(acc1_free(s), RU_EW); }
    mbar_init(y_ready, RU_EW); mbar_init(y_free, 1); mbar_init(acc2_full, 1); mbar_init(acc2_free, RU_EW);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = threadIdx.x; i < p.C; i += RU_THREADS) {
    s_b1[i] = __ldg(p.bias1 + i);
    const float a2 = __ldg(p.alpha2 + i);
    s_a2[i] = a2; s_i2[i] = 1.0f / (a2 + 1e-9f);
    s_b2[i] = __ldg(p.bias2 + i);
    const float an = __ldg(p.alpha_next + i);
    s_an[i] = an; s_in[i] = 1.0f / (an + 1e-9f);
  }
  // Y's K padding (C = 96: channels 96..127 of the second chunk) is never written by E1 and multiplied by TMA's zero
  // fill of the weight tile: it only has to be finite.  Zero all of Y once.
  for (uint32_t i = threadIdx.x * 16u; i < (uint32_t)p.kchunks * y_chunk; i += RU_THREADS * 16u) *reinterpret_cast<uint4*>(y_ptr + i) = make_uint4(0, 0, 0, 0);
  ru_fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot_ptr;
  const uint32_t acc2_col = 384u;                                // TMEM columns: [0,192) acc1[0], [192,384) acc1[1], [384,480) acc2

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer: mirrors the MMA warp's consumption order: P1(it), then P2(it - 1) =====
      int slot = 0, st = 0;
      uint32_t a_par = 1, b_par = 1;
      const uint32_t a_tx = (uint32_t)(p.n_aload * p.box_rows) * 128u;
      for (int it = 0; it <= n_mine; ++it) {
        if (it < n_mine) {
          const int tile = blockIdx.x + it * gridDim.x;
          const int ub = tile / p.m_tiles, mi = tile - ub * p.m_tiles;
          const int row0 = mi * (int)tile_rows - p.pad;
          for (int kc = 0; kc < p.kchunks; ++kc) {
            mbar_wait(a_empty(slot), a_par);
            const uint32_t a_dst = a_base + (uint32_t)slot * (uint32_t)p.a_bytes;
            mbar_expect_tx(a_full(slot), a_tx);
            tma_load_3d(a_dst, &map_x, kc * TC_BK, row0, ub, a_full(slot));
            if (p.n_aload == 2) tma_load_3d(a_dst + (uint32_t)p.box_rows * 128u, &map_x, kc * TC_BK, row0 + p.box_rows, ub, a_full(slot));
            if (++slot == p.a_slots) { slot = 0; a_par ^= 1u; }
            for (int tap = 0; tap < 7; ++tap) {
              mbar_wait(b_empty(st), b_par);
              mbar_expect_tx(b_full(st), b_bytes);
              tma_load_2d(b_base + (uint32_t)st * b_bytes, &map_w1, kc * TC_BK, tap * p.C, b_full(st));
              if (++st == p.b_stages) { st = 0; b_par ^= 1u; }
            }
          }
        }
        if (it >= 1) {
          for (int ch = 0; ch < 2; ++ch) {                       // chunk = (row half, 96-channel block): mt * C / 96 == 2 of them
            const int c2 = p.mt == 2 ? 0 : ch;
            for (int kc = 0; kc < p.kchunks; ++kc) {
              mbar_wait(b_empty(st), b_par);
              mbar_expect_tx(b_full(st), RU_N2 * 128u);
              tma_load_2d(b_base + (uint32_t)st * b_bytes, &map_w2, kc * TC_BK, c2 * RU_N2, b_full(st));
              if (++st == p.b_stages) { st = 0; b_par ^= 1u; }
            }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      const uint32_t idesc1 = umma_idesc_f16(p.C), idesc2 = umma_idesc_f16(RU_N2);
      const uint64_t desc_hi = umma_desc_sw128(0);
      int slot = 0, st = 0;
      uint32_t a_par = 0, b_par = 0, f1_par[2] = {1, 1}, yr_par = 0, a2f_par = 1;
      for (int it = 0; it <= n_mine; ++it) {
        if (it < n_mine) {
          const int buf = it & 1;
          mbar_wait(acc1_free(buf), f1_par[buf]);
          f1_par[buf] ^= 1u;
          tc_fence_after();
          const uint32_t d0 = tmem_d + (uint32_t)(buf * 192);
          uint32_t first = 0;
          for (int kc = 0; kc < p.kchunks; ++kc) {
            mbar_wait(a_full(slot), a_par);
            uint64_t adesc = desc_hi | (uint64_t)((a_base + (uint32_t)slot * (uint32_t)p.a_bytes) >> 4);
            for (int tap = 0; tap < 7; ++tap) {
              mbar_wait(b_full(st), b_par);
              tc_fence_after();
              const uint64_t bdesc = desc_hi | (uint64_t)((b_base + (uint32_t)st * b_bytes) >> 4);
#pragma unroll
              for (int k = 0; k < TC_BK / 16; ++k) {
                tc_mma_f16(d0, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc1, first | (uint32_t)k);
                if (p.mt == 2) tc_mma_f16(d0 + (uint32_t)p.C, adesc + (uint64_t)(TC_BM * 8 + 2 * k), bdesc + (uint64_t)(2 * k), idesc1, first | (uint32_t)k);
              }
              first = 1;
              tc_commit(b_empty(st));
              if (++st == p.b_stages) { st = 0; b_par ^= 1u; }
              adesc += (uint64_t)(p.dil * 8);                    // next tap: window rows shifted by the dilation
            }
            tc_commit(a_empty(slot));
            if (++slot == p.a_slots) { slot = 0; a_par ^= 1u; }
          }
          tc_commit(acc1_full(buf));
        }
        if (it >= 1) {
          mbar_wait(y_ready, yr_par);                            // E1(it - 1) wrote Y
          yr_par ^= 1u;
          tc_fence_after();
          for (int ch = 0; ch < 2; ++ch) {
            const int half = p.mt == 2 ? ch : 0;
            mbar_wait(acc2_free, a2f_par);
            a2f_par ^= 1u;
            tc_fence_after();
            for (int kc = 0; kc < p.kchunks; ++kc) {
              mbar_wait(b_full(st), b_par);
              tc_fence_after();
              const uint64_t ydesc = desc_hi | (uint64_t)((y_base + (uint32_t)kc * y_chunk + (uint32_t)half * (TC_BM * 128u)) >> 4);
              const uint64_t bdesc = desc_hi | (uint64_t)((b_base + (uint32_t)st * b_bytes) >> 4);
#pragma unroll
              for (int k = 0; k < TC_BK / 16; ++k) tc_mma_f16(tmem_d + acc2_col, ydesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc2, (uint32_t)(kc | k));
              tc_commit(b_empty(st));
              if (++st == p.b_stages) { st = 0; b_par ^= 1u; }
            }
            tc_commit(acc2_full);
          }
          tc_commit(y_free);                                     // every MMA that reads Y has retired when this fires
        }
      }
    }
    __syncwarp();
  } else {
    // ===== epilogue warps: E2(it - 1), then E1(it) =====
    const int lg = warp & 3;                                     // TMEM lane quarter this warp may read
    const int qi = (warp - 2) >> 2;                              // 0 / 1: which of the quarter's two warps
    const bool odd = lane & 1;
    uint32_t f_par[2] = {0, 0}, yf_par = 1, a2_par = 0;
    for (int it = 0; it <= n_mine; ++it) {
      if (it >= 1) {
        // ---- E2: out = acc2 + b2 + x, raw + next Snake
        const int tile = blockIdx.x + (it - 1) * gridDim.x;
        const int ub = tile / p.m_tiles, mi = tile - ub * p.m_tiles;
        for (int ch = 0; ch < 2; ++ch) {
          const int half = p.mt == 2 ? ch : 0, c2 = p.mt == 2 ? 0 : ch;
          const int m = mi * (int)tile_rows + half * TC_BM + lg * 32 + lane;
          const bool valid = m < p.T;
          const int vA = __shfl_sync(0xffffffffu, (int)valid, lane & ~1), vB = __shfl_sync(0xffffffffu, (int)valid, lane | 1);
          const int mA = __shfl_sync(0xffffffffu, m, lane & ~1), mB = __shfl_sync(0xffffffffu, m, lane | 1);
          const size_t ubase = (size_t)ub * p.T;
          const size_t rowA = (ubase + (size_t)(vA ? mA : 0)) * p.C + c2 * RU_N2, rowB = (ubase + (size_t)(vB ? mB : 0)) * p.C + c2 * RU_N2;
          uint4 res[2][4];
#pragma unroll
          for (int ci = 0; ci < 2; ++ci) {
            const int c0 = (qi + 2 * ci) * 32;
            if (c0 < RU_N2) tc_res_load(p.residual, rowA + c0, rowB + c0, vA != 0, vB != 0, odd, res[ci]);
          }
          mbar_wait(acc2_full, a2_par);
          a2_par ^= 1u;
          tc_fence_after();
#pragma unroll
          for (int ci = 0; ci < 2; ++ci) {
            const int c0 = (qi + 2 * ci) * 32;
            if (c0 < RU_N2) {
              uint32_t v[32];
              tmem_ld32(tmem_d + ((uint32_t)(lg * 32) << 16) + acc2_col + (uint32_t)c0, v);
              const int cg = c2 * RU_N2 + c0;
              tc_epilogue_chunk(v, s_b2 + cg, s_an + cg, s_in + cg, true, res[ci], p.out_raw, p.out_act, rowA + c0, rowB + c0, vA != 0, vB != 0, odd);
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acc2_free);
        }
      }
      if (it < n_mine) {
        // ---- E1: Y = fp16(snake2(acc1 + b1)) -> shared memory, K-major SWIZZLE_128B (the A operand of P2)
        const int buf = it & 1;
        mbar_wait(acc1_full(buf), f_par[buf]);
        f_par[buf] ^= 1u;
        mbar_wait(y_free, yf_par);                               // P2(it - 1) no longer reads Y
        yf_par ^= 1u;
        tc_fence_after();
        const int units = p.mt * (p.C / 32);                     // 6 (half, 32-column chunk) units per quarter, 3 per warp
        for (int u = qi; u < units; u += 2) {
          const int half = u / (p.C / 32), cc = u - half * (p.C / 32);
          uint32_t v[32];
          tmem_ld32(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)(buf * 192 + half * p.C + cc * 32), v);
          const int r = half * TC_BM + lg * 32 + lane;             // row inside the tile
          const int c0 = cc * 32;
          unsigned char* yrow = y_ptr + (size_t)(c0 >> 6) * y_chunk + (size_t)r * 128;
          const int ju0 = (c0 & 63) >> 3;                          // first 16-byte unit of this chunk inside the 128-byte row
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 o;
            __half2* h = reinterpret_cast<__half2*>(&o);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int c = c0 + q * 8 + 2 * j;
              const float x0 = __uint_as_float(v[q * 8 + 2 * j]) + s_b1[c], x1 = __uint_as_float(v[q * 8 + 2 * j + 1]) + s_b1[c + 1];
              h[j] = __floats2half2_rn(snake_fast(x0, s_a2[c], s_i2[c]), snake_fast(x1, s_a2[c + 1], s_i2[c + 1]));
            }
            *reinterpret_cast<uint4*>(yrow + (((ju0 + q) ^ (r & 7)) << 4)) = o;
          }
        }
        ru_fence_async_smem();                                   // generic-proxy stores -> visible to the tensor core's async proxy
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { mbar_arrive(y_ready); mbar_arrive(acc1_free(buf)); }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512u) : "memory");
  }
}
