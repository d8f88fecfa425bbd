// DAC 44 kHz decoder — persistent tcgen05 implicit-GEMM convolution, second generation (impl=0 default).
//
// Same arithmetic as k_dac_conv_tc (dac_tc.cuh): D[t, co] = sum_tap sum_ci X[t + off(tap), ci] * W[tap][co, ci], fp16
// operands, fp32 accumulation in TMEM, fused bias / residual / next-layer Snake epilogue.  What changed is how the
// operands reach the tensor core.  The first kernel re-loaded the (almost identical) 128-row activation tile once per
// tap and its B tile once per 128 rows; ncu launch lists put every k=7 layer at 12-14 TB/s of L2->shared-memory fill,
// i.e. bound by operand delivery, not by the MMA pipe (profiles/r01_summary.md, "DAC").  Here
//   * the activation WINDOW (tile rows + the halo all taps need) is loaded once per 64-channel chunk and every tap is a
//     row-shifted view of it: a SWIZZLE_128B K-major descriptor may start at any 128-byte row of the window because the
//     swizzle is a function of the absolute shared-memory address (scripts/desc_offset_test.cu, verified on B200);
//   * a tile is 256 time rows: each weight tile feeds two accumulators, halving the weight traffic per output row;
//   * CTAs are persistent (one per SM) and the accumulators are double-buffered in TMEM (2 stages x 2 x <=128 columns),
//     so the epilogue of tile i overlaps the MMAs of tile i+1 while the TMA producer runs ahead across tile borders.
// Warp roles: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2.. = EW epilogue warps.  A warp can
// only read the TMEM lane quarter (warp % 4); the EW/4 warps of a quarter split the accumulators and their 32-column
// chunks.  EW = 8 is the default everywhere; the k=1 layers are bound by their residual / output streams and prefetch
// all residual rows of a tile into registers before the accumulator wait.  EW = 16 (zb_dac_set_option "ew_multi" /
// "ew_k1") was measured in the same process with scripts/dac_ab.py: no gain on the multi-tap layers, a loss on k=1.
#pragma once
#include "dac_tc.cuh"

#define TC2_MAX_COUT 1536
#define TC2_TAB_BYTES (3 * TC2_MAX_COUT * 4)
#define TC2_MAX_A_SLOTS 4
#define TC2_MAX_B_STAGES 12

struct TcConv2Params {
  int ntaps, dil, stride, pad;
  int kchunks;               // ceil(Cin / 64)
  int m_rows;                // GEMM rows (conv: T; convT: T_in + 1)
  int T_out, Cout;
  int n_tile, n_tiles;       // output channels per tile (multiple of 32, <= 128 when acc_stages == 2)
  int m_tiles, phases;       // tiles per utterance = m_tiles * n_tiles * phases
  int nb;                    // utterances in this launch: total tiles = nb * tiles per utterance, the activation map is 3-D {C, T, nb}
  int mt;                    // accumulators (128-row halves) per tile: 1 or 2
  int lead;                  // window row 0 = tile row 0 - lead  (conv: pad; convT: ntaps - 1)
  int box_rows, n_aload;     // the window is n_aload TMA boxes of box_rows rows (box_rows % 8 == 0)
  int a_bytes, a_slots;      // bytes per window slot (multiple of 1024), ring depth
  int b_stages;              // weight-tile ring depth (n_tile * 128 bytes each)
  int acc_cols, acc_stages;  // TMEM columns per accumulator, accumulator stages (mt * acc_cols * acc_stages <= 512)
  const float* bias;
  const __half* residual;
  __half* out_raw;
  __half* out_act;
  const float* alpha_next;
};

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// EW = epilogue warps (8 or 16).  grid = min(tiles, SMs); block = 64 + 32*EW;
// dyn smem = 1024 + a_slots*a_bytes + b_stages*n_tile*128 + 512 + TC2_TAB_BYTES
template <int EW>
__global__ void __launch_bounds__(64 + 32 * EW, 1) k_dac_conv_tc2(const __grid_constant__ CUtensorMap map_x,
                                                                 const __grid_constant__ CUtensorMap map_w, const TcConv2Params p) {
  extern __shared__ unsigned char tc2_smem_raw[];
  const uint32_t raw = smem_u32(tc2_smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t b_bytes = (uint32_t)p.n_tile * 128u;
  const uint32_t a_base = base;
  const uint32_t b_base = a_base + (uint32_t)p.a_slots * (uint32_t)p.a_bytes;
  const uint32_t bar_base = b_base + (uint32_t)p.b_stages * b_bytes;
  auto a_full = [&](int s) { return bar_base + 8u * (uint32_t)s; };
  auto a_empty = [&](int s) { return bar_base + 8u * (uint32_t)(TC2_MAX_A_SLOTS + s); };
  auto b_full = [&](int s) { return bar_base + 8u * (uint32_t)(2 * TC2_MAX_A_SLOTS + s); };
  auto b_empty = [&](int s) { return bar_base + 8u * (uint32_t)(2 * TC2_MAX_A_SLOTS + TC2_MAX_B_STAGES + s); };
  auto t_full = [&](int s) { return bar_base + 8u * (uint32_t)(2 * TC2_MAX_A_SLOTS + 2 * TC2_MAX_B_STAGES + s); };
  auto t_empty = [&](int s) { return bar_base + 8u * (uint32_t)(2 * TC2_MAX_A_SLOTS + 2 * TC2_MAX_B_STAGES + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (uint32_t)(2 * TC2_MAX_A_SLOTS + 2 * TC2_MAX_B_STAGES + 4);
  float* s_bias = reinterpret_cast<float*>(tc2_smem_raw + ((bar_base + 512u) - raw));
  float* s_alpha = s_bias + TC2_MAX_COUT;
  float* s_inv = s_alpha + TC2_MAX_COUT;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(tc2_smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_per_utt = p.m_tiles * p.n_tiles * p.phases;
  const int total_tiles = p.nb * tiles_per_utt;
  const int tile_rows = p.mt * TC_BM;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
    for (int s = 0; s < p.a_slots; ++s) { mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1); }
    for (int s = 0; s < p.b_stages; ++s) { mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(t_full(s), 1); mbar_init(t_empty(s), EW); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = threadIdx.x; i < p.Cout; i += 64 + 32 * EW) {
    s_bias[i] = __ldg(p.bias + i);
    const float a = p.out_act ? __ldg(p.alpha_next + i) : 1.f;
    s_alpha[i] = a;
    s_inv[i] = 1.0f / (a + 1e-9f);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot_ptr;

  // Ring bookkeeping is kept to an increment, a compare and a parity flip per step: the issuing threads are single
  // threads, and integer division by a runtime ring depth per stage made the first version issue-bound (3x slower
  // than the MMAs it was feeding).
  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer: windows and weight tiles, running ahead across tile borders =====
      // a wait on the parity "before" a fresh barrier's first phase returns at once, so no "first round" special case
      int slot = 0, st = 0;
      uint32_t a_par = 1, b_par = 1;
      const uint32_t a_tx = (uint32_t)(p.n_aload * p.box_rows) * 128u;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int ub = tile / tiles_per_utt, tl = tile - ub * tiles_per_utt;
        const int mi = tl % p.m_tiles, rest = tl / p.m_tiles;
        const int n0 = (rest % p.n_tiles) * p.n_tile, phase = rest / p.n_tiles;
        const int row0 = mi * tile_rows - p.lead;
        const int w_base = (p.stride > 1 ? phase : 0) * p.Cout + n0;
        const int w_step = (p.stride > 1 ? p.stride : 1) * p.Cout;
        for (int kc = 0; kc < p.kchunks; ++kc) {
          mbar_wait(a_empty(slot), a_par);
          const uint32_t a_dst = a_base + (uint32_t)slot * (uint32_t)p.a_bytes;
          mbar_expect_tx(a_full(slot), a_tx);
          tma_load_3d(a_dst, &map_x, kc * TC_BK, row0, ub, a_full(slot));
          if (p.n_aload == 2) tma_load_3d(a_dst + (uint32_t)p.box_rows * 128u, &map_x, kc * TC_BK, row0 + p.box_rows, ub, a_full(slot));
          if (++slot == p.a_slots) { slot = 0; a_par ^= 1u; }
          int w_row = w_base;
          for (int tap = 0; tap < p.ntaps; ++tap, w_row += w_step) {
            mbar_wait(b_empty(st), b_par);
            mbar_expect_tx(b_full(st), b_bytes);
            tma_load_2d(b_base + (uint32_t)st * b_bytes, &map_w, kc * TC_BK, w_row, b_full(st));
            if (++st == p.b_stages) { st = 0; b_par ^= 1u; }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      const uint32_t idesc = umma_idesc_f16(p.n_tile);
      const uint64_t desc_hi = umma_desc_sw128(0);             // descriptor with a zero start address
      int slot = 0, st = 0, as = 0;
      uint32_t a_par = 0, b_par = 0, t_par = 1;
      const int shift0 = p.stride > 1 ? p.ntaps - 1 : 0;
      const int shift_step = p.stride > 1 ? -1 : p.dil;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(t_empty(as), t_par);
        tc_fence_after();
        const uint32_t d0 = tmem_d + (uint32_t)(as * p.mt * p.acc_cols);
        uint32_t first = 0;                                    // 0 for the very first MMA of the tile (overwrite)
        for (int kc = 0; kc < p.kchunks; ++kc) {
          mbar_wait(a_full(slot), a_par);
          // conv: window row r holds time m0 - pad + r, tap reads m + tap*dil - pad  -> shift tap*dil
          // convT: window row r holds time m0 - (ntaps-1) + r, tap reads m - tap     -> shift ntaps-1-tap
          uint64_t adesc = desc_hi | (uint64_t)(((a_base + (uint32_t)slot * (uint32_t)p.a_bytes) >> 4) + (uint32_t)(shift0 * 8));
          for (int tap = 0; tap < p.ntaps; ++tap) {
            mbar_wait(b_full(st), b_par);
            tc_fence_after();
            const uint64_t bdesc = desc_hi | (uint64_t)((b_base + (uint32_t)st * b_bytes) >> 4);
#pragma unroll
            for (int k = 0; k < TC_BK / 16; ++k) {
              tc_mma_f16(d0, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, first | (uint32_t)k);
              if (p.mt == 2)
                tc_mma_f16(d0 + (uint32_t)p.acc_cols, adesc + (uint64_t)(TC_BM * 8 + 2 * k), bdesc + (uint64_t)(2 * k), idesc, first | (uint32_t)k);
            }
            first = 1;
            tc_commit(b_empty(st));
            if (++st == p.b_stages) { st = 0; b_par ^= 1u; }
            adesc += (uint64_t)(int64_t)(shift_step * 8);
          }
          tc_commit(a_empty(slot));
          if (++slot == p.a_slots) { slot = 0; a_par ^= 1u; }
        }
        tc_commit(t_full(as));
        if (++as == p.acc_stages) { as = 0; t_par ^= 1u; }
      }
    }
    __syncwarp();
  } else {
    // ===== epilogue =====
    const int lg = warp & 3;
    const int ew = warp - 2;                                   // 0..EW-1; quarter-mates = EW/4 warps
    constexpr int QM = EW / 4;                                 // warps per TMEM lane quarter
    constexpr int NCH = 8 / QM;                                // max 32-column chunks per warp (n_tile <= 128, <= 2 accumulators)
    const int qi = ew >> 2;                                    // index among the quarter-mates
    const int acc = p.mt == 2 ? (qi & 1) : 0;
    const int c_first = (p.mt == 2 ? (qi >> 1) : qi) * 32;
    const int c_step = (p.mt == 2 ? QM / 2 : QM) * 32;
    int as = 0;
    uint32_t t_par = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int ub = tile / tiles_per_utt, tl = tile - ub * tiles_per_utt;
      const int mi = tl % p.m_tiles, rest = tl / p.m_tiles;
      const int n0 = (rest % p.n_tiles) * p.n_tile, phase = rest / p.n_tiles;
      const int m = mi * tile_rows + acc * TC_BM + lg * 32 + lane;
      const int out_row = p.stride > 1 ? m * p.stride + phase - p.pad : m;
      const bool valid = (m < p.m_rows) && (out_row >= 0) && (out_row < p.T_out);
      const bool odd = lane & 1;
      const int vA = __shfl_sync(0xffffffffu, (int)valid, lane & ~1), vB = __shfl_sync(0xffffffffu, (int)valid, lane | 1);
      const int orA = __shfl_sync(0xffffffffu, out_row, lane & ~1), orB = __shfl_sync(0xffffffffu, out_row, lane | 1);
      const size_t ubase = (size_t)ub * p.T_out;
      const size_t rowA = (ubase + (size_t)(vA ? orA : 0)) * p.Cout + n0, rowB = (ubase + (size_t)(vB ? orB : 0)) * p.Cout + n0;
      ZB_ASSERT(ub >= 0 && ub < p.nb && n0 >= 0 && n0 < p.n_tiles * p.n_tile);
      ZB_ASSERT(!vA || (orA >= 0 && orA < p.T_out));
      ZB_ASSERT(!vB || (orB >= 0 && orB < p.T_out));
      ZB_ASSERT((as * p.mt + acc + 1) * p.acc_cols <= 512);
      // the warp's residual rows of the whole tile are requested before the accumulator wait
      uint4 res[NCH][4];
      if (p.residual) {
#pragma unroll
        for (int ci = 0; ci < NCH; ++ci) {
          const int c0 = c_first + ci * c_step;
          if (c0 < p.n_tile) tc_res_load(p.residual, rowA + c0, rowB + c0, vA != 0, vB != 0, odd, res[ci]);
        }
      }
      mbar_wait(t_full(as), t_par);
      tc_fence_after();
      const uint32_t t0 = tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)((as * p.mt + acc) * p.acc_cols);
#pragma unroll
      for (int ci = 0; ci < NCH; ++ci) {
        const int c0 = c_first + ci * c_step;
        if (c0 < p.n_tile) {
          uint32_t v[32];
          tmem_ld32(t0 + (uint32_t)c0, v);
          tc_epilogue_chunk(v, s_bias + n0 + c0, s_alpha + n0 + c0, s_inv + n0 + c0, p.residual != nullptr, res[ci], p.out_raw, p.out_act,
                            rowA + c0, rowB + c0, vA != 0, vB != 0, odd);
        }
      }
      // accumulator stage drained (tcgen05.ld completed inside tmem_ld32): hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_empty(as));
      if (++as == p.acc_stages) { as = 0; t_par ^= 1u; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512u) : "memory");
  }
}
