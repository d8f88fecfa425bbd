// Persistent decode kernel: the WHOLE generation step (embedding gather-sum, n_layer x [LN+QKV+RoPE+KV-append, split-KV
// GQA attention, out_proj+residual, LN+fc1+SwiGLU, fc2+residual], final LN + 9 heads + CFG, sampling, EOS wind-down and
// frame write-back) — and any number of consecutive steps — in ONE cooperative launch of one CTA per SM.
// Reference arithmetic: zonos/model.py:97-116,276-301, zonos/backbone/_torch.py:83-152, zonos/sampling.py:117-182.
//
// Why: at batch 1 the step is a chain of 113 GEMVs that together stream 3.2 GB of weights.  As separate
// kernels each GEMV pays launch + ramp-up + tail (measured r01: 5.8-19.5 us per launch where the bytes alone need
// 1.3-10 us; the step ran at 32 % of HBM peak).  Nothing that is streamed depends on the activations, so here the
// stream NEVER stops:
//   * warp 8 of every CTA is a TMA producer.  It walks the fixed, activation-independent schedule of everything this
//     CTA will need from HBM — its 8-row weight tiles (one 4 KB `cp.async.bulk` per row into a padded,
//     bank-conflict-free shared-memory stage; 2 KB copies ran at only 4.0 TB/s, 4 KB copies at 7.0 TB/s:
//     profiles/r01_ring_bench.md), the LayerNorm parameters, the RoPE rows and its slice of the KV cache — through a
//     ring of 33 KB stages with full/empty mbarriers, running ahead of the math across phase, layer and step
//     boundaries.  With the HBM queues kept full by that stream, a cold demand load costs 5-10 us
//     (profiles/r01_mega_trace_v1.md), so NO consumer ever touches HBM-cold data directly;
//   * warps 0..7 are consumers: per phase they stage the (tiny, L2-resident) activations in shared memory (LayerNorm
//     fused), then eat stages with `mma.sync.m16n8k16` (activations = A rows 0..R-1, 8 weight rows = B columns; K is
//     split over the 8 warps and reduced through shared memory), run store-only fused epilogues (the CTA keeps its own
//     slice of the residual stream in shared memory), and meet the other CTAs at a grid barrier;
//   * rows are dealt to CTAs in contiguous, balanced ranges (<= 1 unit difference), whole rows per CTA, so there is
//     no cross-CTA split-K and the fp32 summation order — hence every token — is reproducible.
// tcgen05 is deliberately not used here: with M = 2..8 activation rows the contraction is a GEMV, ~100x below the
// tensor roofline and bound by the HBM stream; `mma.sync` from shared memory costs ~0.4 issue slots per weight element.
#pragma once
#include "attention_kernels.cuh"
#include "decode_kernels.cuh"
#include "gemv_mma.cuh"
#include "sample_kernels.cuh"

#define MG_CONSUMERS 256
#define MG_THREADS 320               // 8 MMA consumer warps + 1 producer warp + 1 epilogue warp
#define MG_EPI_WARP 9
#define MG_HANDOFF 288               // threads on the tile hand-off barriers: the 8 MMA warps + the epilogue warp
// named barriers: 0 __syncthreads, 1 SyncConsumers (8 MMA warps), 2/3 partial buffer 0/1 full, 4/5 partial buffer 0/1 free,
// 7 step start
enum { MG_BAR_FULL = 2, MG_BAR_FREE = 4, MG_BAR_STEP = 7 };
#define MG_WARPS 8
#define MG_KC 2048                   // weight columns per ring stage (4 KB per row copy)
#define MG_ROW_BYTES (MG_KC * 2 + 64)   // stage row stride: +64 B => 8 rows x 16 B quarter-warp reads hit 32 distinct banks
#define MG_STAGE_BYTES (8 * MG_ROW_BYTES)
#define MG_MAX_STAGES 8
#define MG_XPAD 32                   // bf16 elements of padding per staged activation row
#define MG_KV_TOK 64                 // tokens per KV stage: K rows at [0, 16 KB), V rows at [16 KB, 32 KB)
#define MG_MAX_SPLIT 16             // KV splits per (row, kv head): the merge stages all partials in the s_o scratch ([16][G][128])
#define MG_OWN 16                    // residual-stream columns a CTA can own (2 * ceil(d/2 / grid) <= 16)
#define MG_ROPE_BYTES (ZB_DH / 2 * 2 * 4)   // one RoPE table row: 64 x (cos, sin) fp32

enum { MG_MODE_HIDDEN = 0, MG_MODE_FORCED = 1, MG_MODE_LOGITS = 2, MG_MODE_DECODE = 3 };

struct MegaLayer {
  const bf16 *ln1_w, *ln1_b, *w_in, *w_out, *ln2_w, *ln2_b, *w_fc1, *w_fc2;
};

struct MegaParams {
  MegaLayer l0;                // layer 0; layer l = l0 + l * layer_stride bytes (the arena is layer-major with a constant
  long long layer_stride;      // stride), so no role ever loads a weight POINTER from memory: under a saturated HBM stream
                               // even an L1 miss on a pointer table costs microseconds
  int n_layer, d, q_size, kv_size, F, H, Hkv;
  const bf16 *normf_w, *normf_b, *heads, *emb;
  const float* rope;           // [table_len][Dh/2][2]
  int n_q, V, eos, mask_tok;
  float eps, scale_log2;
  // activations: global scratch, L2 resident, read with ld.cg (they are written by other CTAs of this launch)
  bf16 *x, *q, *y, *h;         // [R][d], [R][q_size], [R][q_size], [R][F]
  float* logits;               // [B][n_q][V]
  bf16* kv;                    // [n_layer][page][2][Hkv][ZB_PAGE][ZB_DH]
  size_t kv_layer_elems;
  const int* page_table;       // [R][max_pages]
  int max_pages;
  unsigned long long* att_ll;  // split-KV partials as (fp32, epoch) words: [R][Hkv][MG_MAX_SPLIT][G * (Dh + 2)]
  int nsplit, min_chunk, single_ctx;
  int R, B;                    // activation rows (= 2B in the CFG modes) and utterances
  ZbDevState* st;
  int* rope_pos;               // [R]
  long long* delayed;          // [B][n_q][L]
  long long* tokens;           // [B][n_q]
  int *remaining, *stopping, *out_len;
  unsigned* bar_counter;       // grid barrier (zero at launch)
  unsigned long long* x_ll;    // [2][8][d/2] residual stream as (bf16 pair, epoch) 8-byte words: data that carries its own 'ready' flag
  unsigned long long* qkv_ll;  // [8][(q_size + 2 kv_size)/2]: this step's q, k, v rows of the current layer, tagged the same way
  unsigned long long* y_ll;    // [8][q_size/2]: attention output
  unsigned long long* h_ll;    // [8][F/2]: SwiGLU output
  int use_ll;                  // 1: LayerNorm staging polls x_ll and the grid barriers after out_proj / fc2 are skipped;
                               // 2: q / k / v and y travel tagged too: no grid barrier around the attention; 3: and h: no barrier in the layer
  unsigned epoch_base;         // launch number << 20: epochs never repeat while an x_ll word can still be around
  int mode, n_steps;
  const bf16* hidden_in;       // MG_MODE_HIDDEN: x <- hidden_in[r, t0 + i, :], row stride hidden_S tokens
  bf16* hidden_out;            // optional: post-norm_f hidden states, same indexing
  int hidden_S, hidden_t0;
  int n_stages;                // ring stages
  unsigned producer_sleep_ns;  // producer back-off between polls of a full ring (0 = hot spin)
  int scratch_bytes;           // activation staging / attention / sampling scratch
  int chain;                   // use_ll >= 3: 1 = the MMA warps do not wait for their epilogue warp between the phases of the layer stack
                               // (measured SLOWER, 918 -> 969 us per step: the pollers' loads delay the epilogue warp's stores; default 0)
  int stream_evict_first;      // 1: weight tiles are loaded with the L2 evict_first policy (A/B switch "evict")
  unsigned long long* dbg;     // optional trace: [grid][MG_TRACE_SLOTS] globaltimer stamps of step dbg_step (thread 0)
  int dbg_step;
  int dbg_kind;                // stage-level trace (dbg_fine == 1): which phase kind of layer 1 is stamped (EPI_SWIGLU = fc1, EPI_QKV, EPI_RESID)
  int dbg_fine;                // trace variant: stage-level stamps of ONE fc1 phase (layer 1 of step dbg_step) instead of the per-phase timeline
};
#define MG_TRACE_SLOTS 64
__device__ __forceinline__ unsigned long long mg_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define MG_STAMP()                                                                                         \
  do {                                                                                                     \
    if (cx.trace && threadIdx.x == 0 && cx.trace_n < MG_TRACE_SLOTS) cx.trace_out[cx.trace_n++] = mg_globaltimer(); \
  } while (0)

#define MG_ASTAMP()                                                                                        \
  do {                                                                                                     \
    if (cx.afine && threadIdx.x == 0 && cx.trace_n < MG_TRACE_SLOTS) cx.trace_out[cx.trace_n++] = mg_globaltimer(); \
  } while (0)

#define MG_LSTAMP()                                                                                        \
  do {                                                                                                     \
    if (cx.lfine && threadIdx.x == 0 && cx.trace_n < MG_TRACE_SLOTS) cx.trace_out[cx.trace_n++] = mg_globaltimer(); \
  } while (0)
#define MG_TSTAMP()                                                                                        \
  do {                                                                                                     \
    if (cx.tfine && threadIdx.x == 0 && cx.trace_n < MG_TRACE_SLOTS) cx.trace_out[cx.trace_n++] = mg_globaltimer(); \
  } while (0)

// ---- PTX helpers ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t mg_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mg_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mg_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mg_mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mg_mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mg_mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mg_mbar_try_wait(bar, parity)) {}
}
// 1-D bulk copy global -> shared, completion counted on an mbarrier (TMA engine, no tensor map needed)
__device__ __forceinline__ void mg_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar)
               : "memory");
}
// The same with an L2 eviction-priority hint.  The weight stream (3.2 GB per step, each byte used once) is marked
// evict_first so that it does not push everything else out of the 126 MB L2 every step: the kernel's own instructions
// (311 KB of SASS, refetched by every SM), the KV cache of the previous steps, the activations and the embedding rows.
// The step-tail trace (--opt trace=4) showed code that runs once per step (sampling, state machine, embedding gather)
// taking 9-22 us for a handful of dependent L2 round trips: its instructions came from DRAM every step.
__device__ __forceinline__ void mg_bulk_g2s_hint(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t policy) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar), "l"(policy)
               : "memory");
}
__device__ __forceinline__ uint64_t mg_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void mg_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void mg_bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ uint4 mg_lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ uint4 mg_ldcg16(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ int mg_ld_volatile(const int* p) { return *reinterpret_cast<const volatile int*>(p); }

// N groups of 8 consecutive bf16 (= 4 tagged words (bf16 pair, epoch) each) from an epoch-tagged buffer: all loads of a
// round are in flight together; the round is repeated until every word carries `epoch`.
template <int N>
__device__ __forceinline__ void mg_ll_load(const unsigned long long* const (&wp)[N], unsigned epoch, uint4 (&out)[N]) {
  // (Re-reading only the groups that are still incomplete keeps every group's registers alive around the loop: with the
  // 168 registers a thread of this 10-warp CTA can have, that version spilled ~1 KB per thread.)
  unsigned long long w[N][4];
  for (;;) {
#pragma unroll
    for (int i = 0; i < N; ++i) {
      asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[i][0]), "=l"(w[i][1]) : "l"(wp[i]) : "memory");
      asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[i][2]), "=l"(w[i][3]) : "l"(wp[i] + 2) : "memory");
    }
    bool ok = true;
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) ok = ok && (unsigned)(w[i][j] >> 32) == epoch;
    if (ok) break;
  }
#pragma unroll
  for (int i = 0; i < N; ++i) out[i] = make_uint4((unsigned)w[i][0], (unsigned)w[i][1], (unsigned)w[i][2], (unsigned)w[i][3]);
}
__device__ __forceinline__ void mg_ll_store(unsigned long long* wp, unsigned epoch, unsigned pair_bits) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(wp), "l"(((unsigned long long)epoch << 32) | (unsigned long long)pair_bits) : "memory");
}

// ---- grid barrier over the consumer threads of all CTAs -----------------------------------------------------------
__device__ __forceinline__ void mg_grid_sync(unsigned* counter, unsigned& target) {
  SyncConsumers::sync();
  if (threadIdx.x == 0) {
    target += gridDim.x;
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
    unsigned v;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
    } while (v < target);
  }
  SyncConsumers::sync();
}

// Epoch of the x_ll words written by out_proj (which = 0) / fc2 (which = 1) of layer l in step `step` of this launch.
// Never 0 (the buffer is zero-filled at creation); the two phases write different halves of x_ll.
__device__ __forceinline__ unsigned mg_ll_epoch(const MegaParams& p, int step, int l, int which) {
  return p.epoch_base | (((unsigned)step & 0x1fffu) << 7) | (unsigned)(2 * l + which + 1);
}

// ---- per-CTA context ----------------------------------------------------------------------------------------------
struct MgCtx {
  unsigned char* ring;       // n_stages x MG_STAGE_BYTES
  uint32_t ring_addr;        // the same, as a shared-space address
  unsigned char* scratch;    // activation staging etc.
  float* part;               // [2][2][8 warps][8 r][8 n]
  float* red;                // [160]: LayerNorm partials [2][8 warps][8 rows] + 32 spare
  float* s_rope;             // [8][128]: (cos, sin) rows of this step's positions
  float* s_xown;             // [8][MG_OWN]: this CTA's columns of the residual stream
  int* s_page;               // [8]: KV page of the current position per row
  uint32_t full0, empty0;    // smem addresses of the barrier arrays
  int n_stages;
  int stage, parity;         // ring cursor of this role
  int tile_seq;              // tiles handed to the epilogue warp since the last drain (partial buffer = tile_seq & 1); both roles count
  bool prof;                 // debug profile on
  bool trace;                // this step is being traced
  bool fine;                 // stage-level trace of one fc1 phase (MegaParams::dbg_fine == 1)
  bool afine;                // step-level trace of one attention phase (dbg_fine == 2, layer 1)
  bool lfine;                // trace of one LayerNorm staging (dbg_fine == 4: LN2 of layer 1)
  bool tfine;                // trace of the step tail (dbg_fine == 3): final LN + heads, sampling, state machine, next embedding
  int trace_n;
  unsigned long long* trace_out;
  long long t_full;          // clock64 ticks thread 0 spent waiting on full barriers (prof)
  long long t_fc1, t_fc1_full, t_fc1_sync;   // fc1 phases only: total / waiting on full barriers / tile reduce + epilogue
};

__device__ __forceinline__ void mg_advance(MgCtx& cx) {
  if (++cx.stage == cx.n_stages) { cx.stage = 0; cx.parity ^= 1; }
}

// Balanced contiguous split of `n_units` over the grid.
__device__ __forceinline__ void mg_range(int n_units, int& lo, int& hi) {
  lo = (int)(((long long)n_units * blockIdx.x) / gridDim.x);
  hi = (int)(((long long)n_units * (blockIdx.x + 1)) / gridDim.x);
}

// The attention work item of this CTA for context length pos+1: (row, kv head, split) and its token range.
struct MgAttItem {
  bool mine;
  int row, kvh, split, t0, t1, chunk, nact;
};
__device__ __forceinline__ MgAttItem mg_att_item(const MegaParams& p, int pos) {
  MgAttItem it;
  const int nsplit = p.nsplit;
  it.split = blockIdx.x % nsplit;
  const int pair = blockIdx.x / nsplit;
  it.kvh = pair % p.Hkv;
  it.row = pair / p.Hkv;
  const int ctx = pos + 1;
  int chunk = (ctx + nsplit - 1) / nsplit;
  chunk = (chunk + 7) & ~7;
  if (chunk < p.min_chunk) chunk = p.min_chunk;        // tokens per split at least
  // short contexts use ONE split and skip the partial merge: a split costs ~2.5 us of ticket + merge latency, which
  // 128 tokens of K/V per CTA (one 16-token group per warp, ~1.2 us) repay only from ~3 groups per warp on
  if (ctx <= p.single_ctx) chunk = (ctx + 7) & ~7;
  it.chunk = chunk;
  it.nact = (ctx + chunk - 1) / chunk;                 // splits that own tokens (<= nsplit)
  it.mine = (int)blockIdx.x < p.R * p.Hkv * nsplit && it.split < it.nact;
  it.t0 = it.split * chunk;
  it.t1 = min(ctx, it.t0 + chunk);
  return it;
}

// =================================================================================================================
// producer side: the activation-independent schedule of everything this CTA reads from HBM, as an iterator that
// yields one ring stage per call (cp.async.bulk into the ring, gated by the empty barriers).  The consumers enumerate
// the same order with plain nested loops (mg_gemv_phase / mg_stage_ln / mg_attention / the RoPE stage).  (An L2-prefetch
// cursor running ahead of the load cursor was measured at 0/4/10/20 stages and made no difference: removed.)
// =================================================================================================================

enum { MG_PH_ROPE = 0, MG_PH_LN1, MG_PH_QKV, MG_PH_KV, MG_PH_OUT, MG_PH_LN2, MG_PH_FC1, MG_PH_FC2, MG_PH_LNF, MG_PH_HEADS, MG_PH_END };

struct MgDesc {              // per-lane view of one stage
  const void* src;           // nullptr: this lane issues nothing
  uint32_t bytes;
  uint32_t dst_off;          // offset inside the stage
  uint32_t total;            // bytes of the whole stage (warp-uniform)
  uint32_t tag;              // tile phases: phase | layer << 8 | step << 16 (trace only)
};

struct MgSched {
  int step, l, ph;
  int n0, nt, kc0, ts;       // cursors inside the current phase
  bool fresh;                // phase cursors not yet initialised
  int lo, hi;                // row range of the current tile phase
  int pos0, rope0;           // KV position at launch; (lane < R) RoPE position of row `lane` at launch
  int my_page;               // lane i: page-table entry i of this CTA's attention row (max_pages <= 32), else -1: the
                             // producer never loads the page table from memory inside the loop
  bool with_heads, final_ln;
};

__device__ __forceinline__ void mg_sched_init(MgSched& s, const MegaParams& p, int pos0, int rope0, bool with_heads, bool final_ln) {
  s.step = 0; s.l = 0; s.ph = MG_PH_ROPE; s.fresh = true;
  s.n0 = s.nt = s.kc0 = s.ts = 0; s.lo = s.hi = 0;
  s.pos0 = pos0; s.rope0 = rope0; s.with_heads = with_heads; s.final_ln = final_ln;
  const int lane = threadIdx.x & 31;
  const int att_row = ((int)blockIdx.x / p.nsplit) / p.Hkv;
  s.my_page = (p.max_pages <= 32 && att_row < p.R && lane < p.max_pages) ? p.page_table[(size_t)att_row * p.max_pages + lane] : -1;
}

__device__ __forceinline__ const bf16* mg_lw(const bf16* base0, long long stride, int l) {
  return reinterpret_cast<const bf16*>(reinterpret_cast<const char*>(base0) + (long long)l * stride);
}

// tile-phase parameters
__device__ __forceinline__ void mg_phase_weights(const MegaParams& p, int l, int ph, const bf16*& W, int& K, int& n_units, int& gate_off,
                                                 int& NT, int& UNIT) {
  gate_off = 0; NT = 1; UNIT = 2;
  switch (ph) {
    case MG_PH_QKV: W = mg_lw(p.l0.w_in, p.layer_stride, l); K = p.d; n_units = (p.q_size + 2 * p.kv_size) / 2; break;
    case MG_PH_OUT: W = mg_lw(p.l0.w_out, p.layer_stride, l); K = p.q_size; n_units = p.d / 2; break;
    case MG_PH_FC1: W = mg_lw(p.l0.w_fc1, p.layer_stride, l); K = p.d; n_units = p.F / 2; gate_off = p.F; NT = 2; break;
    case MG_PH_FC2: W = mg_lw(p.l0.w_fc2, p.layer_stride, l); K = p.F; n_units = p.d / 2; break;
    default: W = p.heads; K = p.d; n_units = p.n_q * p.V / 2; break;      // MG_PH_HEADS
  }
}

// Advance to the next stage of the schedule.  Warp-uniform control flow; returns false at the end of the launch.
__device__ __forceinline__ bool mg_sched_next(MgSched& s, const MegaParams& p, MgDesc& d) {
  const int lane = threadIdx.x & 31;
  d.src = nullptr; d.bytes = 0; d.dst_off = 0; d.total = 0; d.tag = 0;
  for (;;) {
    if (s.step >= p.n_steps) return false;
    switch (s.ph) {
      case MG_PH_ROPE: {
        if (lane < p.R) {
          d.src = p.rope + (size_t)(s.rope0 + s.step) * (ZB_DH / 2) * 2;
          d.bytes = MG_ROPE_BYTES; d.dst_off = lane * MG_ROW_BYTES;
        }
        d.total = p.R * MG_ROPE_BYTES;
        s.ph = MG_PH_LN1; s.l = 0;
        return true;
      }
      case MG_PH_LN1: case MG_PH_LN2: case MG_PH_LNF: {
        const bf16* w = s.ph == MG_PH_LN1 ? mg_lw(p.l0.ln1_w, p.layer_stride, s.l) : s.ph == MG_PH_LN2 ? mg_lw(p.l0.ln2_w, p.layer_stride, s.l) : p.normf_w;
        const bf16* b = s.ph == MG_PH_LN1 ? mg_lw(p.l0.ln1_b, p.layer_stride, s.l) : s.ph == MG_PH_LN2 ? mg_lw(p.l0.ln2_b, p.layer_stride, s.l) : p.normf_b;
        if (s.ph == MG_PH_LNF && !(s.with_heads || s.final_ln)) { s.ph = MG_PH_END; break; }
        if (lane < 2) { d.src = lane == 0 ? w : b; d.bytes = p.d * 2; d.dst_off = lane * MG_ROW_BYTES; }
        d.total = 2 * p.d * 2;
        s.ph = s.ph == MG_PH_LNF ? (s.with_heads ? MG_PH_HEADS : MG_PH_END) : s.ph + 1;
        s.fresh = true;
        return true;
      }
      case MG_PH_KV: {
        const int pos = s.pos0 + s.step;
        const MgAttItem it = mg_att_item(p, pos);
        const int tp = min(it.t1, pos);
        if (s.fresh) { s.ts = it.t0; s.fresh = false; }
        if (!it.mine || s.ts >= tp) { s.ph = MG_PH_OUT; s.fresh = true; break; }
        const int ts = s.ts;
        const int n = min(MG_KV_TOK, tp - ts);
        const int n1 = min(n, ZB_PAGE - ts % ZB_PAGE);         // tokens before the page boundary
        {
          const int run = (lane >> 1) & 1, isv = lane & 1;
          const int tb = run == 0 ? ts : ts + n1;
          const int nn = run == 0 ? n1 : n - n1;
          int page = __shfl_sync(0xffffffffu, s.my_page, (tb / ZB_PAGE) & 31);
          if (lane < 4 && nn > 0) {
            if (p.max_pages > 32) page = p.page_table[(size_t)it.row * p.max_pages + tb / ZB_PAGE];
            ZB_ASSERT(tb / ZB_PAGE < p.max_pages && page >= 0 && (size_t)(page + 1) * 2 * p.Hkv * ZB_PAGE * ZB_DH <= (size_t)p.kv_layer_elems);
            ZB_ASSERT(tb % ZB_PAGE + nn <= ZB_PAGE && (tb - ts) + nn <= MG_KV_TOK);
            const size_t head_stride = (size_t)ZB_PAGE * ZB_DH;
            const size_t kv_stride = (size_t)p.Hkv * head_stride;
            d.src = p.kv + (size_t)s.l * p.kv_layer_elems + (size_t)page * 2 * kv_stride + (size_t)it.kvh * head_stride +
                    (size_t)(tb % ZB_PAGE) * ZB_DH + (isv ? kv_stride : 0);
            d.bytes = nn * ZB_DH * 2;
            d.dst_off = isv * (MG_KV_TOK * ZB_DH * 2) + (tb - ts) * ZB_DH * 2;
          }
        }
        d.total = n * ZB_DH * 2 * 2;
        s.ts += MG_KV_TOK;
        return true;
      }
      case MG_PH_QKV: case MG_PH_OUT: case MG_PH_FC1: case MG_PH_FC2: case MG_PH_HEADS: {
        const bf16* W; int K, n_units, gate_off, NT, UNIT;
        mg_phase_weights(p, s.l, s.ph, W, K, n_units, gate_off, NT, UNIT);
        if (s.fresh) {
          mg_range(n_units, s.lo, s.hi);
          s.lo *= UNIT; s.hi *= UNIT;
          s.n0 = s.lo; s.nt = 0; s.kc0 = 0; s.fresh = false;
        }
        if (s.n0 >= s.hi) {
          if (s.ph == MG_PH_FC2) {
            if (++s.l < p.n_layer) s.ph = MG_PH_LN1; else { s.l = p.n_layer - 1; s.ph = MG_PH_LNF; }
          } else if (s.ph == MG_PH_HEADS) s.ph = MG_PH_END;
          else s.ph = s.ph + 1;
          s.fresh = true;
          break;
        }
        const int cnt = min(8, s.hi - s.n0);
        const int kcs = min(MG_KC, K - s.kc0);
        if (lane < cnt) {
          d.src = W + (size_t)(s.n0 + s.nt * gate_off + lane) * K + s.kc0;
          d.bytes = kcs * 2; d.dst_off = lane * MG_ROW_BYTES;
        }
        d.total = cnt * kcs * 2;
        d.tag = (uint32_t)s.ph | ((uint32_t)s.l << 8) | ((uint32_t)s.step << 16);
        s.kc0 += MG_KC;
        if (s.kc0 >= K) { s.kc0 = 0; if (++s.nt == NT) { s.nt = 0; s.n0 += 8; } }
        return true;
      }
      default: {   // MG_PH_END
        ++s.step; s.l = 0; s.ph = MG_PH_ROPE; s.fresh = true;
        break;
      }
    }
  }
}

// =================================================================================================================
// consumers: ring access
// =================================================================================================================
__device__ __forceinline__ const unsigned char* mg_cons_wait(MgCtx& cx) {
  // one polling lane per warp (+ __syncwarp): 256 threads spinning on try_wait contend with the producer's own
  // barrier traffic and the TMA's complete_tx updates (scripts/ring_bench.cu: idle all-lane pollers cost 30 %)
  const long long c0 = (cx.prof && threadIdx.x == 0) ? clock64() : 0;
  if ((threadIdx.x & 31) == 0) mg_mbar_wait(cx.full0 + 8u * cx.stage, cx.parity);
  __syncwarp();
  if (cx.prof && threadIdx.x == 0) cx.t_full += clock64() - c0;
  return cx.ring + (size_t)cx.stage * MG_STAGE_BYTES;
}
__device__ __forceinline__ uint32_t mg_cons_wait_addr(MgCtx& cx) {        // same, as a shared-space address
  mg_cons_wait(cx);
  return cx.ring_addr + (uint32_t)cx.stage * MG_STAGE_BYTES;
}
__device__ __forceinline__ void mg_cons_release(MgCtx& cx) {
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mg_mbar_arrive(cx.empty0 + 8u * cx.stage);
  mg_advance(cx);
}

// =================================================================================================================
// consumers: activation staging
// =================================================================================================================
// xs[r][0..K) <- bf16 row r of src (plain copy), rows r < R
__device__ __forceinline__ void mg_stage_copy(bf16* xs, int ld, const bf16* src, int K, int R) {
  const int tid = threadIdx.x;
  for (int r = 0; r < R; ++r)
    for (int e = tid * 8; e < K; e += MG_CONSUMERS * 8)
      *reinterpret_cast<uint4*>(xs + (size_t)r * ld + e) = mg_ldcg16(src + (size_t)r * K + e);
  SyncConsumers::sync();
}

// the same from an epoch-tagged buffer [R][K/2] words: the poll IS the cross-CTA dependency (no grid barrier before it)
__device__ __forceinline__ void mg_stage_ll(bf16* xs, int ld, const unsigned long long* ll, int K, int R, unsigned epoch) {
  const int tid = threadIdx.x;
  const int e0 = tid * 8;
  if (K <= MG_CONSUMERS * 8) {                                      // y: one column group, two rows per round trip
    if (e0 < K) {
#pragma unroll 1
      for (int r0 = 0; r0 < R; r0 += 2) {
        const bool r2 = r0 + 1 < R;
        const unsigned long long* w00 = ll + (size_t)r0 * (K >> 1) + (e0 >> 1);
        const unsigned long long* const wp[2] = {w00, r2 ? w00 + (K >> 1) : w00};
        uint4 v[2];
        mg_ll_load<2>(wp, epoch, v);
        *reinterpret_cast<uint4*>(xs + (size_t)r0 * ld + e0) = v[0];
        if (r2) *reinterpret_cast<uint4*>(xs + (size_t)(r0 + 1) * ld + e0) = v[1];
      }
    }
  } else {                                                          // h: K % 4096 == 0, two column groups x two rows per
#pragma unroll 1                                                    //    round trip (8 groups in flight spill: 168-register cap)
    for (int e = e0; e < K; e += MG_CONSUMERS * 8 * 2) {
#pragma unroll 1
      for (int r0 = 0; r0 < R; r0 += 2) {
        const bool r2 = r0 + 1 < R;
        const unsigned long long* w00 = ll + (size_t)r0 * (K >> 1) + (e >> 1);
        const unsigned long long* w01 = w00 + MG_CONSUMERS * 4;
        const unsigned long long* const wp[4] = {w00, w01, r2 ? w00 + (K >> 1) : w00, r2 ? w01 + (K >> 1) : w01};
        uint4 v[4];
        mg_ll_load<4>(wp, epoch, v);
        *reinterpret_cast<uint4*>(xs + (size_t)r0 * ld + e) = v[0];
        *reinterpret_cast<uint4*>(xs + (size_t)r0 * ld + e + MG_CONSUMERS * 8) = v[1];
        if (r2) {
          *reinterpret_cast<uint4*>(xs + (size_t)(r0 + 1) * ld + e) = v[2];
          *reinterpret_cast<uint4*>(xs + (size_t)(r0 + 1) * ld + e + MG_CONSUMERS * 8) = v[3];
        }
      }
    }
  }
  SyncConsumers::sync();
}

__device__ __forceinline__ void mg_unpack8(const uint4& raw, float (&v)[8]) {
  v[0] = bf16lo(raw.x); v[1] = bf16hi(raw.x); v[2] = bf16lo(raw.y); v[3] = bf16hi(raw.y);
  v[4] = bf16lo(raw.z); v[5] = bf16hi(raw.z); v[6] = bf16lo(raw.w); v[7] = bf16hi(raw.w);
}

// LayerNorm (F.layer_norm: biased variance, eps inside the sqrt, fp32 statistics, bf16 result) of rows r < R of src
// into xs.  The affine parameters arrive through the ring (stage rows 0 / 1).  K <= 2048 (one uint4 per thread per row).
// ONE statistics pass: every thread accumulates sum and sum of squares of its 8 values straight from the registers the
// loads land in, one warp reduction round, ONE CTA barrier, then every thread folds the 8 warp partials itself (warp
// order: reproducible) and normalises.  The LayerNorm trace (--opt trace=7) of the two-pass version showed 1.7 us of
// latency-bound scalar work (three barrier-separated passes over 2 x 2048 values) after the 0.9 us the data takes to
// arrive.  var = E[x^2] - mean^2 in fp32: with K = 2048 its rounding error is ~1e-5 relative, 200x below the bf16
// rounding of the result.
__device__ __forceinline__ void mg_ln_row_stats(const uint4& raw, int r, int lane, int warp, float* red1, float* red2) {
  float v[8];
  mg_unpack8(raw, v);
  float a = 0.f, q = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) { a += v[j]; q = fmaf(v[j], v[j], q); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); q += __shfl_xor_sync(0xffffffffu, q, o); }
  if (lane == 0) { red1[warp * 8 + r] = a; red2[warp * 8 + r] = q; }
}

__device__ __forceinline__ void mg_stage_ln(MgCtx& cx, bf16* xs, int ld, const bf16* src, int K, int R, float eps,
                                            const unsigned long long* ll = nullptr, unsigned ll_epoch = 0) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int e = tid * 8;
  const bool active = e < K;                        // warp-uniform (K % 256 == 0)
  float* red1 = cx.red;             // [8 warps][8 rows] partial sums
  float* red2 = cx.red + 64;        // [8 warps][8 rows] partial sums of squares (NOT the GEMV partial buffer: with use_ll >= 3
                                    // the epilogue warp may still be reducing the previous phase's last tiles)
  const float fk = (float)K;
  const uint4 z4 = make_uint4(0, 0, 0, 0);
  MG_LSTAMP();                                                      // L0: entered
  // raw rows -> xs and the statistics partials, two rows per round trip
#pragma unroll 1
  for (int r0 = 0; r0 < R; r0 += 2) {
    const bool two = r0 + 1 < R;
    uint4 raw0 = z4, raw1 = z4;
    if (active) {
      if (ll != nullptr) {
        // epoch-tagged source (x_ll): every 8-byte word is (bf16 pair, epoch).  Poll the words themselves until all of
        // them carry this phase's epoch - this replaces "grid barrier, then load" (two dependent L2 round trips plus the
        // barrier's own fence + atomic) by the load alone.  A thread only spins on its own 4 words per row.
        const unsigned long long* wp0 = ll + (size_t)r0 * (K >> 1) + (e >> 1);
        const unsigned long long* wp1 = two ? wp0 + (K >> 1) : wp0;
        unsigned long long w[8];
        for (;;) {
          asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[0]), "=l"(w[1]) : "l"(wp0) : "memory");
          asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[2]), "=l"(w[3]) : "l"(wp0 + 2) : "memory");
          asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[4]), "=l"(w[5]) : "l"(wp1) : "memory");
          asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[6]), "=l"(w[7]) : "l"(wp1 + 2) : "memory");
          bool ok = true;
#pragma unroll
          for (int i = 0; i < 8; ++i) ok = ok && (unsigned)(w[i] >> 32) == ll_epoch;
          if (ok) break;
        }
        raw0 = make_uint4((unsigned)w[0], (unsigned)w[1], (unsigned)w[2], (unsigned)w[3]);
        raw1 = make_uint4((unsigned)w[4], (unsigned)w[5], (unsigned)w[6], (unsigned)w[7]);
      } else {
        raw0 = mg_ldcg16(src + (size_t)r0 * K + e);
        if (two) raw1 = mg_ldcg16(src + (size_t)(r0 + 1) * K + e);
      }
      *reinterpret_cast<uint4*>(xs + (size_t)r0 * ld + e) = raw0;
      if (two) *reinterpret_cast<uint4*>(xs + (size_t)(r0 + 1) * ld + e) = raw1;
    }
    mg_ln_row_stats(raw0, r0, lane, warp, red1, red2);
    if (two) mg_ln_row_stats(raw1, r0 + 1, lane, warp, red1, red2);
  }
  MG_LSTAMP();                                                      // L1: thread 0's own words arrived, partials written
  // affine parameters from the ring
  const uint32_t sp = mg_cons_wait_addr(cx);
  uint4 wr = z4, br = z4;
  if (active) {
    wr = mg_lds128(sp + e * 2);
    br = mg_lds128(sp + MG_ROW_BYTES + e * 2);
  }
  mg_cons_release(cx);
  MG_LSTAMP();                                                      // L2: affine parameters read
  SyncConsumers::sync();
  MG_LSTAMP();                                                      // L3: every thread's words arrived (CTA barrier)
  float g[8], bb[8];
  mg_unpack8(wr, g);
  mg_unpack8(br, bb);
#pragma unroll 1
  for (int r = 0; r < R; ++r) {
    float m = 0.f, q = 0.f;
#pragma unroll
    for (int w = 0; w < MG_WARPS; ++w) { m += red1[w * 8 + r]; q += red2[w * 8 + r]; }
    const float mean = m / fk;
    const float rstd = rsqrtf(fmaxf(q / fk - mean * mean, 0.f) + eps);
    if (active) {
      float v[8];
      mg_unpack8(*reinterpret_cast<const uint4*>(xs + (size_t)r * ld + e), v);
      __nv_bfloat162 o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float a0 = (v[2 * j] - mean) * rstd * g[2 * j] + bb[2 * j];
        const float a1 = (v[2 * j + 1] - mean) * rstd * g[2 * j + 1] + bb[2 * j + 1];
        o[j] = __floats2bfloat162_rn(a0, a1);
      }
      *reinterpret_cast<uint4*>(xs + (size_t)r * ld + e) = *reinterpret_cast<uint4*>(o);
    }
  }
  SyncConsumers::sync();
  MG_LSTAMP();                                                      // L5: normalised rows staged
}

// =================================================================================================================
// consumers: epilogues (same arithmetic as gm_epilogue_* in gemv_mma.cuh).  Store-only: nothing here waits on memory.
// =================================================================================================================
struct MgEpi {
  bf16* out;             // RESID: x; SWIGLU: h; QKV: q
  int out_stride;
  bf16* kv_layer;
  int pos;
  int own_lo;            // RESID: first residual column owned by this CTA
  float cfg_scale;
  int eos_bias;
  float* logits;
  int layer;
  unsigned long long* ll;  // RESID: x_ll buffer of this phase
  unsigned ll_epoch;
};

__device__ __forceinline__ void mg_epi_resid(const MgCtx& cx, const MgEpi& e, int a, int n0, float s0, float s1) {
  // x <- bf16(x + bf16(W a))   (_torch.py:100-101, bf16 residual stream); the old value lives in shared memory
  float* xo = cx.s_xown + a * MG_OWN + (n0 - e.own_lo);
  const __nv_bfloat162 o = __floats2bfloat162_rn(xo[0] + bf16_round(s0), xo[1] + bf16_round(s1));
  *reinterpret_cast<__nv_bfloat162*>(e.out + (size_t)a * e.out_stride + n0) = o;
  {
    // the same pair again, tagged with this phase's epoch, for the LayerNorm staging of the next phase: its readers
    // poll the data itself instead of meeting at a grid barrier first (8-byte stores are single-copy atomic)
    const unsigned long long w = ((unsigned long long)e.ll_epoch << 32) | (unsigned long long)(*reinterpret_cast<const unsigned*>(&o));
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(e.ll + (size_t)a * (e.out_stride >> 1) + (n0 >> 1)), "l"(w) : "memory");
  }
  const float2 f = __bfloat1622float2(o);
  xo[0] = f.x; xo[1] = f.y;
}

__device__ __forceinline__ void mg_epi_qkv(const MegaParams& p, const MgCtx& cx, const MgEpi& e, int a, int n0, float s0, float s1) {
  const float v0 = bf16_round(s0), v1 = bf16_round(s1);
  const int pos = e.pos;
  __nv_bfloat162 ob;
  if (n0 < p.q_size + p.kv_size) {
    const int i = (n0 % ZB_DH) >> 1;
    const float2 cs = *reinterpret_cast<const float2*>(cx.s_rope + a * (ZB_DH) + i * 2);
    const float o0 = __fsub_rn(__fmul_rn(v0, cs.x), __fmul_rn(v1, cs.y));
    const float o1 = __fadd_rn(__fmul_rn(v1, cs.x), __fmul_rn(v0, cs.y));
    const __nv_bfloat162 o = __floats2bfloat162_rn(o0, o1);
    ob = o;
    if (n0 < p.q_size) {
      if (e.ll == nullptr) *reinterpret_cast<__nv_bfloat162*>(e.out + (size_t)a * e.out_stride + n0) = o;
    } else {
      const int m = n0 - p.q_size;
      const int page = cx.s_page[a];
      bf16* dst = e.kv_layer + ((((size_t)page * 2 + 0) * p.Hkv + m / ZB_DH) * ZB_PAGE + pos % ZB_PAGE) * ZB_DH + m % ZB_DH;
      *reinterpret_cast<__nv_bfloat162*>(dst) = o;
    }
  } else {
    const int m = n0 - p.q_size - p.kv_size;
    const int page = cx.s_page[a];
    bf16* dst = e.kv_layer + ((((size_t)page * 2 + 1) * p.Hkv + m / ZB_DH) * ZB_PAGE + pos % ZB_PAGE) * ZB_DH + m % ZB_DH;
    ob = __floats2bfloat162_rn(v0, v1);
    *reinterpret_cast<__nv_bfloat162*>(dst) = ob;
  }
  // use_ll >= 2: the attention of THIS step polls q and the fresh k / v row from here (the cache rows above are for later steps)
  if (e.ll != nullptr) mg_ll_store(e.ll + (size_t)a * ((p.q_size + 2 * p.kv_size) >> 1) + (n0 >> 1), e.ll_epoch, *reinterpret_cast<const unsigned*>(&ob));
}

__device__ __forceinline__ void mg_epi_heads(const MegaParams& p, const MgEpi& e, int b, int n0, float c0, float c1, float u0, float u1) {
  const int k = n0 / p.V, n = n0 % p.V;
  float out[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const float c = bf16_round(j == 0 ? c0 : c1);
    const float u = bf16_round(j == 0 ? u0 : u1);
    float l = __fadd_rn(u, __fmul_rn(__fsub_rn(c, u), e.cfg_scale));
    const int nn = n + j;
    if (nn >= 1025) l = -INFINITY;
    if (e.eos_bias && k >= 1 && nn == p.eos) l = -INFINITY;
    out[j] = l;
  }
  *reinterpret_cast<float2*>(e.logits + ((size_t)b * p.n_q + k) * p.V + n) = make_float2(out[0], out[1]);
}

// =================================================================================================================
// consumers: one GEMV phase over this CTA's row range.  xs must be staged (and synced) before the call.  ONE copy of
// this code serves all five weight matrices (runtime epilogue switch): code size matters more than a few branches.
// =================================================================================================================
// PK (compile time): packed-K MMA layout — at most 4 activation rows and every contraction length a multiple of 512 (the
// host checks both); only ONE of the two MMA loop variants exists in a given instantiation (code size: see k_decode_mega).
template <bool PK>
__device__ __forceinline__ void mg_gemv_phase(const MegaParams& p, MgCtx& cx, const bf16* xs, int ld, int K, int n_units, int gate_off,
                                           int kind, int layer, bool drain) {
  const int NT = kind == EPI_SWIGLU ? 2 : 1;
  const int UNIT = 2;                            // every phase deals rows in pairs (bf16 pairs are the store / tag granularity)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int R = p.R;
  int lo, hi;
  mg_range(n_units, lo, hi);
  lo *= UNIT; hi *= UNIT;
  const uint32_t xs_addr = mg_smem_u32(xs);
  const bool pf1 = cx.prof && kind == EPI_SWIGLU && threadIdx.x == 0;
  const long long pt0 = pf1 ? clock64() : 0, pfull0 = cx.t_full;
  for (int n0 = lo; n0 < hi; n0 += 8) {
    const int cnt = min(8, hi - n0);
    // FOUR independent accumulator sets per tile: HMMA results chain through the accumulator, and one chain of 16
    // dependent HMMAs per stage left the tensor pipe idle most of the time (r01 trace: the consumers, not the HBM
    // stream, were the bottleneck).  The sets are summed once per tile.
    float acc[2][4][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int a = 0; a < 4; ++a) { acc[nt][a][0] = acc[nt][a][1] = acc[nt][a][2] = acc[nt][a][3] = 0.f; }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      if (nt < NT) {
        for (int kc0 = 0; kc0 < K; kc0 += MG_KC) {
          const int kcs = min(MG_KC, K - kc0);
          const uint32_t st_addr = mg_cons_wait_addr(cx);
          const bool fstamp = cx.fine && kind == p.dbg_kind && layer == 1 && threadIdx.x == 0 && cx.trace_n < 32;
          if (fstamp) cx.trace_out[cx.trace_n++] = mg_globaltimer();        // stage data seen by warp 0
          if (PK) {
            // R <= 4: the WEIGHTS are the MMA's A operand.  A rows 0..7 = the 8 weight rows over the first half of the
            // chunk's columns, A rows 8..15 = the same weight rows over the second half; B columns 0..3 = the
            // activation rows over the first half, columns 4..7 = over the second half.  D[i<8][n<4] and
            // D[i>=8][n>=4] are the two K-half partial sums, the cross blocks are ignored.  One HMMA now covers
            // 8 rows x 32 columns instead of 8 x 16: the tensor pipe (1 HMMA / 16 clk / sub-partition on sm_100,
            // scripts/consume_bench.cu) was the bottleneck of the fc1/fc2 phases with the activations as A.
            const int hs = kcs >> 1;                     // columns per half
            const int slice = hs >> 3;                   // this warp's columns in each half (multiple of 32)
            const int nkb = slice >> 5;
            const uint32_t wa0 = st_addr + g * MG_ROW_BYTES + (warp * slice + t * 8) * 2;
            const uint32_t wa1 = wa0 + hs * 2;
            const int xn = g & 3, xh = g >> 2;
            const uint32_t xa = xs_addr + ((uint32_t)xn * ld + kc0 + xh * hs + warp * slice + t * 8) * 2;
            int j = 0;
#pragma unroll 2
            for (; j + 1 < nkb; j += 2) {
              const uint4 w0 = mg_lds128(wa0 + j * 64), w1 = mg_lds128(wa1 + j * 64);
              const uint4 w2 = mg_lds128(wa0 + j * 64 + 64), w3 = mg_lds128(wa1 + j * 64 + 64);
              uint4 x0 = make_uint4(0, 0, 0, 0), x1 = make_uint4(0, 0, 0, 0);
              if (xn < R) { x0 = mg_lds128(xa + j * 64); x1 = mg_lds128(xa + j * 64 + 64); }
              mma_bf16_16816_full(acc[nt][0], w0.x, w1.x, w0.y, w1.y, x0.x, x0.y);
              mma_bf16_16816_full(acc[nt][1], w0.z, w1.z, w0.w, w1.w, x0.z, x0.w);
              mma_bf16_16816_full(acc[nt][2], w2.x, w3.x, w2.y, w3.y, x1.x, x1.y);
              mma_bf16_16816_full(acc[nt][3], w2.z, w3.z, w2.w, w3.w, x1.z, x1.w);
            }
            if (j < nkb) {
              const uint4 w0 = mg_lds128(wa0 + j * 64), w1 = mg_lds128(wa1 + j * 64);
              uint4 x0 = make_uint4(0, 0, 0, 0);
              if (xn < R) x0 = mg_lds128(xa + j * 64);
              mma_bf16_16816_full(acc[nt][0], w0.x, w1.x, w0.y, w1.y, x0.x, x0.y);
              mma_bf16_16816_full(acc[nt][1], w0.z, w1.z, w0.w, w1.w, x0.z, x0.w);
            }
          } else {
            const int slice = kcs >> 3;                  // this warp's columns of the chunk (multiple of 32)
            const int nkb = slice >> 5;
            // explicit shared-space addresses: through generic pointers the compiler emitted LD.E.128 for the ring
            const uint32_t wa = st_addr + g * MG_ROW_BYTES + (warp * slice + t * 8) * 2;
            const uint32_t xa = xs_addr + ((uint32_t)g * ld + kc0 + warp * slice + t * 8) * 2;
            int j = 0;
#pragma unroll 2
            for (; j + 1 < nkb; j += 2) {
              const uint4 w0 = mg_lds128(wa + j * 64), w1 = mg_lds128(wa + j * 64 + 64);
              uint4 x0 = make_uint4(0, 0, 0, 0), x1 = make_uint4(0, 0, 0, 0);
              if (g < R) { x0 = mg_lds128(xa + j * 64); x1 = mg_lds128(xa + j * 64 + 64); }
              mma_bf16_16816(acc[nt][0], x0.x, x0.y, w0.x, w0.y);
              mma_bf16_16816(acc[nt][1], x0.z, x0.w, w0.z, w0.w);
              mma_bf16_16816(acc[nt][2], x1.x, x1.y, w1.x, w1.y);
              mma_bf16_16816(acc[nt][3], x1.z, x1.w, w1.z, w1.w);
            }
            if (j < nkb) {
              const uint4 w0 = mg_lds128(wa + j * 64);
              uint4 x0 = make_uint4(0, 0, 0, 0);
              if (g < R) x0 = mg_lds128(xa + j * 64);
              mma_bf16_16816(acc[nt][0], x0.x, x0.y, w0.x, w0.y);
              mma_bf16_16816(acc[nt][1], x0.z, x0.w, w0.z, w0.w);
            }
          }
          mg_cons_release(cx);
          if (fstamp && cx.trace_n < 32) cx.trace_out[cx.trace_n++] = mg_globaltimer();      // warp 0 done with the stage
        }
      }
    }
    // ---- hand the tile to the epilogue warp.  The stage-level trace of the version in which the MMA warps did the
    // cross-warp reduce + epilogue themselves (profiles/r01_fc1_stage_trace.txt) showed a 0.6 us tail per 8-row tile
    // (partial store, CTA barrier, 8-warp reduce, SwiGLU on warp 0) against 2 x 0.35 us of MMA work: the GEMV phases were
    // bound by that serial tail, not by HBM.  Here the MMA warps only store their partials and arrive on a named
    // barrier; warp 9 reduces and runs the epilogue while they are already in the next tile.
    const int f = cx.tile_seq & 1;
    float* pb = cx.part + (size_t)f * 2 * MG_WARPS * 64;
    if (cx.tile_seq >= 2) mg_bar_sync(MG_BAR_FREE + f, MG_HANDOFF);        // the epilogue warp is done with tile tile_seq - 2
    ++cx.tile_seq;
    if (cx.fine && kind == p.dbg_kind && layer == 1 && threadIdx.x == 0 && cx.trace_n < 32) cx.trace_out[cx.trace_n++] = mg_globaltimer();   // partial buffer free
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      if (nt < NT) {
        float s4[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) s4[i] = (acc[nt][0][i] + acc[nt][1][i]) + (acc[nt][2][i] + acc[nt][3][i]);
        float* pw = pb + (nt * MG_WARPS + warp) * 64;          // [8 activation rows][8 weight rows]
        if (PK) {
          // lane (g, t<2): D[g][2t, 2t+1] (first K half) + lane (g, t+2): D[g+8][2t+4, 2t+5] (second K half)
          const float v0 = s4[0] + __shfl_xor_sync(0xffffffffu, s4[2], 2);
          const float v1 = s4[1] + __shfl_xor_sync(0xffffffffu, s4[3], 2);
          if (t < 2) { pw[(2 * t) * 8 + g] = v0; pw[(2 * t + 1) * 8 + g] = v1; }
        } else {
          *reinterpret_cast<float2*>(pw + g * 8 + 2 * t) = make_float2(s4[0], s4[1]);
        }
      }
    }
    __threadfence_block();
    mg_bar_arrive(MG_BAR_FULL + f, MG_HANDOFF);
    if (cx.fine && kind == p.dbg_kind && layer == 1 && threadIdx.x == 0 && cx.trace_n < 32) cx.trace_out[cx.trace_n++] = mg_globaltimer();   // tile handed over
  }
  // Drain: wait until the epilogue warp has issued every store of the (up to two) tiles still in flight - needed where a
  // grid barrier publishes plain stores.  Where the phase's output travels epoch-tagged (use_ll >= 3: every phase of the
  // layer stack) nobody waits: the MMA warps go straight on to the next phase's staging and the tile counter runs on.
  if (drain) {
    if (cx.tile_seq >= 2) mg_bar_sync(MG_BAR_FREE + (cx.tile_seq & 1), MG_HANDOFF);
    if (cx.tile_seq >= 1) mg_bar_sync(MG_BAR_FREE + ((cx.tile_seq - 1) & 1), MG_HANDOFF);
    cx.tile_seq = 0;
  }
  if (cx.fine && kind == p.dbg_kind && layer == 1 && threadIdx.x == 0 && cx.trace_n < 32) cx.trace_out[cx.trace_n++] = mg_globaltimer();     // phase done
  if (pf1) { cx.t_fc1 += clock64() - pt0; cx.t_fc1_full += cx.t_full - pfull0; }
}

// The epilogue warp's side of one GEMV phase: per tile, wait for the 8 MMA warps' partials, reduce them in warp order
// (fixed summation order: reproducible tokens) and run the fused epilogue; same arithmetic as gm_epilogue_* in gemv_mma.cuh.
__device__ __forceinline__ void mg_gemv_epi(const MegaParams& p, MgCtx& cx, int n_units, int kind, const MgEpi& epi, bool drain) {
  const int UNIT = 2;
  const int lane = threadIdx.x & 31;
  const int R = p.R;
  int lo, hi;
  mg_range(n_units, lo, hi);
  lo *= UNIT; hi *= UNIT;
  for (int n0 = lo; n0 < hi; n0 += 8) {
    const int cnt = min(8, hi - n0);
    const int f = cx.tile_seq & 1;
    ++cx.tile_seq;
    const float* pb = cx.part + (size_t)f * 2 * MG_WARPS * 64;
    mg_bar_sync(MG_BAR_FULL + f, MG_HANDOFF);
    if (kind == EPI_SWIGLU) {
      for (int i0 = 0; i0 < R * 8; i0 += 32) {                      // all 32 lanes run every trip (the shuffle below)
        const int i = i0 + lane;
        const int r = i >> 3, n = i & 7;                            // r < 8 always: in bounds of the partial buffer
        float yv = 0.f, gt = 0.f;
#pragma unroll
        for (int w = 0; w < MG_WARPS; ++w) { yv += pb[((0 * MG_WARPS + w) * 8 + r) * 8 + n]; gt += pb[((1 * MG_WARPS + w) * 8 + r) * 8 + n]; }
        yv = bf16_round(yv);
        gt = bf16_round(gt);
        const float sg = bf16_round(gt / (1.0f + expf(-gt)));
        const float hv = yv * sg, ho = __shfl_xor_sync(0xffffffffu, hv, 1);
        if (r < R && n < cnt && !(n & 1)) {                                     // n0 and cnt are even
          const __nv_bfloat162 o2 = __floats2bfloat162_rn(hv, ho);
          if (epi.ll != nullptr) mg_ll_store(epi.ll + (((size_t)r * epi.out_stride + n0 + n) >> 1), epi.ll_epoch, *reinterpret_cast<const unsigned*>(&o2));
          else *reinterpret_cast<__nv_bfloat162*>(epi.out + (size_t)r * epi.out_stride + n0 + n) = o2;
        }
      }
    } else if (kind == EPI_HEADS) {
      const int B = R >> 1;
      for (int i = lane; i < B * 4; i += 32) {
        const int b = i >> 2, j = i & 3;
        float c0 = 0.f, c1 = 0.f, u0 = 0.f, u1 = 0.f;
#pragma unroll
        for (int w = 0; w < MG_WARPS; ++w) {
          c0 += pb[(w * 8 + b) * 8 + 2 * j]; c1 += pb[(w * 8 + b) * 8 + 2 * j + 1];
          u0 += pb[(w * 8 + b + B) * 8 + 2 * j]; u1 += pb[(w * 8 + b + B) * 8 + 2 * j + 1];
        }
        if (2 * j < cnt) mg_epi_heads(p, epi, b, n0 + 2 * j, c0, c1, u0, u1);
      }
    } else {
      for (int i = lane; i < R * 4; i += 32) {
        const int r = i >> 2, j = i & 3;
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int w = 0; w < MG_WARPS; ++w) { s0 += pb[(w * 8 + r) * 8 + 2 * j]; s1 += pb[(w * 8 + r) * 8 + 2 * j + 1]; }
        if (2 * j < cnt) {
          if (kind == EPI_QKV) mg_epi_qkv(p, cx, epi, r, n0 + 2 * j, s0, s1);
          else mg_epi_resid(cx, epi, r, n0 + 2 * j, s0, s1);
        }
      }
    }
    __syncwarp();
    __threadfence_block();
    mg_bar_arrive(MG_BAR_FREE + f, MG_HANDOFF);   // consumed by the MMA warps before tile + 2, or by their drain
  }
  if (drain) cx.tile_seq = 0;
}

// =================================================================================================================
// consumers: split-KV GQA decode attention for this CTA's (row, kv head, split) item; same arithmetic as
// k_attn_decode.  K/V of earlier steps arrive through the ring; the row appended in this step is read from L2.
// =================================================================================================================
template <int G>
__device__ __forceinline__ void mg_att_token(const uint4& kraw, const uint4& vraw, bool valid, const float (&q)[G][8], float (&m)[G], float (&l)[G],
                                             float (&acc)[G][8], float scale_log2) {
  float kf[8], vf[8];
  mg_unpack8(kraw, kf);
  mg_unpack8(vraw, vf);
#pragma unroll
  for (int g = 0; g < G; ++g) {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s = fmaf(q[g][j], kf[j], s);
    s += __shfl_xor_sync(0xffffffffu, s, 8);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    if (valid) {
      s *= scale_log2;
      const float mn = fmaxf(m[g], s);
      const float corr = exp2f(m[g] - mn);       // m = -inf on the first token -> 0
      const float pr = exp2f(s - mn);
      l[g] = l[g] * corr + pr;
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[g][j] = fmaf(pr, vf[j], acc[g][j] * corr);
      m[g] = mn;
    }
  }
}

template <int G>
__device__ __forceinline__ void mg_attention(const MegaParams& p, MgCtx& cx, const bf16* kv_layer, int pos, unsigned epoch, bool tagged) {
  const MgAttItem it = mg_att_item(p, pos);
  if (!it.mine) return;
  MG_ASTAMP();                                                      // a0: phase entered                               // CTA-uniform
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;                // mma.sync fragment coordinates (attention_kernels.cuh, att_group16)
  const int nsplit = p.nsplit, row = it.row, kvh = it.kvh, split = it.split;

  // q and (warp 0) the K/V row appended in this step come from L2 in ONE round trip: both were written by other CTAs
  // before the grid barrier.  Every later read of this phase is shared memory.  Both are loaded straight into the
  // fragment layout of the tensor-core contraction.
  const int tp = min(it.t1, pos);
  const bool has_new = it.t0 <= pos && pos < it.t1;
  const uint4 z4 = make_uint4(0, 0, 0, 0);
  uint32_t qa[8][2];
  {
    uint4 qw[4];
    AttFrag fn;                                         // the fresh row is token 0 of a one-token group
    if (tagged) {
      // use_ll >= 2: q and the fresh k / v row are polled from the in_proj epilogues' tagged words - no grid barrier
      // between in_proj and here.  Lanes without a fragment of their own poll lane 0's words (always valid addresses).
      const unsigned long long* qrow = p.qkv_ll + (size_t)row * ((p.q_size + 2 * p.kv_size) >> 1);
      const unsigned long long* qb = qrow + (((size_t)(kvh * G + (g < G ? g : 0)) * ZB_DH + 8 * t) >> 1);
      const unsigned long long* const wq[4] = {qb, qb + 16, qb + 32, qb + 48};
      mg_ll_load<4>(wq, epoch, qw);
      if (g >= G) { qw[0] = z4; qw[1] = z4; qw[2] = z4; qw[3] = z4; }
      if (has_new && warp == 0) {
        const unsigned long long* kb = qrow + ((p.q_size + (size_t)kvh * ZB_DH + 8 * t) >> 1);
        const unsigned long long* vb = qrow + ((p.q_size + p.kv_size + (size_t)kvh * ZB_DH + 16 * g) >> 1);
        const unsigned long long* const wf[6] = {kb, kb + 16, kb + 32, kb + 48, vb, vb + 4};
        uint4 f6[6];
        mg_ll_load<6>(wf, epoch, f6);
#pragma unroll
        for (int s = 0; s < 4; ++s) { fn.k[0][s] = g == 0 ? f6[s] : z4; fn.k[1][s] = z4; }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          fn.v[0][h] = t == 0 ? f6[4 + h] : z4;
          fn.v[1][h] = z4; fn.v[2][h] = z4; fn.v[3][h] = z4;
        }
      }
    } else {
#pragma unroll
      for (int s = 0; s < 4; ++s)
        qw[s] = g < G ? mg_ldcg16(p.q + (size_t)row * p.H * ZB_DH + (size_t)(kvh * G + g) * ZB_DH + 32 * s + 8 * t) : z4;
      if (has_new && warp == 0) {
        const size_t head_stride = (size_t)ZB_PAGE * ZB_DH;
        const size_t kv_stride = (size_t)p.Hkv * head_stride;
        const bf16* kp = kv_layer + (size_t)cx.s_page[row] * 2 * kv_stride + (size_t)kvh * head_stride + (size_t)(pos % ZB_PAGE) * ZB_DH;
#pragma unroll
        for (int s = 0; s < 4; ++s) { fn.k[0][s] = g == 0 ? mg_ldcg16(kp + 32 * s + 8 * t) : z4; fn.k[1][s] = z4; }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          fn.v[0][h] = t == 0 ? mg_ldcg16(kp + kv_stride + 16 * g + 8 * h) : z4;
          fn.v[1][h] = z4; fn.v[2][h] = z4; fn.v[3][h] = z4;
        }
      }
    }
#pragma unroll
    for (int s = 0; s < 4; ++s) { qa[2 * s][0] = qw[s].x; qa[2 * s][1] = qw[s].y; qa[2 * s + 1][0] = qw[s].z; qa[2 * s + 1][1] = qw[s].w; }
    if (cx.afine && qa[0][0] == 0x12345678u) cx.trace_n = 0;        // (trace only) make the next stamp wait for the q load
    MG_ASTAMP();                                                    // a1: q arrived from L2
    float m = -INFINITY, l = 0.f, o[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) { o[j][0] = 0.f; o[j][1] = 0.f; o[j][2] = 0.f; o[j][3] = 0.f; }
    // ---- tokens of earlier steps: MG_KV_TOK = 64 per ring stage = four 16-token groups.  Warps 0-3 take the groups of
    // the even stages, warps 4-7 those of the odd ones (every warp still passes every stage's barriers).
    const int n_kv_stages = tp > it.t0 ? (tp - it.t0 + MG_KV_TOK - 1) / MG_KV_TOK : 0;
    for (int si = 0; si < n_kv_stages; ++si) {
      const int ts = it.t0 + si * MG_KV_TOK;
      const int n = min(MG_KV_TOK, tp - ts);
      const uint32_t sp = mg_cons_wait_addr(cx);
      MG_ASTAMP();                                                  // per KV stage: data seen by warp 0
      const int tl = 16 * (warp & 3);                   // first token of this warp's group inside the stage
      if ((si & 1) == (warp >> 2) && tl < n) {          // warp-uniform
        AttFrag f;
#pragma unroll
        for (int nn = 0; nn < 2; ++nn) {
          const int tok = tl + 8 * nn + g;
#pragma unroll
          for (int s = 0; s < 4; ++s) f.k[nn][s] = tok < n ? mg_lds128(sp + tok * (ZB_DH * 2) + (32 * s + 8 * t) * 2) : z4;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int tok = tl + 2 * t + (u & 1) + 8 * (u >> 1);
#pragma unroll
          for (int h = 0; h < 2; ++h)
            f.v[u][h] = tok < n ? mg_lds128(sp + MG_KV_TOK * ZB_DH * 2 + tok * (ZB_DH * 2) + (16 * g + 8 * h) * 2) : z4;
        }
        att_group16<G>(f, ts + tl, ts + n, qa, m, l, o, p.scale_log2, lane);
      }
      mg_cons_release(cx);
      MG_ASTAMP();                                                  // per KV stage: warp 0 done
    }
    if (has_new && warp == 0) att_group16<G>(fn, pos, pos + 1, qa, m, l, o, p.scale_log2, lane);
    MG_ASTAMP();                                                    // fresh row done
    // per-warp states -> shared memory
    float* s_o = reinterpret_cast<float*>(cx.scratch);              // [MG_MAX_SPLIT >= 8][G][128]
    float* s_m = s_o + MG_MAX_SPLIT * G * ZB_DH;                    // [8][G]
    float* s_l = s_m + MG_WARPS * G;                                // [8][G]
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int d = 16 * g + 2 * j;
      if (2 * t < G) { s_o[(warp * G + 2 * t) * ZB_DH + d] = o[j][0]; s_o[(warp * G + 2 * t) * ZB_DH + d + 1] = o[j][2]; }
      if (2 * t + 1 < G) { s_o[(warp * G + 2 * t + 1) * ZB_DH + d] = o[j][1]; s_o[(warp * G + 2 * t + 1) * ZB_DH + d + 1] = o[j][3]; }
    }
    if (t == 0 && g < G) { s_m[warp * G + g] = m; s_l[warp * G + g] = l; }
  }
  float* s_o = reinterpret_cast<float*>(cx.scratch);              // [8 warps][G][128]
  float* s_m = s_o + MG_MAX_SPLIT * G * ZB_DH;                    // [8][G]
  float* s_l = s_m + MG_WARPS * G;                                // [8][G]
  SyncConsumers::sync();
  const int nact = it.nact;
  // Split partials travel as epoch-tagged 8-byte words (fp32 value, epoch) like the residual stream (x_ll): the merging
  // CTA polls the DATA.  The first version stored plain partials, took a ticket with an acq_rel atomic (a round trip,
  // after a CTA barrier), and the last arriver then loaded all partials (another round trip) into shared memory and
  // merged from there: 4.3 us from "partials stored" to "y stored" in the attention trace, on the critical path of the
  // grid barrier that 100+ idle CTAs wait at.  Now: stores are fire-and-forget, the merger (the CTA of split 0) sees
  // the last partial one round trip after it was stored and merges straight from the registers the polls land in.
  const int wps = G * (ZB_DH + 2);                                // words per split: [G][128] o, then [G][2] (m, l)
  unsigned long long* pair_ll = p.att_ll + ((size_t)row * p.Hkv + kvh) * MG_MAX_SPLIT * wps;
  unsigned long long* my_ll = pair_ll + (size_t)split * wps;
  // rescale factor of every (warp, head) once, by one thread each (the exp2f chain used to sit in every output's loop)
  float* s_c = s_l + MG_WARPS * G;                                // [8][G]
  if (tid < MG_WARPS * G) {
    const int g = tid % G;
    float mn = -INFINITY;
#pragma unroll
    for (int w = 0; w < MG_WARPS; ++w) mn = fmaxf(mn, s_m[w * G + g]);
    const float ms = s_m[tid];
    s_c[tid] = (ms == -INFINITY) ? 0.f : exp2f(ms - mn);
  }
  SyncConsumers::sync();
#pragma unroll 1
  for (int oi = tid; oi < G * ZB_DH; oi += MG_CONSUMERS) {
    const int g = oi >> 7, dim = oi & (ZB_DH - 1);
    float lt = 0.f, ot = 0.f, mn = -INFINITY;
#pragma unroll
    for (int w = 0; w < MG_WARPS; ++w) {
      const float c = s_c[w * G + g];
      lt += s_l[w * G + g] * c;
      ot += s_o[(w * G + g) * ZB_DH + dim] * c;
    }
    if (dim == 0) {
#pragma unroll
      for (int w = 0; w < MG_WARPS; ++w) mn = fmaxf(mn, s_m[w * G + g]);
    }
    if (nact == 1) {                                              // the only split: this IS the attention output
      const float yv = ot / lt;
      if (tagged) {                                               // (pair, epoch) words: the even lane stores both values
        const float yo = __shfl_xor_sync(0xffffffffu, yv, 1);
        const __nv_bfloat162 o2 = __floats2bfloat162_rn(yv, yo);
        if (!(dim & 1)) mg_ll_store(p.y_ll + (((size_t)row * p.H * ZB_DH + (size_t)(kvh * G + g) * ZB_DH + dim) >> 1), epoch, *reinterpret_cast<const unsigned*>(&o2));
      } else {
        p.y[(size_t)row * p.H * ZB_DH + (size_t)(kvh * G + g) * ZB_DH + dim] = __float2bfloat16_rn(yv);
      }
    } else {
      const unsigned long long tag = (unsigned long long)epoch << 32;
      asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(my_ll + oi), "l"(tag | __float_as_uint(ot)) : "memory");
      if (dim == 0) {
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(my_ll + G * ZB_DH + 2 * g), "l"(tag | __float_as_uint(mn)) : "memory");
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(my_ll + G * ZB_DH + 2 * g + 1), "l"(tag | __float_as_uint(lt)) : "memory");
      }
    }
  }
  MG_ASTAMP();                                                      // partials (or y) stored
  if (nact == 1) return;                                          // CTA-uniform
  // ---- merge, spread over the nact CTAs of this (row, kv head): CTA `split` owns a balanced slice of the G * 64 output
  // pairs and polls exactly the words it needs - (o[2 dp], o[2 dp + 1]) of its pairs from every split, and the (m, l)
  // pairs of all (split, head) - about one 16-byte load per thread.  (One CTA pulling all 58 KB of partials through its
  // own L2 port needs > 1 us for the transfer alone, and re-polling multiplies that: measured 3.5-5 us.)
  const int npairs = G * (ZB_DH / 2);
  const int lo = split * npairs / nact, cnt = (split + 1) * npairs / nact - lo;
  const int items = cnt * nact;
  float2* s_po = reinterpret_cast<float2*>(s_o + MG_WARPS * G * ZB_DH);       // [cnt][nact]  (upper half of the s_o region)
  float2* s_pml = s_po + (MG_MAX_SPLIT * G * ZB_DH - MG_WARPS * G * ZB_DH) / 2 - MG_MAX_SPLIT * G;    // [nact][G] (m, l)
#pragma unroll 1
  for (int item = tid; item < items + nact * G; item += MG_CONSUMERS) {
    const unsigned long long* wp;
    float2* dst;
    if (item < items) {
      const int pl = item / nact, u = item - pl * nact, P = lo + pl;
      wp = pair_ll + (size_t)u * wps + (P >> 6) * ZB_DH + 2 * (P & 63);
      dst = s_po + item;
    } else {
      const int j = item - items, u = j / G, g = j - u * G;
      wp = pair_ll + (size_t)u * wps + G * ZB_DH + 2 * g;
      dst = s_pml + j;
    }
    unsigned long long w0, w1;
    do {
      asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(wp) : "memory");
    } while ((unsigned)(w0 >> 32) != epoch || (unsigned)(w1 >> 32) != epoch);
    *dst = make_float2(__uint_as_float((unsigned)w0), __uint_as_float((unsigned)w1));
  }
  SyncConsumers::sync();
  MG_ASTAMP();                                                      // every partial word of this CTA's slice arrived
  if (tid < nact * G) {                                           // (split, head): rescale factor and rescaled l, once
    const int g = tid % G;
    float mn = -INFINITY;
#pragma unroll 1
    for (int u = 0; u < nact; ++u) mn = fmaxf(mn, s_pml[u * G + g].x);
    const float2 ml = s_pml[tid];
    const float c = (ml.x == -INFINITY) ? 0.f : exp2f(ml.x - mn);
    s_c[tid] = c;
    s_c[MG_MAX_SPLIT * G + tid] = ml.y * c;
  }
  SyncConsumers::sync();
#pragma unroll 1
  for (int pl = tid; pl < cnt; pl += MG_CONSUMERS) {
    const int P = lo + pl, g = P >> 6, dp = P & 63;
    float lt = 0.f, o0 = 0.f, o1 = 0.f;
#pragma unroll 2
    for (int u = 0; u < nact; ++u) {
      const float c = s_c[u * G + g];
      const float2 o = s_po[pl * nact + u];
      lt += s_c[MG_MAX_SPLIT * G + u * G + g];
      o0 += o.x * c;
      o1 += o.y * c;
    }
    const __nv_bfloat162 o2 = __floats2bfloat162_rn(o0 / lt, o1 / lt);
    const size_t yi = (size_t)row * p.H * ZB_DH + (size_t)(kvh * G + g) * ZB_DH + 2 * dp;
    if (tagged) mg_ll_store(p.y_ll + (yi >> 1), epoch, *reinterpret_cast<const unsigned*>(&o2));
    else *reinterpret_cast<__nv_bfloat162*>(p.y + yi) = o2;
  }
  MG_ASTAMP();                                                      // this CTA's slice of y stored
}

// =================================================================================================================
// consumers: step input
// =================================================================================================================
__device__ __noinline__ void mg_embed_row(const MegaParams& p, int b, int col, int L) {
  // model.py:97-98,141-142: x[b] = x[B+b] = ((E0[c0] + E1[c1]) + ...), sequential bf16 adds.  Runs once per step and is
  // instruction-cache cold every time: the fully unrolled first version (16 predicated rows, 14 KB of SASS) took 18 us
  // for two L2 round trips.  Lane k fetches id k, the ids travel by shuffle, rows go three at a time.
  const int d = p.d, lane = threadIdx.x & 31;
  int myid = 0;
  if (lane < p.n_q) {
    long long id = __ldcg(p.delayed + ((size_t)b * p.n_q + lane) * L + col);
    myid = (int)(id < 0 ? 0 : (id >= p.V ? p.V - 1 : id));
  }
#pragma unroll 1
  for (int e = threadIdx.x * 8; e < d; e += MG_CONSUMERS * 8) {        // warp-uniform trip count (d % 256 == 0)
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll 1
    for (int k0 = 0; k0 < p.n_q; k0 += 3) {
      uint4 raw[3];
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int k = k0 + u;
        const int id = __shfl_sync(0xffffffffu, myid, k & 31);
        raw[u] = make_uint4(0, 0, 0, 0);
        ZB_ASSERT(k >= p.n_q || (id >= 0 && id < p.V));
        if (k < p.n_q) raw[u] = *reinterpret_cast<const uint4*>(p.emb + ((size_t)k * p.V + id) * d + e);
      }
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        if (k0 + u < p.n_q) {
          float v[8];
          mg_unpack8(raw[u], v);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = (k0 + u) == 0 ? v[j] : bf16_round(acc[j] + v[j]);
        }
      }
    }
    __nv_bfloat162 o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = __floats2bfloat162_rn(acc[2 * j], acc[2 * j + 1]);
    *reinterpret_cast<uint4*>(p.x + (size_t)b * d + e) = *reinterpret_cast<uint4*>(o);
    *reinterpret_cast<uint4*>(p.x + (size_t)(b + p.B) * d + e) = *reinterpret_cast<uint4*>(o);
  }
}

// Once-per-step code kept out of line so that it does not inflate the per-layer instruction footprint.
__device__ __noinline__ void mg_sample(const SampleParams& sp, int k, int b, unsigned char* scratch) {
  sample_body<SyncConsumers>(sp, k, b, sample_smem_carve(scratch));
}
__device__ __noinline__ void mg_update(const UpdateParams& up, int* s_alive) { update_body<SyncConsumers>(up, s_alive, MG_CONSUMERS); }

// =================================================================================================================
// the kernel
// =================================================================================================================
// TR = false (production): every trace / profile stamp is compiled out.  A build with seven more stamps in the step tail
// benched 3.7 % slower: instrumentation in this kernel is not free even when it is switched off at run time.
template <int G, bool TR, bool PK>
__global__ void __launch_bounds__(MG_THREADS, 1) k_decode_mega(const __grid_constant__ MegaParams p) {
  extern __shared__ __align__(128) unsigned char mg_smem[];
  MgCtx cx;
  cx.n_stages = p.n_stages;
  cx.ring = mg_smem;
  cx.ring_addr = mg_smem_u32(mg_smem);
  cx.scratch = cx.ring + (size_t)p.n_stages * MG_STAGE_BYTES;
  cx.part = reinterpret_cast<float*>(cx.scratch + p.scratch_bytes);
  cx.red = cx.part + 2 * 2 * MG_WARPS * 64;
  cx.s_rope = cx.red + 160;
  cx.s_xown = cx.s_rope + 8 * ZB_DH;
  cx.s_page = reinterpret_cast<int*>(cx.s_xown + 8 * MG_OWN);
  unsigned char* bars = reinterpret_cast<unsigned char*>(cx.s_page + 16);
  cx.full0 = mg_smem_u32(bars);
  cx.empty0 = cx.full0 + 8u * MG_MAX_STAGES;
  volatile int* s_flags = reinterpret_cast<volatile int*>(bars + 16 * MG_MAX_STAGES);   // [0] stop, [1] consumer cursor, [2] misc, [3] step mailbox for the epilogue warp (KV position, -1 = stop)
  cx.stage = 0; cx.parity = 0; cx.tile_seq = 0;
  cx.prof = TR && p.dbg != nullptr; cx.t_full = 0; cx.t_fc1 = cx.t_fc1_full = cx.t_fc1_sync = 0;
  cx.trace = false; cx.fine = false; cx.afine = false; cx.tfine = false; cx.lfine = false; cx.trace_n = 0; cx.trace_out = p.dbg ? p.dbg + (size_t)blockIdx.x * MG_TRACE_SLOTS : nullptr;

  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    for (int s = 0; s < p.n_stages; ++s) { mg_mbar_init(cx.full0 + 8u * s, 1); mg_mbar_init(cx.empty0 + 8u * s, MG_WARPS); }
    s_flags[0] = 0; s_flags[1] = 0; s_flags[2] = 0; s_flags[3] = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const bool with_heads = p.mode == MG_MODE_LOGITS || p.mode == MG_MODE_DECODE;
  const bool final_ln = p.mode == MG_MODE_HIDDEN && p.hidden_out != nullptr && (int)blockIdx.x < p.R;
  const int qkv_rows = p.q_size + 2 * p.kv_size;
  ZbDevState* st = p.st;

  if (warp == MG_WARPS) {
    // ================================================ producer warp ================================================
    const int lane = tid & 31;
    const int pos0 = mg_ld_volatile(&st->pos);
    const int rope0 = lane < p.R ? p.rope_pos[lane] : 0;
    if (!mg_ld_volatile(&st->done)) {
      MgSched ld;
      mg_sched_init(ld, p, pos0, rope0, with_heads, final_ln);
      const uint64_t pol_stream = mg_policy_evict_first();
      MgDesc d;
      // optional profile of this warp (clock64 ticks): [0] schedule arithmetic, [1] waiting for a free slot, [2] issue, [3] stages
      long long t_sched = 0, t_wait = 0, t_issue = 0, n_ld = 0;
      int n_fine = 0;
      const bool prof = TR && p.dbg != nullptr;
      for (;;) {
        const long long c0 = prof ? clock64() : 0;
        if (!mg_sched_next(ld, p, d)) break;
        const long long c1 = prof ? clock64() : 0;
        // wait for a free ring slot
        int go = 1;
        if (lane == 0) {
          // back off between polls: the ring is full most of the time, and a hot spin here steals issue slots from
          // consumer warps 0 and 4, which share this warp's scheduler (they became the stragglers of every tile)
          while (!mg_mbar_try_wait(cx.empty0 + 8u * cx.stage, cx.parity ^ 1)) {
            if (*s_flags) { go = 0; break; }
            if (p.producer_sleep_ns) __nanosleep(p.producer_sleep_ns);
          }
          if (go) mg_mbar_expect_tx(cx.full0 + 8u * cx.stage, d.total);
        }
        if (!__shfl_sync(0xffffffffu, go, 0)) break;
        const long long c2 = prof ? clock64() : 0;
        if (prof && p.dbg_fine == 1 && lane == 0 && d.tag == ((uint32_t)(p.dbg_kind == EPI_QKV ? MG_PH_QKV : p.dbg_kind == EPI_RESID ? MG_PH_OUT : MG_PH_FC1) | (1u << 8) | ((uint32_t)p.dbg_step << 16)) && n_fine < 32)
          p.dbg[(size_t)blockIdx.x * MG_TRACE_SLOTS + 32 + n_fine++] = mg_globaltimer();     // issue time of this fc1 stage
        if (d.src != nullptr) {
          ZB_ASSERT(cx.stage >= 0 && cx.stage < p.n_stages && d.dst_off + d.bytes <= MG_STAGE_BYTES && (d.bytes & 15) == 0);
          const uint32_t dst = mg_smem_u32(cx.ring + (size_t)cx.stage * MG_STAGE_BYTES + d.dst_off);
          if (d.tag != 0 && p.stream_evict_first) mg_bulk_g2s_hint(dst, d.src, d.bytes, cx.full0 + 8u * cx.stage, pol_stream);   // weight tiles
          else mg_bulk_g2s(dst, d.src, d.bytes, cx.full0 + 8u * cx.stage);
        }
        mg_advance(cx);
        if (prof) { const long long c3 = clock64(); t_sched += c1 - c0; t_wait += c2 - c1; t_issue += c3 - c2; ++n_ld; }
      }
      if (prof && lane == 0) {
        unsigned long long* o = p.dbg + (size_t)gridDim.x * MG_TRACE_SLOTS + (size_t)blockIdx.x * 8;
        o[0] = t_sched; o[1] = t_wait; o[2] = t_issue; o[3] = n_ld;
      }
    }
    // ---- drain: wait until the consumers are done, then until every copy this warp issued has landed
    if (lane == 0) {
      while (!s_flags[0]) {}
      __threadfence_block();
      // stages [consumer cursor, producer cursor) were issued but never consumed
      int cs = s_flags[1] & 0xffff, cp = (s_flags[1] >> 16) & 1;
      while (cs != cx.stage || cp != cx.parity) {
        mg_mbar_wait(cx.full0 + 8u * cs, cp);
        if (++cs == cx.n_stages) { cs = 0; cp ^= 1; }
      }
    }
    return;
  }

  if (warp == MG_EPI_WARP) {
    // ================================================ epilogue warp ================================================
    // Mirrors the consumers' step / layer / phase nest; blocks on the hand-off barriers in between (LayerNorm staging,
    // attention, grid barriers and sampling are the MMA warps' business).
    MgEpi epi;
    { int own_hi; mg_range(p.d / 2, epi.own_lo, own_hi); epi.own_lo *= 2; }
    epi.logits = p.logits; epi.ll = p.x_ll; epi.ll_epoch = 0;
    for (int step = 0; step < p.n_steps; ++step) {
      mg_bar_sync(MG_BAR_STEP, MG_HANDOFF);
      const int pos = s_flags[3];
      if (pos < 0) break;
      epi.pos = pos; epi.cfg_scale = 0.f; epi.eos_bias = 0;
      // ONE mg_gemv_epi call site for all five phase kinds (layer n_layer = the heads pseudo-layer): every copy of this
      // code that the layer loop does not contain is instruction-cache footprint the SM does not have to refetch 26 times
      // per step (cold straight-line code runs at ~0.2 us per 128-byte line here, profiles/r02_summary.md)
      for (int l = 0; l <= p.n_layer; ++l) {
        const bool heads_l = l == p.n_layer;
        if (heads_l && !with_heads) break;
        epi.layer = l;
        epi.kv_layer = p.kv + (size_t)l * p.kv_layer_elems;
        if (heads_l) {
          epi.cfg_scale = *reinterpret_cast<const volatile float*>(&st->cfg_scale);
          epi.eos_bias = mg_ld_volatile(&st->apply_eos_bias);
        }
        const int n_ph = heads_l ? 1 : 4;
#pragma unroll 1
        for (int ph = 0; ph < n_ph; ++ph) {
          int n_units, kind;
          if (ph == 0) {
            kind = heads_l ? EPI_HEADS : EPI_QKV; n_units = heads_l ? p.n_q * p.V / 2 : qkv_rows / 2; epi.out = p.q; epi.out_stride = p.q_size;
            epi.ll = p.use_ll >= 2 ? p.qkv_ll : nullptr; epi.ll_epoch = mg_ll_epoch(p, step, l, 0);
          }
          else if (ph == 2) {
            kind = EPI_SWIGLU; n_units = p.F / 2; epi.out = p.h; epi.out_stride = p.F;
            epi.ll = p.use_ll >= 3 ? p.h_ll : nullptr; epi.ll_epoch = mg_ll_epoch(p, step, l, 0);
          }
          else {
            kind = EPI_RESID; n_units = p.d / 2; epi.out = p.x; epi.out_stride = p.d;
            epi.ll = p.x_ll + (ph == 3 ? (size_t)8 * (p.d >> 1) : 0); epi.ll_epoch = mg_ll_epoch(p, step, l, ph == 3);
          }
          mg_gemv_epi(p, cx, n_units, kind, epi, !(p.use_ll >= 3 && p.chain) || heads_l || (!with_heads && l == p.n_layer - 1 && ph == 3));
        }
      }
    }
    return;
  }

  // ================================================== consumer warps ==================================================
  unsigned bar_target = 0;
  const long long t_kernel0 = clock64();
  bf16* xs = reinterpret_cast<bf16*>(cx.scratch);
  const int ld_d = p.d + MG_XPAD;
  int own_lo, own_hi;                                 // residual-stream columns owned by this CTA (out_proj / fc2 rows)
  mg_range(p.d / 2, own_lo, own_hi);
  own_lo *= 2; own_hi *= 2;

  const bool embed_ahead = p.mode == MG_MODE_DECODE || p.mode == MG_MODE_FORCED;
  for (int step = 0; step < p.n_steps; ++step) {
    {
      // step mailbox: the epilogue warp learns the KV position of this step (or that the launch is over)
      const int done = mg_ld_volatile(&st->done);
      if (tid == 0) s_flags[3] = done ? -1 : mg_ld_volatile(&st->pos);
      mg_bar_sync(MG_BAR_STEP, MG_HANDOFF);
      if (done) break;
    }
    cx.trace = TR && p.dbg != nullptr && step == p.dbg_step && !p.dbg_fine;
    cx.fine = TR && p.dbg != nullptr && step == p.dbg_step && p.dbg_fine == 1;
    const bool afine_step = TR && p.dbg != nullptr && step == p.dbg_step && p.dbg_fine == 2;
    cx.tfine = TR && p.dbg != nullptr && step == p.dbg_step && p.dbg_fine == 3;
    const bool lfine_step = TR && p.dbg != nullptr && step == p.dbg_step && p.dbg_fine == 4;
    cx.trace_n = 0;
    MG_TSTAMP();                                                  // t0: step start
    MG_STAMP();                                                   // 0: step start
    const int pos = mg_ld_volatile(&st->pos), col = mg_ld_volatile(&st->col), Lbuf = mg_ld_volatile(&st->L);

    // ---- step input.  In the embedding modes only the first step of a launch embeds here: later steps get x from
    // CTA 0 at the end of the previous step (right after the frame write-back), which saves a grid barrier per step.
    const bool input_here = p.mode == MG_MODE_HIDDEN || step == 0 || !embed_ahead;
    if (p.mode == MG_MODE_HIDDEN) {
      if ((int)blockIdx.x < p.R) {
        const int r = blockIdx.x;
        const bf16* src = p.hidden_in + ((size_t)r * p.hidden_S + p.hidden_t0 + step) * p.d;
        for (int e = tid * 8; e < p.d; e += MG_CONSUMERS * 8)
          *reinterpret_cast<uint4*>(p.x + (size_t)r * p.d + e) = *reinterpret_cast<const uint4*>(src + e);
      }
    } else if (input_here && (int)blockIdx.x < p.B) {
      mg_embed_row(p, blockIdx.x, col, Lbuf);
    }
    // RoPE rows of this step (ring) and the KV page of the current position
    {
      const uint32_t sp = mg_cons_wait_addr(cx);
      if (tid < p.R * (MG_ROPE_BYTES / 16)) {
        const int r = tid / (MG_ROPE_BYTES / 16), c = tid % (MG_ROPE_BYTES / 16);
        *reinterpret_cast<uint4*>(cx.s_rope + r * ZB_DH + c * 4) = mg_lds128(sp + r * MG_ROW_BYTES + c * 16);
      }
      ZB_ASSERT(pos >= 0 && pos / ZB_PAGE < p.max_pages);
      if (tid < p.R) cx.s_page[tid] = p.page_table[tid * p.max_pages + pos / ZB_PAGE];
      mg_cons_release(cx);
    }
    if (input_here) mg_grid_sync(p.bar_counter, bar_target);
    MG_STAMP();                                                   // 1: input written + barrier
    // this CTA's columns of the residual stream (kept in shared memory for the whole step)
    if (tid < p.R * MG_OWN) {
      const int r = tid / MG_OWN, i = tid % MG_OWN;
      if (own_lo + i < own_hi) cx.s_xown[r * MG_OWN + i] = __bfloat162float(__ldcg(p.x + (size_t)r * p.d + own_lo + i));
    }

    const bool ll = p.use_ll != 0, ll2 = p.use_ll >= 2, ll3 = p.use_ll >= 3;
    const unsigned long long* x_ll1 = p.x_ll + (size_t)8 * (p.d >> 1);      // fc2's half of x_ll
    // Layer loop with ONE call site each for the LayerNorm staging, the plain staging, the attention and the GEMV phase
    // (layer n_layer = final LN + heads).  The first version had every phase inlined separately (5 GEMV phases, 8
    // LayerNorm stagings): ~94 KB of SASS per trip of the layer loop, refetched by every SM 26 times per step.
#pragma unroll 1
    for (int l = 0; l <= p.n_layer; ++l) {
      const bool heads_l = l == p.n_layer;
      if (heads_l) {
        MG_TSTAMP();                                              // t1: layers done
        if (!with_heads) {
          // No heads phase follows: without the epoch-tagged stream the last fc2 still needs its barrier (final_ln reads
          // p.x), and with it CTA 0's embed-ahead below must not overwrite p.x while other CTAs' fc2 epilogues are still
          // storing their columns of this step into it (MG_MODE_FORCED: the audio-prefix columns of zb_gen_begin).
          if (!ll || embed_ahead) mg_grid_sync(p.bar_counter, bar_target);
          break;
        }
      }
      bf16* kv_layer = p.kv + (size_t)l * p.kv_layer_elems;
      const int n_ph = heads_l ? 1 : 4;
#pragma unroll 1
      for (int ph = 0; ph < n_ph; ++ph) {
        // ---- cross-CTA dependencies of this phase
        if (ph == 1) {
          // use_ll >= 2: q / k / v and y carry their own epoch tags (qkv_ll, y_ll), so neither barrier is needed: the
          // CTAs with an attention item poll the in_proj outputs, everyone polls y in the staging below.  Overwriting a
          // tagged buffer in layer l + 1 is safe without a barrier because every phase consumes ALL of the previous
          // phase's output: a CTA that reaches phase n + 2 has seen every CTA finish phase n + 1, i.e. finish reading
          // phase n's output (the epoch of a word is only ever polled for, never waited past).
          if (!ll2) mg_grid_sync(p.bar_counter, bar_target);
          MG_STAMP();                                             // +2: barrier
          cx.afine = afine_step && l == 1;
          mg_attention<G>(p, cx, kv_layer, pos, mg_ll_epoch(p, step, l, 0), ll2);
          cx.afine = false;
          MG_STAMP();                                             // +3: attention done
          if (!ll2) mg_grid_sync(p.bar_counter, bar_target);
          MG_STAMP();                                             // +4: barrier
        } else if ((ph == 3 && !ll3) || (!ll && (ph == 2 || l > 0))) {
          // fc1 -> fc2 meets at a barrier unless h travels tagged (use_ll >= 3); out_proj -> LN2 and fc2 -> next LN1 only
          // without the epoch-tagged stream
          mg_grid_sync(p.bar_counter, bar_target);
          MG_STAMP();
        }
        // ---- activation staging + GEMV phase parameters
        int K, n_units, gate_off = 0, kind;
        if (ph == 0 || ph == 2) {
          K = p.d;
          const unsigned long long* src_ll = nullptr;
          unsigned epoch = 0;
          if (ph == 2) { kind = EPI_SWIGLU; n_units = p.F / 2; gate_off = p.F; if (ll) { src_ll = p.x_ll; epoch = mg_ll_epoch(p, step, l, 0); } }
          else {
            kind = heads_l ? EPI_HEADS : EPI_QKV; n_units = heads_l ? p.n_q * p.V / 2 : qkv_rows / 2;
            if (ll && l > 0) { src_ll = x_ll1; epoch = mg_ll_epoch(p, step, l - 1, 1); }
          }
          cx.lfine = TR && lfine_step && l == 1 && ph == 2;
          mg_stage_ln(cx, xs, K + MG_XPAD, p.x, K, p.R, p.eps, src_ll, epoch);
          cx.lfine = false;
        } else {
          K = ph == 1 ? p.q_size : p.F;
          kind = EPI_RESID; n_units = p.d / 2;
          if (ph == 1 ? ll2 : ll3) mg_stage_ll(xs, K + MG_XPAD, ph == 1 ? p.y_ll : p.h_ll, K, p.R, mg_ll_epoch(p, step, l, 0));
          else mg_stage_copy(xs, K + MG_XPAD, ph == 1 ? p.y : p.h, K, p.R);
        }
        MG_STAMP();                                               // activations staged
        if (heads_l) MG_TSTAMP();                                 // t2: final LN staged
        mg_gemv_phase<PK>(p, cx, xs, K + MG_XPAD, K, n_units, gate_off, kind, l, !(ll3 && p.chain) || heads_l || (!with_heads && l == p.n_layer - 1 && ph == 3));
        MG_STAMP();                                               // GEMV phase done
      }
    }

    if (with_heads) {
      MG_TSTAMP();                                                // t3: heads done
      mg_grid_sync(p.bar_counter, bar_target);
      MG_TSTAMP();                                                // t4: barrier
      if (p.mode == MG_MODE_DECODE) {
        if ((int)blockIdx.x < p.n_q * p.B) {
          SampleParams sp;
          sp.logits = p.logits; sp.history = p.delayed; sp.hist_stride = Lbuf;
          sp.hist_len = mg_ld_volatile(&st->use_history) ? col + 1 : 0;
          sp.tokens = p.tokens; sp.probs_out = nullptr; sp.n_q = p.n_q; sp.V = p.V;
          const volatile zb_sampling* vs = &st->sp;
          sp.sp.temperature = vs->temperature; sp.sp.top_p = vs->top_p; sp.sp.top_k = vs->top_k; sp.sp.min_p = vs->min_p;
          sp.sp.linear = vs->linear; sp.sp.conf = vs->conf; sp.sp.quad = vs->quad;
          sp.sp.repetition_penalty = vs->repetition_penalty; sp.sp.repetition_penalty_window = vs->repetition_penalty_window;
          sp.seed = *reinterpret_cast<const volatile unsigned long long*>(&st->seed);
          sp.step = (unsigned long long)mg_ld_volatile(&st->step);
          sp.st = nullptr;
          sp.dbg = (cx.tfine && blockIdx.x == 0) ? p.dbg + (size_t)gridDim.x * MG_TRACE_SLOTS + (size_t)gridDim.x * 8 : nullptr;
          mg_sample(sp, blockIdx.x % p.n_q, blockIdx.x / p.n_q, cx.scratch);
          if (tid == 0) {
            // hint: pull the embedding row of the sampled token into L2 now; CTA 0 gathers it after the write-back
            long long id = p.tokens[blockIdx.x];
            id = id < 0 ? 0 : (id >= p.V ? p.V - 1 : id);
            ZB_ASSERT(id >= 0 && id < p.V);
            const bf16* rowp = p.emb + ((size_t)(blockIdx.x % p.n_q) * p.V + id) * p.d;
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(rowp), "r"(p.d * 2) : "memory");
          }
        }
        MG_TSTAMP();                                              // t5: sampled
        mg_grid_sync(p.bar_counter, bar_target);
        MG_TSTAMP();                                              // t6: barrier
      }
    } else if (final_ln) {
      // plugin-level forward: norm_f of the residual stream (backbone/_torch.py:80)
      const int r = blockIdx.x;
      if (ll) mg_stage_ln(cx, xs, ld_d, p.x + (size_t)r * p.d, p.d, 1, p.eps, x_ll1 + (size_t)r * (p.d >> 1), mg_ll_epoch(p, step, p.n_layer - 1, 1));
      else mg_stage_ln(cx, xs, ld_d, p.x + (size_t)r * p.d, p.d, 1, p.eps);
      bf16* dst = p.hidden_out + ((size_t)r * p.hidden_S + p.hidden_t0 + step) * p.d;
      for (int e = tid * 8; e < p.d; e += MG_CONSUMERS * 8) *reinterpret_cast<uint4*>(dst + e) = *reinterpret_cast<const uint4*>(xs + e);
    }

    // ---- advance the device state (CTA 0), then make it visible to everyone
    if (blockIdx.x == 0) {
      if (p.mode == MG_MODE_DECODE) {
        UpdateParams up;
        up.st = st; up.delayed = p.delayed; up.tokens = p.tokens; up.remaining = p.remaining; up.stopping = p.stopping;
        up.out_len = p.out_len; up.rope_pos = p.rope_pos; up.rows = p.R; up.n_q = p.n_q; up.eos = p.eos; up.mask = p.mask_tok;
        mg_update(up, const_cast<int*>(&s_flags[2]));
        MG_TSTAMP();                                              // t7 (CTA 0): state machine done
      } else {
        if (tid == 0) {
          st->pos = pos + 1;
          if (p.mode == MG_MODE_FORCED) st->col = col + 1;
        }
        for (int r = tid; r < p.R; r += MG_CONSUMERS) p.rope_pos[r] += 1;
      }
      if (embed_ahead && step + 1 < p.n_steps) {
        SyncConsumers::sync();                                    // the frame write-back above is visible CTA-wide
        if (!mg_ld_volatile(&st->done))
          for (int b = 0; b < p.B; ++b) mg_embed_row(p, b, col + 1, Lbuf);
        MG_TSTAMP();                                              // t8 (CTA 0): next input embedded
      }
    }
    mg_grid_sync(p.bar_counter, bar_target);
    MG_TSTAMP();                                                  // end-of-step barrier
  }

  // ---- tell the producer where the consumers stopped
  SyncConsumers::sync();
  if (cx.prof && tid == 0) {
    unsigned long long* o = p.dbg + (size_t)gridDim.x * MG_TRACE_SLOTS + (size_t)blockIdx.x * 8;
    o[4] = cx.t_full; o[5] = clock64() - t_kernel0; o[6] = cx.t_fc1; o[7] = (unsigned long long)cx.t_fc1_full | ((unsigned long long)cx.t_fc1_sync << 32);
  }
  if (tid == 0) {
    s_flags[1] = cx.stage | (cx.parity << 16);
    __threadfence_block();
    s_flags[0] = 1;
  }
}
