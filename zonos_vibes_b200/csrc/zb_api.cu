// C ABI of libzonos_b200.so — backbone/generation side.  See include/zonos_b200.h for the contract.
#include <stdarg.h>
#include <string.h>

#include <map>
#include <string>
#include <algorithm>
#include <vector>

#include "attention_kernels.cuh"
#include "bt_kernels.cuh"
#include "decode_kernels.cuh"
#include "gemv_mma.cuh"
#include "hybrid_kernels.cuh"
#include "mega_kernel.cuh"
#include "sample_kernels.cuh"
#include "prefix_kernels.cuh"

int ZbErr::fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  msg = buf;
  return code;
}

static std::string g_create_error;

namespace {

struct Slot {
  void* ptr = nullptr;
  int64_t numel = 0;
  int dtype = ZB_BF16;
  bool provided = false;
  int64_t alt_numel = 0;   // heads: 1025 rows accepted, pad row zeroed
};

struct LayerW {
  bf16 *ln1_w, *ln1_b, *w_in, *w_out, *ln2_w, *ln2_b, *w_fc1, *w_fc2;
};

struct LayerM {   // Mamba2 layer (hybrid backbone)
  bf16 *ln_w, *ln_b, *w_in, *conv_w, *conv_b, *norm_w, *w_out;
  float *dt_bias, *A_log, *D;
};

struct Pass {   // one GEMV launch covers template rows mapped by (base0, base1, lim0, lim1)
  int R, base0, base1, lim0, lim1;
};

}  // namespace

struct zb_ctx {
  zb_config cfg{};
  int device = 0;
  ZbErr err;
  // weights
  char* arena = nullptr;
  size_t arena_bytes = 0;
  std::map<std::string, Slot> slots;
  std::vector<LayerW> layers;
  bf16 *normf_w = nullptr, *normf_b = nullptr, *emb = nullptr, *heads = nullptr;
  float* rope = nullptr;
  bool finalized = false;
  int64_t step_weight_bytes = 0;
  // kv cache
  bf16* kv = nullptr;
  size_t kv_layer_elems = 0;
  int* page_table = nullptr;
  int rows = 0, max_pages = 0, max_seqlen = 0;
  // scratch
  bf16 *x = nullptr, *q = nullptr, *y = nullptr, *h = nullptr;
  float *logits = nullptr, *part_o = nullptr, *part_ml = nullptr;
  int* counters = nullptr;
  int nsplit = 1;
  long long* tokens = nullptr;
  int *remaining = nullptr, *stopping = nullptr, *out_len = nullptr, *rope_pos = nullptr;
  ZbDevState* st = nullptr;
  // generation
  long long* delayed = nullptr;
  int B = 0, L = 0;
  cudaGraphExec_t graph = nullptr;
  int graph_nodes = 0;
  int graph_B = -1, graph_L = -1;
  long long* graph_delayed = nullptr;
  // options
  int opt_pdl = 1, opt_l2_prefetch_mb = 0, opt_graph = 1, opt_gemv = 1;
  int opt_attn_mma = 1;     // 1: mma.sync attention kernel, 0: SIMT (A/B and cross-check)
  int opt_nsplit = 0;       // > 0: KV splits of the multi-kernel / batched attention (default: 4 CTAs per SM worth), read by zb_kv_alloc
  int opt_attn_tma = 1;     // batched path: K/V staged through shared memory by cp.async.bulk (k_attn_decode_tma) instead of ld.global fragments
                            // (k_attn_decode_mma).  0 = never, 1 = when rows >= 64 (measured: batch 64 -9 %, batch 32 -3 %, batch 16 +-0 at
                            // ctx 1 300; hybrid batch 16 at short context +2 %), 2 = always, 3 = always, ring slot released before the math
  int64_t launches = 0;
  bool counting = true;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;     // zb_bench_kernel
  // zb_gen_steps calls in flight: each has its own event pair and a pinned snapshot of the device state taken right
  // after it, so that the host can enqueue call k+1 before it waits for call k (zb_gen_poll pops the oldest)
  struct GenSlot { cudaEvent_t e0 = nullptr, e1 = nullptr, e2 = nullptr; ZbDevState* h_st = nullptr; int32_t* h_len = nullptr; };
  static constexpr int kGenSlots = 4, kGenMaxB = 1024;
  GenSlot gen[kGenSlots];
  int gen_head = 0, gen_tail = 0;               // issued / polled counters (slot = counter % kGenSlots)
  float last_ms = -1.f;
  cudaStream_t cap_stream = nullptr;   // capture happens here: the caller's stream may be the legacy default stream
  // hybrid backbone (hybrid_kernels.cuh)
  bool hybrid = false;
  zb_hybrid hyb{};
  std::vector<int> is_mamba;           // per layer
  std::vector<int> kv_slot;            // attention layers: index into the KV cache; Mamba2 layers: index into the states
  int n_attn = 0, n_mamba = 0;
  std::vector<LayerM> mlayers;
  int m_dinner = 0, m_nheads = 0, m_convdim = 0, m_dinproj = 0;
  float *m_conv_state = nullptr, *m_ssm_state = nullptr, *m_xbc = nullptr;
  // layer-major prefill workspace of the hybrid backbone (token-major activations of the whole prefix, grown on demand)
  bf16 *pf_x = nullptr, *pf_zx = nullptr, *pf_y = nullptr, *pf_yn = nullptr;
  float* pf_xbc = nullptr;
  int pf_M = 0;
  int opt_chunked_prefill = 1;         // hybrid: prefill Mamba2 layers with the chunked SSD scan (0 = step token by token)
  bf16 *m_zx = nullptr, *m_y = nullptr, *m_ynorm = nullptr;
  // batched tcgen05 path (bt_kernels.cuh): rows >= opt_bt_min_rows
  typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  EncodeTiledFn encode = nullptr;
  int opt_bt_min_rows = 9;
  int opt_bt_prefill = 1;              // transformer: prefill blocks of tokens as extra rows of the batched tcgen05 path
  bf16* pf_embed = nullptr;            // block prefill of an audio prefix: embedded prefix columns [rows][P][d]
  size_t pf_embed_elems = 0;
  size_t bt_partial_elems = 0;         // floats behind bt_partial
  int bt_vM = 0, bt_vrows = 0;         // block prefill in progress: bt_vM activation rows = tokens x bt_vrows cache rows
  bool bt_ready = false;
  int bt_mpad = 0;
  bf16* bt_xn = nullptr;
  float* bt_partial = nullptr;
  CUtensorMap bt_map_xn{}, bt_map_y{}, bt_map_h{}, bt_map_heads{}, bt_map_yn{};   // yn: gated-norm output of a Mamba2 layer (hybrid)
  CUtensorMap bt_maps_xn{}, bt_maps_y{}, bt_maps_h{}, bt_maps_yn{};                // the same activations with a box of 128 / bt_cs rows (cluster multicast slices)
  int bt_cs = 1;                       // cluster size the slice maps were built for
  CUtensorMap bt_map2_xn{}, bt_map2_y{}, bt_map2_h{}, bt_map2_yn{};                // ... with a box of 64 rows: the half of the N operand one CTA of a pair loads (k_bt_gemm2)
  int opt_bt_cta2 = 1;                 // batched path, k_bt_gemm2 (CTA pairs, tcgen05.mma.cta_group::2, weights as the M operand): bit 0 = the
                                       // split-K GEMMs (default: 1 817 -> 1 667 us per batch-64 step at short context), bit 1 = fc1
                                       // (SwiGLU epilogue; measured slower than k_bt_gemm: 1 854, profiles/r02_bt_cta2_ab.txt)
  int opt_bt_cluster = 1;              // batched path: CTAs per cluster sharing (multicasting) the activation tile: 1, 2 or 4.  Measured
                                       // at batch 64: 1 837 / 1 891 / 2 060 us per step for 1 / 2 / 4 - the GEMMs are bound by what is delivered INTO the
                                       // SMs (weights + activation tile per CTA), which multicast does not reduce; off by default
  std::vector<CUtensorMap> bt_map_in, bt_map_out, bt_map_fc1, bt_map_fc2;
  std::vector<CUtensorMap> bt_map_fc1p;                                             // fc1 as value / gate row pairs (k_bt_gemm2)
  // persistent decode kernel (mega_kernel.cuh)
  MegaLayer mega_l0{};
  long long mega_layer_stride = 0;
  bool mega_ok = false;
  unsigned* bar_counter = nullptr;
  unsigned long long* x_ll = nullptr;  // persistent kernel: epoch-tagged residual stream [2][8][d/2]
  unsigned long long* att_ll = nullptr;   // persistent kernel: epoch-tagged split-KV partials [8 rows][Hkv][MG_MAX_SPLIT][G * (Dh + 2)]
  unsigned long long *qkv_ll = nullptr, *y_ll = nullptr, *h_ll = nullptr;   // tagged q/k/v, attention output, SwiGLU output (same allocation as x_ll)
  size_t ll_bytes = 0;                 // size of that allocation
  int opt_chain = 0;                   // persistent kernel, ll = 3: MMA warps run on into the next phase without waiting for their epilogue warp
  unsigned launch_seq = 0;             // persistent-kernel launches so far (epoch base of x_ll)
  int opt_ll = 3;                      // persistent kernel: 1 = poll x_ll instead of the grid barriers after out_proj / fc2; 2 = also q/k/v and y
                                       // (no barrier around the attention); 0 = grid barriers everywhere
  int n_sms = ZB_NUM_SMS;
  int smem_optin = 0;
  int opt_mega = 1;
  int opt_ring = 0;                    // cap on the ring depth of the persistent kernel (0 = as many stages as fit)
  int opt_psleep = 100;                // persistent kernel: producer back-off (ns) between polls of a full ring
  int opt_att_single = 256;            // persistent kernel: contexts up to this many tokens are not split over CTAs
  int opt_att_chunk = 96;             // persistent kernel: tokens per KV split at least (short contexts use fewer splits)
  int opt_evict = 1;                   // persistent kernel: weight stream carries the L2 evict_first hint
  int opt_trace = 0;                   // debug: print a per-phase timeline of one step of every persistent launch
  unsigned long long* dbg = nullptr;
  int opt_trace_quiet = 0;             // 1: traced launches keep their stamps for zb_last_trace instead of printing them
  std::vector<unsigned long long> last_trace;   // %globaltimer stamps of CTA 0 in the traced step of the last traced launch
};

namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// -------------------------------------------------------------------------------------------
// launch helper: optional programmatic-dependent-launch attribute on every kernel
// -------------------------------------------------------------------------------------------
template <typename... KArgs, typename... Args>
cudaError_t launch(zb_ctx* c, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  if (c->opt_pdl) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  if (c->counting) c->launches += 1;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

typedef void (*GemvKernel)(const GemvParams);
bool use_bt(const zb_ctx* c, int rows);
bool use_mega(const zb_ctx* c, int rows);
cudaError_t enqueue_bt_layers(zb_ctx* c, cudaStream_t s);
cudaError_t enqueue_bt_heads(zb_ctx* c, float* logits, cudaStream_t s);
int bt_setup(zb_ctx* c, int rows_pad);

template <int EPI>
GemvKernel gemv_kernel_for(int R) {
  switch (R) {
    case 2: return k_gemv<2, EPI>;
    case 4: return k_gemv<4, EPI>;
    case 6: return k_gemv<6, EPI>;
    case 8: return k_gemv<8, EPI>;
  }
  return nullptr;
}
GemvKernel gemv_kernel(int R, int epi) {
  switch (epi) {
    case EPI_QKV: return gemv_kernel_for<EPI_QKV>(R);
    case EPI_RESID: return gemv_kernel_for<EPI_RESID>(R);
    case EPI_SWIGLU: return gemv_kernel_for<EPI_SWIGLU>(R);
    case EPI_HEADS: return gemv_kernel_for<EPI_HEADS>(R);
  }
  return nullptr;
}
template <int EPI>
GemvKernel gemv_mma_kernel_for(int R) {
  switch (R) {
    case 2: return k_gemv_mma<2, EPI>;
    case 4: return k_gemv_mma<4, EPI>;
    case 6: return k_gemv_mma<6, EPI>;
    case 8: return k_gemv_mma<8, EPI>;
  }
  return nullptr;
}
GemvKernel gemv_mma_kernel(int R, int epi) {
  switch (epi) {
    case EPI_QKV: return gemv_mma_kernel_for<EPI_QKV>(R);
    case EPI_RESID: return gemv_mma_kernel_for<EPI_RESID>(R);
    case EPI_SWIGLU: return gemv_mma_kernel_for<EPI_SWIGLU>(R);
    case EPI_HEADS: return gemv_mma_kernel_for<EPI_HEADS>(R);
    case EPI_PLAIN: return gemv_mma_kernel_for<EPI_PLAIN>(R);
  }
  return nullptr;
}
size_t gemv_mma_smem(int R, int K, int epi) {
  const int NT = epi == EPI_SWIGLU ? 2 : 1;
  return (size_t)R * (K + GM_XPAD) * sizeof(bf16) + (size_t)2 * NT * GM_WARPS * 64 * sizeof(float) + (size_t)GM_WARPS * R * sizeof(float);
}
size_t gemv_smem(int R, int K) { return (size_t)R * K * sizeof(bf16) + (size_t)(GEMV_THREADS / 32) * R * sizeof(float); }

typedef void (*AttnKernel)(const AttnParams);
AttnKernel attn_tma_kernel(int G, bool early) {
  switch (G) {
    case 1: return early ? k_attn_decode_tma<1, true> : k_attn_decode_tma<1, false>;
    case 2: return early ? k_attn_decode_tma<2, true> : k_attn_decode_tma<2, false>;
    case 4: return early ? k_attn_decode_tma<4, true> : k_attn_decode_tma<4, false>;
    case 8: return early ? k_attn_decode_tma<8, true> : k_attn_decode_tma<8, false>;
  }
  return nullptr;
}
AttnKernel attn_kernel(int G, bool mma = true) {
  switch (G) {
    case 1: return mma ? k_attn_decode_mma<1> : k_attn_decode<1>;
    case 2: return mma ? k_attn_decode_mma<2> : k_attn_decode<2>;
    case 4: return mma ? k_attn_decode_mma<4> : k_attn_decode<4>;
    case 8: return mma ? k_attn_decode_mma<8> : k_attn_decode<8>;
  }
  return nullptr;
}

std::vector<Pass> cfg_passes(int B) {       // cond/uncond pairs of up to 4 utterances per pass
  std::vector<Pass> v;
  for (int b0 = 0; b0 < B; b0 += 4) {
    const int nb = B - b0 < 4 ? B - b0 : 4;
    v.push_back(Pass{2 * nb, b0, B + b0, B, 2 * B});
  }
  return v;
}
std::vector<Pass> plain_passes(int rows) {  // consecutive rows, 8 per pass
  std::vector<Pass> v;
  for (int r0 = 0; r0 < rows; r0 += 8) {
    int n = rows - r0 < 8 ? rows - r0 : 8;
    int R = (n + 1) & ~1;
    v.push_back(Pass{R, r0, r0 + R / 2, rows, rows});
  }
  return v;
}

struct NextW {   // head of the next weight matrix, for the optional L2 prefetch
  const bf16* ptr = nullptr;
  size_t bytes = 0;
  size_t second_off = 0;   // byte offset of the paired half (fc1 gate rows), 0 = single range
};

void set_prefetch(zb_ctx* c, GemvParams& p, const NextW& nx) {
  p.pf0 = p.pf1 = nullptr;
  p.pf_bytes = 0;
  if (c->opt_l2_prefetch_mb <= 0 || nx.ptr == nullptr) return;
  size_t budget = (size_t)c->opt_l2_prefetch_mb << 20;
  const int nr = nx.second_off ? 2 : 1;
  size_t per = nx.bytes / nr;
  if (per > budget / nr) per = budget / nr;
  per = per / (16 * ZB_NUM_SMS * GEMV_CTAS_PER_SM) * (16 * ZB_NUM_SMS * GEMV_CTAS_PER_SM);
  if (per == 0) return;
  p.pf0 = reinterpret_cast<const char*>(nx.ptr);
  p.pf1 = nx.second_off ? p.pf0 + nx.second_off : nullptr;
  p.pf_bytes = (unsigned)per;
}

GemvParams base_params(zb_ctx* c, const Pass& ps) {
  GemvParams p;
  memset(&p, 0, sizeof(p));
  p.base0 = ps.base0; p.base1 = ps.base1; p.lim0 = ps.lim0; p.lim1 = ps.lim1;
  p.st = c->st;
  p.eps = c->cfg.norm_eps;
  return p;
}

cudaError_t launch_gemv(zb_ctx* c, int epi, const Pass& ps, const GemvParams& pin, cudaStream_t s) {
  GemvParams p = pin;
  // rows that produce outputs: SwiGLU pairs value row i with gate row i+F (n_tasks == F); the others use adjacent pairs
  p.n_rows = epi == EPI_SWIGLU ? pin.n_tasks : pin.n_tasks * 2;
  if (c->opt_gemv == 0 && epi != EPI_PLAIN) {     // SIMT cross-check path
    return launch(c, gemv_kernel(ps.R, epi), dim3(ZB_NUM_SMS * GEMV_CTAS_PER_SM), dim3(GEMV_THREADS), gemv_smem(ps.R, p.K), s, p);
  }
  p.n_tasks = (p.n_rows + GM_ROWS - 1) / GM_ROWS;
  return launch(c, gemv_mma_kernel(ps.R, epi), dim3(ZB_NUM_SMS * 2), dim3(GM_THREADS), gemv_mma_smem(ps.R, p.K, epi), s, p);
}

// One Mamba2 layer for the current token (hybrid backbone): x <- x + out_proj(gated_norm(ssm(conv(in_proj(LN(x))))))
cudaError_t enqueue_mamba_layer(zb_ctx* c, int l, const std::vector<Pass>& passes, cudaStream_t s) {
  const zb_config& g = c->cfg;
  const LayerM& w = c->mlayers[l];
  const int d = g.d_model, rows = c->rows;
  cudaError_t e;
  for (const Pass& ps : passes) {   // LN + in_proj -> zxbcdt
    GemvParams p = base_params(c, ps);
    p.W = w.w_in; p.K = d; p.n_tasks = c->m_dinproj / 2; p.row0_mul = 2; p.row1_off = 1;
    p.x = c->x; p.ln_w = w.ln_w; p.ln_b = w.ln_b;
    p.out = c->m_zx; p.out_stride = c->m_dinproj;
    if ((e = launch_gemv(c, EPI_PLAIN, ps, p, s)) != cudaSuccess) return e;
  }
  M2Params m;
  memset(&m, 0, sizeof(m));
  m.zx = c->m_zx; m.d_in_proj = c->m_dinproj; m.d_inner = c->m_dinner; m.conv_dim = c->m_convdim; m.d_conv = c->hyb.d_conv;
  m.nheads = c->m_nheads; m.headdim = c->hyb.headdim; m.d_state = c->hyb.d_state; m.ngroups = c->hyb.ngroups;
  m.conv_w = w.conv_w; m.conv_b = w.conv_b; m.dt_bias = w.dt_bias; m.A_log = w.A_log; m.D = w.D; m.norm_w = w.norm_w;
  const int slot = c->kv_slot[l];
  m.conv_state = c->m_conv_state + (size_t)slot * rows * c->m_convdim * c->hyb.d_conv;
  m.ssm_state = c->m_ssm_state + (size_t)slot * rows * c->m_nheads * c->hyb.headdim * c->hyb.d_state;
  m.xbc = c->m_xbc; m.y = c->m_y; m.ynorm = c->m_ynorm; m.eps = 1e-5f; m.st = c->st;
  if ((e = launch(c, k_m2_conv, dim3((c->m_convdim + 255) / 256, rows), dim3(256), 0, s, m)) != cudaSuccess) return e;
  if ((e = launch(c, k_m2_ssm, dim3(c->m_nheads, rows), dim3(256), 0, s, m)) != cudaSuccess) return e;
  if ((e = launch(c, k_m2_gated_norm, dim3(rows), dim3(256), 0, s, m)) != cudaSuccess) return e;
  for (const Pass& ps : passes) {   // out_proj + residual
    GemvParams p = base_params(c, ps);
    p.W = w.w_out; p.K = c->m_dinner; p.n_tasks = d / 2; p.row0_mul = 2; p.row1_off = 1;
    p.x = c->m_ynorm; p.out = c->x; p.out_stride = d;
    if ((e = launch_gemv(c, EPI_RESID, ps, p, s)) != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// one attention + GatedMLP layer for the current token (x in c->x, positions in device state)
cudaError_t enqueue_attn_layer(zb_ctx* c, int l, const std::vector<Pass>& passes, cudaStream_t s, bool heads_follow) {
  const zb_config& g = c->cfg;
  const int d = g.d_model, H = g.num_heads, Hkv = g.num_heads_kv, F = g.ffn;
  const int q_size = H * ZB_DH, kv_size = Hkv * ZB_DH;
  cudaError_t e;
  {
    const LayerW& w = c->layers[l];
    bf16* kv_layer = c->kv + (size_t)c->kv_slot[l] * c->kv_layer_elems;
    for (const Pass& ps : passes) {   // LN1 + in_proj + RoPE + KV append
      GemvParams p = base_params(c, ps);
      p.W = w.w_in; p.K = d; p.n_tasks = (q_size + 2 * kv_size) / 2; p.row0_mul = 2; p.row1_off = 1;
      p.x = c->x; p.ln_w = w.ln1_w; p.ln_b = w.ln1_b;
      p.out = c->q; p.out_stride = q_size;
      p.kv_layer = kv_layer; p.page_table = c->page_table; p.max_pages = c->max_pages; p.Hkv = Hkv;
      p.q_size = q_size; p.kv_size = kv_size; p.rope = c->rope; p.rope_pos = c->rope_pos;
      set_prefetch(c, p, NextW{w.w_out, (size_t)d * q_size * 2, 0});
      if ((e = launch_gemv(c, EPI_QKV, ps, p, s)) != cudaSuccess) return e;
    }
    {
      AttnParams a;
      memset(&a, 0, sizeof(a));
      a.q = c->q; a.kv_layer = kv_layer; a.page_table = c->page_table; a.max_pages = c->max_pages;
      a.H = H; a.Hkv = Hkv; a.y = c->y; a.part_o = c->part_o; a.part_ml = c->part_ml; a.counters = c->counters;
      a.st = c->st; a.scale_log2 = (1.0f / sqrtf((float)ZB_DH)) * 1.4426950408889634f; a.total_pages = c->rows * c->max_pages;
      if ((e = launch(c, attn_kernel(H / Hkv, c->opt_attn_mma != 0), dim3(c->nsplit, Hkv, c->rows), dim3(ATT_THREADS), 0, s, a)) != cudaSuccess) return e;
    }
    for (const Pass& ps : passes) {   // out_proj + residual
      GemvParams p = base_params(c, ps);
      p.W = w.w_out; p.K = q_size; p.n_tasks = d / 2; p.row0_mul = 2; p.row1_off = 1;
      p.x = c->y; p.out = c->x; p.out_stride = d;
      set_prefetch(c, p, NextW{w.w_fc1, (size_t)2 * F * d * 2, (size_t)F * d * 2});
      if ((e = launch_gemv(c, EPI_RESID, ps, p, s)) != cudaSuccess) return e;
    }
    for (const Pass& ps : passes) {   // LN2 + fc1 + SwiGLU
      GemvParams p = base_params(c, ps);
      p.W = w.w_fc1; p.K = d; p.n_tasks = F; p.row0_mul = 1; p.row1_off = F;
      p.x = c->x; p.ln_w = w.ln2_w; p.ln_b = w.ln2_b;
      p.out = c->h; p.out_stride = F;
      set_prefetch(c, p, NextW{w.w_fc2, (size_t)d * F * 2, 0});
      if ((e = launch_gemv(c, EPI_SWIGLU, ps, p, s)) != cudaSuccess) return e;
    }
    for (const Pass& ps : passes) {   // fc2 + residual
      GemvParams p = base_params(c, ps);
      p.W = w.w_fc2; p.K = F; p.n_tasks = d / 2; p.row0_mul = 2; p.row1_off = 1;
      p.x = c->h; p.out = c->x; p.out_stride = d;
      NextW nx;
      if (l + 1 < g.n_layer && !c->is_mamba[l + 1]) nx = NextW{c->layers[l + 1].w_in, (size_t)(q_size + 2 * kv_size) * d * 2, 0};
      else if (heads_follow) nx = NextW{c->heads, (size_t)g.n_codebooks * g.vocab * d * 2, 0};
      set_prefetch(c, p, nx);
      if ((e = launch_gemv(c, EPI_RESID, ps, p, s)) != cudaSuccess) return e;
    }
  }
  return cudaSuccess;
}

// all layers for the current token (x in c->x, positions in device state)
cudaError_t enqueue_layers(zb_ctx* c, const std::vector<Pass>& passes, cudaStream_t s, bool heads_follow) {
  cudaError_t e;
  for (int l = 0; l < c->cfg.n_layer; ++l) {
    e = c->is_mamba[l] ? enqueue_mamba_layer(c, l, passes, s) : enqueue_attn_layer(c, l, passes, s, heads_follow);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t enqueue_heads(zb_ctx* c, const std::vector<Pass>& passes, float* logits, cudaStream_t s) {
  const zb_config& g = c->cfg;
  cudaError_t e;
  for (const Pass& ps : passes) {
    GemvParams p = base_params(c, ps);
    p.W = c->heads; p.K = g.d_model; p.n_tasks = g.n_codebooks * g.vocab / 2; p.row0_mul = 2; p.row1_off = 1;
    p.x = c->x; p.ln_w = c->normf_w; p.ln_b = c->normf_b;
    p.logits = logits; p.n_q = g.n_codebooks; p.V = g.vocab; p.eos = g.eos_token_id;
    const LayerW& w0 = c->layers[0];
    if (!c->is_mamba[0]) set_prefetch(c, p, NextW{w0.w_in, (size_t)(g.num_heads + 2 * g.num_heads_kv) * ZB_DH * g.d_model * 2, 0});
    if ((e = launch_gemv(c, EPI_HEADS, ps, p, s)) != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// one full decode-loop iteration (model.py:276-301) — the body that is graph-captured
cudaError_t enqueue_decode_step(zb_ctx* c, cudaStream_t s) {
  const zb_config& g = c->cfg;
  cudaError_t e;
  const std::vector<Pass> passes = cfg_passes(c->B);
  int thr = g.d_model / 8 < 1024 ? g.d_model / 8 : 1024;
  if ((e = launch(c, k_embed, dim3(c->B), dim3(thr), 0, s, (const bf16*)c->emb, (const long long*)c->delayed, c->x,
                  g.d_model, g.n_codebooks, g.vocab, (const ZbDevState*)c->st)) != cudaSuccess) return e;
  if (use_bt(c, 2 * c->B)) {
    if ((e = enqueue_bt_layers(c, s)) != cudaSuccess) return e;
    if ((e = enqueue_bt_heads(c, c->logits, s)) != cudaSuccess) return e;
  } else {
    if ((e = enqueue_layers(c, passes, s, true)) != cudaSuccess) return e;
    if ((e = enqueue_heads(c, passes, c->logits, s)) != cudaSuccess) return e;
  }
  SampleParams sp;
  memset(&sp, 0, sizeof(sp));
  sp.logits = c->logits; sp.history = c->delayed; sp.hist_stride = c->L; sp.tokens = c->tokens;
  sp.n_q = g.n_codebooks; sp.V = g.vocab; sp.st = c->st;
  if ((e = launch(c, k_sample, dim3(g.n_codebooks, c->B), dim3(SMP_THREADS), 0, s, sp)) != cudaSuccess) return e;
  UpdateParams up;
  memset(&up, 0, sizeof(up));
  up.st = c->st; up.delayed = c->delayed; up.tokens = c->tokens; up.remaining = c->remaining; up.stopping = c->stopping;
  up.out_len = c->out_len; up.rope_pos = c->rope_pos; up.rows = 2 * c->B; up.n_q = g.n_codebooks;
  up.eos = g.eos_token_id; up.mask = g.masked_token_id;
  const int ub = c->B <= 8 ? 32 * c->B : 256;        // one warp per utterance
  if ((e = launch(c, k_update, dim3(1), dim3(ub), 0, s, up)) != cudaSuccess) return e;
  return cudaSuccess;
}

// ---------------------------------------------------------------------------------------------------------------
// persistent decode kernel
// ---------------------------------------------------------------------------------------------------------------
typedef void (*MegaKernel)(const MegaParams);
template <int G>
MegaKernel mega_kernel_g(bool trace, bool packed) {
  if (trace) return packed ? k_decode_mega<G, true, true> : k_decode_mega<G, true, false>;
  return packed ? k_decode_mega<G, false, true> : k_decode_mega<G, false, false>;
}
MegaKernel mega_kernel(int G, bool trace, bool packed) {
  switch (G) {
    case 1: return mega_kernel_g<1>(trace, packed);
    case 2: return mega_kernel_g<2>(trace, packed);
    case 4: return mega_kernel_g<4>(trace, packed);
    case 8: return mega_kernel_g<8>(trace, packed);
  }
  return nullptr;
}
// packed-K MMA layout of the persistent kernel's GEMV phases: <= 4 activation rows, every contraction length % 512 == 0
bool mega_packed(const zb_config& g, int R) {
  return R <= 4 && g.d_model % 512 == 0 && (g.num_heads * ZB_DH) % 512 == 0 && g.ffn % 512 == 0;
}

bool use_mega(const zb_ctx* c, int rows) {
  const int own = 2 * ((c->cfg.d_model / 2 + c->n_sms - 1) / c->n_sms);     // residual columns per CTA kept in shared memory
  return c->opt_mega && rows >= 1 && rows <= 8 && c->mega_ok && own <= MG_OWN;
}

// mode: MG_MODE_*.  R activation rows (= 2B in the CFG modes).  One cooperative launch runs `n_steps` steps.
cudaError_t launch_mega(zb_ctx* c, int mode, int n_steps, int R, int B, float* logits, const void* hidden_in, void* hidden_out,
                        int hidden_S, int hidden_t0, cudaStream_t s) {
  const zb_config& g = c->cfg;
  const int G = g.num_heads / g.num_heads_kv;
  MegaParams p;
  memset(&p, 0, sizeof(p));
  p.l0 = c->mega_l0; p.layer_stride = c->mega_layer_stride; p.n_layer = g.n_layer; p.d = g.d_model; p.q_size = g.num_heads * ZB_DH; p.kv_size = g.num_heads_kv * ZB_DH;
  p.F = g.ffn; p.H = g.num_heads; p.Hkv = g.num_heads_kv;
  p.normf_w = c->normf_w; p.normf_b = c->normf_b; p.heads = c->heads; p.emb = c->emb; p.rope = c->rope;
  p.n_q = g.n_codebooks; p.V = g.vocab; p.eos = g.eos_token_id; p.mask_tok = g.masked_token_id;
  p.eps = g.norm_eps; p.scale_log2 = (1.0f / sqrtf((float)ZB_DH)) * 1.4426950408889634f;
  p.x = c->x; p.q = c->q; p.y = c->y; p.h = c->h; p.logits = logits ? logits : c->logits;
  p.kv = c->kv; p.kv_layer_elems = c->kv_layer_elems; p.page_table = c->page_table; p.max_pages = c->max_pages;
  p.att_ll = c->att_ll;
  int ns = c->n_sms / (R * g.num_heads_kv);
  ns = ns < 1 ? 1 : (ns > c->nsplit ? c->nsplit : ns);
  if (ns > MG_MAX_SPLIT) ns = MG_MAX_SPLIT;
  p.nsplit = ns; p.min_chunk = c->opt_att_chunk < 8 ? 8 : (c->opt_att_chunk + 7) & ~7; p.single_ctx = c->opt_att_single;
  p.R = R; p.B = B; p.st = c->st; p.rope_pos = c->rope_pos; p.delayed = c->delayed; p.tokens = c->tokens;
  p.remaining = c->remaining; p.stopping = c->stopping; p.out_len = c->out_len; p.bar_counter = c->bar_counter;
  p.mode = mode; p.n_steps = n_steps;
  p.x_ll = c->x_ll; p.qkv_ll = c->qkv_ll; p.y_ll = c->y_ll; p.h_ll = c->h_ll;
  p.use_ll = (n_steps < 8192 && R <= 4) ? c->opt_ll : 0;     // R > 4: the polling loader needs > 2 round trips, the barrier is cheaper
  if (p.use_ll >= 3 && g.ffn > 2048 && g.ffn % 4096 != 0) p.use_ll = 2;      // mg_stage_ll's wide variant walks 4096 columns per round
  p.epoch_base = ((++c->launch_seq) & 0xfffu) << 20;    // epochs of different launches never coincide
  p.hidden_in = (const bf16*)hidden_in; p.hidden_out = (bf16*)hidden_out; p.hidden_S = hidden_S; p.hidden_t0 = hidden_t0;
  // shared memory plan: [ring][scratch][partials 8 KB][red 256 B][barriers][flags]
  int maxK = g.d_model > p.q_size ? g.d_model : p.q_size;
  if (g.ffn > maxK) maxK = g.ffn;
  size_t scratch = (size_t)R * (maxK + MG_XPAD) * 2;
  const size_t att = ((size_t)MG_MAX_SPLIT * G * ZB_DH + 2 * MG_WARPS * G + 32 * 8 * 2) * 4 + 16;
  if (att > scratch) scratch = att;
  if ((size_t)SMP_SMEM_BYTES > scratch) scratch = SMP_SMEM_BYTES;
  scratch = align_up(scratch, 128);
  const size_t fixed = 2 * 2 * MG_WARPS * 64 * 4 + 160 * 4 + 8 * ZB_DH * 4 + 8 * MG_OWN * 4 + 64 + 16 * MG_MAX_STAGES + 64;
  int stages = (int)(((size_t)c->smem_optin - scratch - fixed) / MG_STAGE_BYTES);
  if (stages > MG_MAX_STAGES) stages = MG_MAX_STAGES;
  if (c->opt_ring > 0 && stages > c->opt_ring) stages = c->opt_ring;
  if (stages < 2) return cudaErrorInvalidConfiguration;
  p.n_stages = stages; p.scratch_bytes = (int)scratch; p.producer_sleep_ns = (unsigned)c->opt_psleep; p.stream_evict_first = c->opt_evict; p.chain = c->opt_chain;
  const size_t smem = (size_t)stages * MG_STAGE_BYTES + scratch + fixed;

  cudaError_t e = cudaMemsetAsync(c->bar_counter, 0, sizeof(unsigned), s);
  if (e != cudaSuccess) return e;
  // the tags of every epoch-tagged word start from zero in every launch: a word that this launch has not written yet can
  // never carry one of its epochs (launch numbers wrap after 4096 launches)
  if ((e = cudaMemsetAsync(c->x_ll, 0, c->ll_bytes, s)) != cudaSuccess) return e;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(c->n_sms);
  cfg.blockDim = dim3(MG_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;       // all CTAs co-resident: they meet at grid barriers
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (c->counting) c->launches += 1;
  if (c->opt_trace) {
    const size_t nb = ((size_t)c->n_sms * (MG_TRACE_SLOTS + 8) + 16) * sizeof(unsigned long long);
    if (!c->dbg && (e = cudaMalloc(&c->dbg, nb)) != cudaSuccess) return e;
    cudaMemsetAsync(c->dbg, 0, nb, s);
    p.dbg = c->dbg; p.dbg_step = n_steps / 2; // trace = 1 per-phase timeline | 2 / 5 / 6 stage level of fc1 / in_proj / out_proj (layer 1) | 3 attention | 4 step tail | 7 LN2 of layer 1
    p.dbg_fine = (c->opt_trace == 2 || c->opt_trace == 5 || c->opt_trace == 6) ? 1 : c->opt_trace == 3 ? 2 : c->opt_trace == 4 ? 3 : c->opt_trace == 7 ? 4 : 0;
    p.dbg_kind = c->opt_trace == 5 ? EPI_QKV : c->opt_trace == 6 ? EPI_RESID : EPI_SWIGLU;
    e = cudaLaunchKernelEx(&cfg, mega_kernel(G, true, mega_packed(g, R)), p);
    if (e != cudaSuccess) return e;
    std::vector<unsigned long long> hbuf((size_t)c->n_sms * (MG_TRACE_SLOTS + 8) + 16);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return e;
    cudaMemcpy(hbuf.data(), c->dbg, nb, cudaMemcpyDeviceToHost);
    c->last_trace.assign(hbuf.begin(), hbuf.begin() + MG_TRACE_SLOTS);
    if (c->opt_trace_quiet) return cudaSuccess;
    const int ctas[3] = {0, p.dbg_fine == 2 ? 5 : c->n_sms / 2, p.dbg_fine == 2 ? 41 : c->n_sms - 1};
    fprintf(stderr, "[mega trace] mode %d steps %d step %d: ns since step start (cta %d / %d / %d)\n", mode, n_steps, p.dbg_step, ctas[0], ctas[1], ctas[2]);
    for (int i = 0; i < MG_TRACE_SLOTS; ++i) {
      if (!hbuf[i] && !p.dbg_fine) break;
      if (!hbuf[i]) continue;
      fprintf(stderr, "  slot %2d:", i);
      for (int j = 0; j < 3; ++j) {
        const unsigned long long* row = hbuf.data() + (size_t)ctas[j] * MG_TRACE_SLOTS;
        fprintf(stderr, " %9lld (+%6lld)", (long long)(row[i] - hbuf[0]), i ? (long long)(row[i] - row[i - 1]) : 0LL);
      }
      fprintf(stderr, "\n");
    }
    if (p.dbg_fine != 1) {
      // skew: for every stamp, the earliest / median / latest CTA among those that took it, and who was last
      fprintf(stderr, "[mega skew] slot: CTAs / min / median / max (ns since CTA 0's first stamp), last CTA\n");
      std::vector<long long> v;
      for (int i = 0; i < MG_TRACE_SLOTS; ++i) {
        v.clear();
        int last = -1;
        long long vmax = 0;
        for (int j = 0; j < c->n_sms; ++j) {
          const unsigned long long t = hbuf[(size_t)j * MG_TRACE_SLOTS + i];
          if (!t) continue;
          const long long d = (long long)(t - hbuf[0]);
          if (last < 0 || d > vmax) { vmax = d; last = j; }
          v.push_back(d);
        }
        if (v.empty()) break;
        std::sort(v.begin(), v.end());
        fprintf(stderr, "  skew %2d: n=%3d %9lld %9lld %9lld  (spread %6lld) last cta %d\n", i, (int)v.size(), v[0], v[v.size() / 2], vmax, vmax - v[0], last);
      }
    }
    if (p.dbg_fine == 3) {
      const unsigned long long* q = hbuf.data() + (size_t)c->n_sms * (MG_TRACE_SLOTS + 8);
      fprintf(stderr, "  [sampler cta 0] ns:");
      for (int i = 1; i < 6 && q[i]; ++i) fprintf(stderr, " +%lld", (long long)(q[i] - q[i - 1]));
      fprintf(stderr, "  (entered -> loaded+penalty -> softmax -> filters -> min-p -> drawn)\n");
    }
    // whole-launch profile of the producer warp / consumer thread 0 (clock64 ticks)
    for (int j = 0; j < 3; ++j) {
      const unsigned long long* o = hbuf.data() + (size_t)c->n_sms * MG_TRACE_SLOTS + (size_t)ctas[j] * 8;
      fprintf(stderr, "  [prof cta %3d] producer: sched %llu wait_empty %llu issue %llu ticks over %llu stages | consumer: wait_full %llu of %llu ticks | fc1: total %llu wait_full %llu reduce+epilogue %llu\n",
              ctas[j], o[0], o[1], o[2], o[3], o[4], o[5], o[6], o[7] & 0xffffffffull, o[7] >> 32);
    }
    return cudaSuccess;
  }
  return cudaLaunchKernelEx(&cfg, mega_kernel(G, false, mega_packed(g, R)), p);
}


// ---------------------------------------------------------------------------------------------------------------
// batched tcgen05 path
// ---------------------------------------------------------------------------------------------------------------
bool use_bt(const zb_ctx* c, int rows) { return c->bt_ready && rows >= c->opt_bt_min_rows && !(use_mega(c, rows)); }

int bt_make_map(zb_ctx* c, CUtensorMap* m, const void* base, uint64_t inner, uint64_t rows, uint32_t box_rows) {
  cuuint64_t dims[2] = {inner, rows};
  cuuint64_t strides[1] = {inner * 2};
  cuuint32_t box[2] = {BT_BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = c->encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return c->err.fail(ZB_ERR_CUDA, "cuTensorMapEncodeTiled failed: %d (inner=%llu rows=%llu box=%u)", (int)r,
                                            (unsigned long long)inner, (unsigned long long)rows, box_rows);
  return ZB_OK;
}

// fc1 [2F][K] as {K, 16 rows, value | gate, F / 16 groups}: one box {64 k, 16, 2, 4} is a 128-row tile in which every 32-row
// quarter is 16 value rows followed by the gate rows of the same ffn columns (k_bt_gemm2's SwiGLU epilogue)
int bt_make_map_fc1_pairs(zb_ctx* c, CUtensorMap* m, const void* base, uint64_t K, uint64_t F) {
  cuuint64_t dims[4] = {K, 16, 2, F / 16};
  cuuint64_t strides[3] = {K * 2, F * K * 2, 16 * K * 2};
  cuuint32_t box[4] = {BT_BK, 16, 2, 4};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = c->encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return c->err.fail(ZB_ERR_CUDA, "cuTensorMapEncodeTiled (fc1 pairs) failed: %d", (int)r);
  return ZB_OK;
}

struct BtPlan { int n_tiles, splits, kb_per_split, num_kb; };
// split K only as far as the weight tiles x activation-row tiles leave SMs idle (a prefill block of 1 024 rows has 8 row tiles:
// no split, no partial traffic; a decode step has one)
BtPlan bt_plan(const zb_ctx* c, int n_tiles, int K, int m_tiles) {
  BtPlan pl;
  pl.n_tiles = n_tiles; pl.num_kb = K / BT_BK;
  int s = c->n_sms / (n_tiles * m_tiles);
  if (s < 1) s = 1;
  int max_s = pl.num_kb / 4;                       // at least 4 k-blocks (32 KB of weights) per CTA
  if (max_s < 1) max_s = 1;
  if (s > max_s) s = max_s;
  if (s > 8) s = 8;
  pl.kb_per_split = (pl.num_kb + s - 1) / s;
  pl.splits = (pl.num_kb + pl.kb_per_split - 1) / pl.kb_per_split;
  return pl;
}

// the activation maps again with a box of 128 / cluster-size rows: the slice one CTA of a cluster loads and multicasts
int bt_make_slices(zb_ctx* c) {
  const zb_config& g = c->cfg;
  const int rows_pad = c->bt_mpad, box = BT_BM / c->opt_bt_cluster;
  int rc;
  c->bt_cs = c->opt_bt_cluster;
  if ((rc = bt_make_map(c, &c->bt_maps_xn, c->bt_xn, g.d_model, rows_pad, box))) return rc;
  if ((rc = bt_make_map(c, &c->bt_maps_y, c->y, g.num_heads * ZB_DH, rows_pad, box))) return rc;
  if ((rc = bt_make_map(c, &c->bt_maps_h, c->h, g.ffn, rows_pad, box))) return rc;
  if (c->hybrid && c->n_mamba > 0 && (rc = bt_make_map(c, &c->bt_maps_yn, c->m_ynorm, c->m_dinner, rows_pad, box))) return rc;
  return ZB_OK;
}

int bt_setup(zb_ctx* c, int rows_pad) {
  const zb_config& g = c->cfg;
  c->bt_ready = false;
  if (!c->encode) return ZB_OK;                                                // driver without cuTensorMapEncodeTiled: per-op path only
  const int d = g.d_model, F = g.ffn, q_size = g.num_heads * ZB_DH, qkv = (g.num_heads + 2 * g.num_heads_kv) * ZB_DH;
  if (d % BT_BK || F % BT_BK || q_size % BT_BK || F % 64) return ZB_OK;
  if (d < BT_BN || qkv < BT_BN || F < 64) return ZB_OK;
  if (c->hybrid && (c->m_dinner % BT_BK || c->m_dinproj % 8 || c->m_dinproj < BT_BN)) return ZB_OK;
  c->bt_mpad = rows_pad;
  int maxN = qkv > d ? qkv : d;
  if (g.n_codebooks * g.vocab > maxN) maxN = g.n_codebooks * g.vocab;
  if (c->hybrid && c->m_dinproj > maxN) maxN = c->m_dinproj;
  ZB_CUDA(c->err, cudaMalloc((void**)&c->bt_xn, (size_t)rows_pad * d * 2));
  ZB_CUDA(c->err, cudaMemset(c->bt_xn, 0, (size_t)rows_pad * d * 2));
  c->bt_partial_elems = (size_t)8 * rows_pad * maxN;
  ZB_CUDA(c->err, cudaMalloc((void**)&c->bt_partial, c->bt_partial_elems * 4));
  int rc;
  if ((rc = bt_make_map(c, &c->bt_map_xn, c->bt_xn, d, rows_pad, BT_BM))) return rc;
  if ((rc = bt_make_map(c, &c->bt_map_y, c->y, q_size, rows_pad, BT_BM))) return rc;
  if ((rc = bt_make_map(c, &c->bt_map_h, c->h, F, rows_pad, BT_BM))) return rc;
  if (c->hybrid && c->n_mamba > 0 && (rc = bt_make_map(c, &c->bt_map_yn, c->m_ynorm, c->m_dinner, rows_pad, BT_BM))) return rc;
  if ((rc = bt_make_slices(c))) return rc;
  if ((rc = bt_make_map(c, &c->bt_map2_xn, c->bt_xn, d, rows_pad, 64))) return rc;
  if ((rc = bt_make_map(c, &c->bt_map2_y, c->y, q_size, rows_pad, 64))) return rc;
  if ((rc = bt_make_map(c, &c->bt_map2_h, c->h, F, rows_pad, 64))) return rc;
  if (c->hybrid && c->n_mamba > 0 && (rc = bt_make_map(c, &c->bt_map2_yn, c->m_ynorm, c->m_dinner, rows_pad, 64))) return rc;
  const int nl = g.n_layer;
  c->bt_map_in.resize(nl); c->bt_map_out.resize(nl); c->bt_map_fc1.resize(nl); c->bt_map_fc2.resize(nl); c->bt_map_fc1p.resize(nl);
  for (int l = 0; l < nl; ++l) {
    if (c->hybrid && c->is_mamba[l]) {                     // Mamba2 layer: in_proj [d_in_proj][d], out_proj [d][d_inner]
      const LayerM& m = c->mlayers[l];
      if ((rc = bt_make_map(c, &c->bt_map_in[l], m.w_in, d, c->m_dinproj, BT_BN))) return rc;
      if ((rc = bt_make_map(c, &c->bt_map_out[l], m.w_out, c->m_dinner, d, BT_BN))) return rc;
      continue;
    }
    const LayerW& w = c->layers[l];
    if ((rc = bt_make_map(c, &c->bt_map_in[l], w.w_in, d, qkv, BT_BN))) return rc;
    if ((rc = bt_make_map(c, &c->bt_map_out[l], w.w_out, q_size, d, BT_BN))) return rc;
    if ((rc = bt_make_map(c, &c->bt_map_fc1[l], w.w_fc1, d, 2 * F, 64))) return rc;       // value box + gate box per tile
    if ((rc = bt_make_map(c, &c->bt_map_fc2[l], w.w_fc2, F, d, BT_BN))) return rc;
    if ((rc = bt_make_map_fc1_pairs(c, &c->bt_map_fc1p[l], w.w_fc1, d, F))) return rc;
  }
  if ((rc = bt_make_map(c, &c->bt_map_heads, c->heads, d, (uint64_t)g.n_codebooks * g.vocab, BT_BN))) return rc;
  c->bt_ready = true;
  return ZB_OK;
}

cudaError_t bt_gemm(zb_ctx* c, const CUtensorMap& mx, const CUtensorMap& mw, int N, int K, int mode, int* splits_out, cudaStream_t s) {
  const int n_tiles = mode == BT_MODE_SWIGLU ? (N + 63) / 64 : (N + BT_BN - 1) / BT_BN;
  const int Mrows = c->bt_vM ? c->bt_vM : c->rows;
  BtPlan pl = bt_plan(c, n_tiles, K, (Mrows + BT_BM - 1) / BT_BM);
  if (mode == BT_MODE_SWIGLU) { pl.splits = 1; pl.kb_per_split = pl.num_kb; }
  BtGemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = Mrows; p.N = N; p.num_kb = pl.num_kb; p.kb_per_split = pl.kb_per_split; p.mode = mode;
  p.partial = c->bt_partial; p.m_pad = c->bt_mpad; p.h = c->h; p.st = c->st; p.partial_elems = c->bt_partial_elems;
  if (splits_out) *splits_out = pl.splits;
  // the activation tile is shared by the CTAs of a cluster (multicast): the slice map that goes with the full-tile map
  const int cs = c->bt_cs;
  const CUtensorMap* mxs = &mx == &c->bt_map_xn ? &c->bt_maps_xn : &mx == &c->bt_map_y ? &c->bt_maps_y : &mx == &c->bt_map_h ? &c->bt_maps_h : &c->bt_maps_yn;
  p.cs = cs;
  if (c->opt_bt_cta2 & (mode == BT_MODE_SWIGLU ? 2 : 1)) {
    const CUtensorMap* mxh = &mx == &c->bt_map_xn ? &c->bt_map2_xn : &mx == &c->bt_map_y ? &c->bt_map2_y : &mx == &c->bt_map_h ? &c->bt_map2_h : &c->bt_map2_yn;
    p.cs = 2;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((n_tiles + 1) / 2 * 2, (Mrows + BT_BM - 1) / BT_BM, pl.splits);   // a padding CTA loads zero weight rows and stores nothing
    cfg.blockDim = dim3(BT_THREADS);
    cfg.dynamicSmemBytes = (size_t)BT2_SMEM_BYTES;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (c->opt_pdl) {
      attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[na].val.programmaticStreamSerializationAllowed = 1;
      ++na;
    }
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
    ++na;
    cfg.attrs = attr;
    cfg.numAttrs = na;
    if (c->counting) c->launches += 1;
    // fc1: the map that interleaves value and gate rows, found by the layer index of the map the caller passed
    const CUtensorMap& mw2 = mode == BT_MODE_SWIGLU ? c->bt_map_fc1p[&mw - c->bt_map_fc1.data()] : mw;
    return cudaLaunchKernelEx(&cfg, k_bt_gemm2, *mxh, mw2, p);
  }
  const int gx = (n_tiles + cs - 1) / cs * cs;        // padding CTAs load out-of-range (zero) weight rows and store nothing
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(gx, (Mrows + BT_BM - 1) / BT_BM, pl.splits);
  cfg.blockDim = dim3(BT_THREADS);
  cfg.dynamicSmemBytes = (size_t)BT_SMEM_BYTES;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (c->opt_pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (cs > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = cs; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  if (c->counting) c->launches += 1;
  return cudaLaunchKernelEx(&cfg, k_bt_gemm, mx, *mxs, mw, p);
}

BtFinishParams bt_finish_params(zb_ctx* c, int splits, int N) {
  const zb_config& g = c->cfg;
  BtFinishParams f;
  memset(&f, 0, sizeof(f));
  f.partial = c->bt_partial; f.splits = splits; f.m_pad = c->bt_mpad; f.N = N;
  f.x = c->x; f.xn = c->bt_xn; f.eps = g.norm_eps; f.d = g.d_model;
  f.q = c->q; f.page_table = c->page_table; f.max_pages = c->max_pages; f.Hkv = g.num_heads_kv;
  f.q_size = g.num_heads * ZB_DH; f.kv_size = g.num_heads_kv * ZB_DH; f.rope = c->rope; f.rope_pos = c->rope_pos;
  f.logits = c->logits; f.n_q = g.n_codebooks; f.V = g.vocab; f.eos = g.eos_token_id; f.B = c->rows / 2; f.st = c->st;
  f.vrows = c->bt_vrows; f.total_pages = c->rows * c->max_pages;
  return f;
}

// LayerNorm that feeds layer l's first projection (l == n_layer: norm_f): a Mamba2 block's norm, or an attention block's norm1
static void bt_ln_of(const zb_ctx* c, int l, const bf16*& w, const bf16*& b) {
  if (l >= c->cfg.n_layer) { w = c->normf_w; b = c->normf_b; }
  else if (c->hybrid && c->is_mamba[l]) { w = c->mlayers[l].ln_w; b = c->mlayers[l].ln_b; }
  else { w = c->layers[l].ln1_w; b = c->layers[l].ln1_b; }
}

// One Mamba2 layer of the hybrid backbone on the batched path: in_proj and out_proj stream their weights ONCE for all rows
// (the per-op path re-streams them per group of 8 rows); conv / SSM step / gated norm are the decode-step kernels.
cudaError_t enqueue_bt_mamba_layer(zb_ctx* c, int l, cudaStream_t s) {
  const zb_config& g = c->cfg;
  const LayerM& w = c->mlayers[l];
  const int d = g.d_model, rows = c->rows;
  cudaError_t e;
  int splits = 1;
  if ((e = bt_gemm(c, c->bt_map_xn, c->bt_map_in[l], c->m_dinproj, d, BT_MODE_PARTIAL, &splits, s)) != cudaSuccess) return e;
  {
    BtFinishParams f = bt_finish_params(c, splits, c->m_dinproj);
    if ((e = launch(c, k_bt_plain_finish, dim3(rows, (c->m_dinproj + 2047) / 2048), dim3(256), 0, s, f, c->m_zx)) != cudaSuccess) return e;
  }
  M2Params m;
  memset(&m, 0, sizeof(m));
  m.zx = c->m_zx; m.d_in_proj = c->m_dinproj; m.d_inner = c->m_dinner; m.conv_dim = c->m_convdim; m.d_conv = c->hyb.d_conv;
  m.nheads = c->m_nheads; m.headdim = c->hyb.headdim; m.d_state = c->hyb.d_state; m.ngroups = c->hyb.ngroups;
  m.conv_w = w.conv_w; m.conv_b = w.conv_b; m.dt_bias = w.dt_bias; m.A_log = w.A_log; m.D = w.D; m.norm_w = w.norm_w;
  const int slot = c->kv_slot[l];
  m.conv_state = c->m_conv_state + (size_t)slot * rows * c->m_convdim * c->hyb.d_conv;
  m.ssm_state = c->m_ssm_state + (size_t)slot * rows * c->m_nheads * c->hyb.headdim * c->hyb.d_state;
  m.xbc = c->m_xbc; m.y = c->m_y; m.ynorm = c->m_ynorm; m.eps = 1e-5f; m.st = c->st;
  if ((e = launch(c, k_m2_conv, dim3((c->m_convdim + 255) / 256, rows), dim3(256), 0, s, m)) != cudaSuccess) return e;
  if ((e = launch(c, k_m2_ssm, dim3(c->m_nheads, rows), dim3(256), 0, s, m)) != cudaSuccess) return e;
  if ((e = launch(c, k_m2_gated_norm, dim3(rows), dim3(256), 0, s, m)) != cudaSuccess) return e;
  if ((e = bt_gemm(c, c->bt_map_yn, c->bt_map_out[l], d, c->m_dinner, BT_MODE_PARTIAL, &splits, s)) != cudaSuccess) return e;
  {
    BtFinishParams f = bt_finish_params(c, splits, d);
    bt_ln_of(c, l + 1, f.ln_w, f.ln_b);
    if ((e = launch(c, k_bt_resid_ln, dim3(rows), dim3(256), (size_t)d * 4, s, f)) != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// x holds the layer-0 input.  Runs LN1(0) + all layers; leaves norm_f(x) in bt_xn.
cudaError_t enqueue_bt_layers(zb_ctx* c, cudaStream_t s) {
  const zb_config& g = c->cfg;
  const int d = g.d_model, H = g.num_heads, Hkv = g.num_heads_kv, F = g.ffn;
  const int q_size = H * ZB_DH, qkv = (H + 2 * Hkv) * ZB_DH;
  const int rows = c->bt_vM ? c->bt_vM : c->rows;
  const size_t ln_smem = (size_t)d * 4;
  cudaError_t e;
  {
    BtFinishParams f = bt_finish_params(c, 0, d);
    f.partial = nullptr;
    bt_ln_of(c, 0, f.ln_w, f.ln_b);
    if ((e = launch(c, k_bt_resid_ln, dim3(rows), dim3(256), ln_smem, s, f)) != cudaSuccess) return e;
  }
  for (int l = 0; l < g.n_layer; ++l) {
    if (c->hybrid && c->is_mamba[l]) {
      if ((e = enqueue_bt_mamba_layer(c, l, s)) != cudaSuccess) return e;
      continue;
    }
    const LayerW& w = c->layers[l];
    bf16* kv_layer = c->kv + (size_t)c->kv_slot[l] * c->kv_layer_elems;
    int splits = 1;
    if ((e = bt_gemm(c, c->bt_map_xn, c->bt_map_in[l], qkv, d, BT_MODE_PARTIAL, &splits, s)) != cudaSuccess) return e;
    {
      BtFinishParams f = bt_finish_params(c, splits, qkv);
      f.kv_layer = kv_layer;
      if ((e = launch(c, k_bt_qkv_finish, dim3(rows, (qkv + 2047) / 2048), dim3(256), 0, s, f)) != cudaSuccess) return e;
    }
    {
      AttnParams a;
      memset(&a, 0, sizeof(a));
      a.q = c->q; a.kv_layer = kv_layer; a.page_table = c->page_table; a.max_pages = c->max_pages;
      a.H = H; a.Hkv = Hkv; a.y = c->y; a.part_o = c->part_o; a.part_ml = c->part_ml; a.counters = c->counters;
      a.st = c->st; a.scale_log2 = (1.0f / sqrtf((float)ZB_DH)) * 1.4426950408889634f; a.total_pages = c->rows * c->max_pages;
      a.vrows = c->bt_vrows;
      // block prefill: one split per (token row, kv head) - no partial buffers or tickets (they are sized for the cache rows)
      if (c->opt_attn_tma >= 2 || (c->opt_attn_tma == 1 && c->rows >= 64)) {
        if ((e = launch(c, attn_tma_kernel(H / Hkv, c->opt_attn_tma == 3), dim3(c->bt_vM ? 1 : c->nsplit, Hkv, rows), dim3(ATT_THREADS), ATT_TMA_SMEM, s, a)) != cudaSuccess) return e;
      } else
      if ((e = launch(c, attn_kernel(H / Hkv, c->opt_attn_mma != 0), dim3(c->bt_vM ? 1 : c->nsplit, Hkv, rows), dim3(ATT_THREADS), 0, s, a)) != cudaSuccess) return e;
    }
    if ((e = bt_gemm(c, c->bt_map_y, c->bt_map_out[l], d, q_size, BT_MODE_PARTIAL, &splits, s)) != cudaSuccess) return e;
    {
      BtFinishParams f = bt_finish_params(c, splits, d);
      f.ln_w = w.ln2_w; f.ln_b = w.ln2_b;
      if ((e = launch(c, k_bt_resid_ln, dim3(rows), dim3(256), ln_smem, s, f)) != cudaSuccess) return e;
    }
    if ((e = bt_gemm(c, c->bt_map_xn, c->bt_map_fc1[l], F, d, BT_MODE_SWIGLU, nullptr, s)) != cudaSuccess) return e;
    if ((e = bt_gemm(c, c->bt_map_h, c->bt_map_fc2[l], d, F, BT_MODE_PARTIAL, &splits, s)) != cudaSuccess) return e;
    {
      BtFinishParams f = bt_finish_params(c, splits, d);
      bt_ln_of(c, l + 1, f.ln_w, f.ln_b);
      if ((e = launch(c, k_bt_resid_ln, dim3(rows), dim3(256), ln_smem, s, f)) != cudaSuccess) return e;
    }
  }
  return cudaSuccess;
}

cudaError_t enqueue_bt_heads(zb_ctx* c, float* logits, cudaStream_t s) {
  const zb_config& g = c->cfg;
  const int NH = g.n_codebooks * g.vocab;
  int splits = 1;
  cudaError_t e;
  if ((e = bt_gemm(c, c->bt_map_xn, c->bt_map_heads, NH, g.d_model, BT_MODE_PARTIAL, &splits, s)) != cudaSuccess) return e;
  BtFinishParams f = bt_finish_params(c, splits, NH);
  f.logits = logits;
  return launch(c, k_bt_heads_finish, dim3(g.n_codebooks, c->rows / 2), dim3(256), 0, s, f);
}

// ---------------------------------------------------------------------------------------------------------------
// hybrid backbone: layer-major prefill of S tokens.  Mamba2 layers run in their SEQUENCE form (in_proj over all token
// rows, depthwise conv over the sequence, chunked SSD scan, gated norm, out_proj); the attention layers (5 of 46 in the
// benchmarked stack) are stepped token by token inside their layer.  Reference: mamba_ssm Mamba2.forward as wired by
// zonos/backbone/_mamba_ssm.py:15-57.  Activations are token-major: row a = t * rows + r.
// ---------------------------------------------------------------------------------------------------------------
int ensure_prefill_ws(zb_ctx* c, int M) {
  if (M <= c->pf_M) return ZB_OK;
  ZB_CUDA(c->err, cudaDeviceSynchronize());
  for (void* q : {(void*)c->pf_x, (void*)c->pf_zx, (void*)c->pf_y, (void*)c->pf_yn, (void*)c->pf_xbc}) if (q) cudaFree(q);
  c->pf_x = c->pf_zx = c->pf_y = c->pf_yn = nullptr; c->pf_xbc = nullptr; c->pf_M = 0;
  const size_t Mp = ((size_t)M + 7) / 8 * 8;
  ZB_CUDA(c->err, cudaMalloc((void**)&c->pf_x, Mp * c->cfg.d_model * 2));
  ZB_CUDA(c->err, cudaMalloc((void**)&c->pf_zx, Mp * c->m_dinproj * 2));
  ZB_CUDA(c->err, cudaMalloc((void**)&c->pf_y, Mp * c->m_dinner * 2));
  ZB_CUDA(c->err, cudaMalloc((void**)&c->pf_yn, Mp * c->m_dinner * 2));
  ZB_CUDA(c->err, cudaMalloc((void**)&c->pf_xbc, Mp * c->m_convdim * 4));
  ZB_CUDA(c->err, cudaMemset(c->pf_yn, 0, Mp * c->m_dinner * 2));
  c->pf_M = M;
  return ZB_OK;
}

int hybrid_prefill(zb_ctx* c, const bf16* d_in, bf16* d_out, int S, int pos0, const int* d_lengths, cudaStream_t s) {
  const zb_config& g = c->cfg;
  const int rows = c->rows, d = g.d_model, M = rows * S;
  int rc = ensure_prefill_ws(c, M);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    ZB_CUDA(c->err, cudaFuncSetAttribute(k_m2_ssd_chunked, cudaFuncAttributeMaxDynamicSharedMemorySize, M2_SSD_SMEM));
    attr_set = true;
  }
  const bool pdl = c->opt_pdl;
  c->opt_pdl = 0;                                   // plain stream order between the sequence kernels
  auto restore = [&](int r) { c->opt_pdl = pdl; return r; };
  ZB_CUDA(c->err, launch(c, k_set_pos, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, rows, pos0, d_lengths, 0));
  k_m2_gather_tm<<<M, 256, 0, s>>>(d_in, c->pf_x, rows, S, d);
  const std::vector<Pass> all = plain_passes(M), tok = plain_passes(rows);
  for (int l = 0; l < g.n_layer; ++l) {
    cudaError_t e = cudaSuccess;
    if (c->is_mamba[l]) {
      const LayerM& w = c->mlayers[l];
      for (const Pass& ps : all) {                  // LN + in_proj over all token rows
        GemvParams p = base_params(c, ps);
        p.W = w.w_in; p.K = d; p.n_tasks = c->m_dinproj / 2; p.row0_mul = 2; p.row1_off = 1;
        p.x = c->pf_x; p.ln_w = w.ln_w; p.ln_b = w.ln_b; p.out = c->pf_zx; p.out_stride = c->m_dinproj;
        if ((e = launch_gemv(c, EPI_PLAIN, ps, p, s)) != cudaSuccess) break;
      }
      if (e != cudaSuccess) return restore(c->err.fail(ZB_ERR_CUDA, "hybrid prefill in_proj: %s", cudaGetErrorString(e)));
      M2SeqParams q;
      memset(&q, 0, sizeof(q));
      q.zx = c->pf_zx; q.rows = rows; q.S = S; q.d_in_proj = c->m_dinproj; q.d_inner = c->m_dinner; q.conv_dim = c->m_convdim;
      q.d_conv = c->hyb.d_conv; q.nheads = c->m_nheads; q.headdim = c->hyb.headdim; q.d_state = c->hyb.d_state; q.ngroups = c->hyb.ngroups;
      q.conv_w = w.conv_w; q.conv_b = w.conv_b; q.dt_bias = w.dt_bias; q.A_log = w.A_log; q.D = w.D;
      const int slot = c->kv_slot[l];
      q.conv_state = c->m_conv_state + (size_t)slot * rows * c->m_convdim * c->hyb.d_conv;
      q.ssm_state = c->m_ssm_state + (size_t)slot * rows * c->m_nheads * c->hyb.headdim * c->hyb.d_state;
      q.xbc = c->pf_xbc; q.y = c->pf_y;
      k_m2_conv_seq<<<dim3((c->m_convdim + 255) / 256, rows), 256, 0, s>>>(q);
      k_m2_ssd_chunked<<<dim3(c->m_nheads, rows), 256, M2_SSD_SMEM, s>>>(q);
      M2Params m;
      memset(&m, 0, sizeof(m));
      m.zx = c->pf_zx; m.d_in_proj = c->m_dinproj; m.d_inner = c->m_dinner; m.y = c->pf_y; m.ynorm = c->pf_yn; m.norm_w = w.norm_w;
      m.eps = 1e-5f; m.st = c->st;
      if ((e = launch(c, k_m2_gated_norm, dim3(M), dim3(256), 0, s, m)) != cudaSuccess)
        return restore(c->err.fail(ZB_ERR_CUDA, "hybrid prefill gated norm: %s", cudaGetErrorString(e)));
      c->launches += 3;
      for (const Pass& ps : all) {                  // out_proj + residual
        GemvParams p = base_params(c, ps);
        p.W = w.w_out; p.K = c->m_dinner; p.n_tasks = d / 2; p.row0_mul = 2; p.row1_off = 1;
        p.x = c->pf_yn; p.out = c->pf_x; p.out_stride = d;
        if ((e = launch_gemv(c, EPI_RESID, ps, p, s)) != cudaSuccess) break;
      }
      if (e != cudaSuccess) return restore(c->err.fail(ZB_ERR_CUDA, "hybrid prefill out_proj: %s", cudaGetErrorString(e)));
    } else {
      for (int t = 0; t < S; ++t) {                 // attention + GatedMLP layer: token by token (KV position = pos0 + t)
        if ((e = launch(c, k_set_pos, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, rows, pos0 + t, d_lengths, t)) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(c->x, c->pf_x + (size_t)t * rows * d, (size_t)rows * d * 2, cudaMemcpyDeviceToDevice, s)) != cudaSuccess) break;
        if ((e = enqueue_attn_layer(c, l, tok, s, false)) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(c->pf_x + (size_t)t * rows * d, c->x, (size_t)rows * d * 2, cudaMemcpyDeviceToDevice, s)) != cudaSuccess) break;
      }
      if (e != cudaSuccess) return restore(c->err.fail(ZB_ERR_CUDA, "hybrid prefill attention layer: %s", cudaGetErrorString(e)));
    }
  }
  if (d_out) {
    for (int t = 0; t < S; ++t)
      ZB_CUDA(c->err, launch(c, k_final_ln, dim3(rows), dim3(256), 0, s, (const bf16*)(c->pf_x + (size_t)t * rows * d), (const bf16*)c->normf_w,
                             (const bf16*)c->normf_b, g.norm_eps, d_out, d, S, t));
  }
  // leave the device state positioned after the appended tokens
  ZB_CUDA(c->err, launch(c, k_set_pos, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, rows, pos0 + S, d_lengths, S));
  ZB_CUDA(c->err, cudaGetLastError());
  return restore(ZB_OK);
}

// ---------------------------------------------------------------------------------------------------------------
// transformer: prefill of S tokens as BLOCKS of extra activation rows on the batched tcgen05 path.  A block of
// Sc = rows_pad / rows tokens is M = Sc * rows activation rows (row a = token a / rows of cache row a % rows): the GEMMs
// stream every weight once per block instead of once per token, k_bt_qkv_finish appends the K/V of all the block's tokens
// at their own positions, and the decode attention kernel run with per-row context lengths IS causal attention over the
// block.  Reference: Zonos._prefill -> one batched forward (model.py:181-196).  25 tokens: 27 ms -> ~1 ms.
// ---------------------------------------------------------------------------------------------------------------
int bt_prefill(zb_ctx* c, const bf16* d_in, bf16* d_out, int S, int pos0, const int* d_lengths, cudaStream_t s) {
  const zb_config& g = c->cfg;
  const int rows = c->rows, d = g.d_model;
  const int Sc = c->bt_mpad / rows;                       // tokens per block
  int thr = d / 8 < 1024 ? d / 8 : 1024;
  for (int t0 = 0; t0 < S; t0 += Sc) {
    const int n = S - t0 < Sc ? S - t0 : Sc, M = n * rows;
    ZB_CUDA(c->err, launch(c, k_set_pos, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, rows, pos0 + t0, d_lengths, t0));
    k_m2_gather_tm<<<M, thr, 0, s>>>(d_in + (size_t)t0 * d, c->x, rows, S, d);      // x[t * rows + r] <- in[r][t0 + t]
    c->bt_vM = M; c->bt_vrows = rows;
    cudaError_t e = enqueue_bt_layers(c, s);             // leaves norm_f(x) in bt_xn
    c->bt_vM = 0; c->bt_vrows = 0;
    ZB_CUDA(c->err, e);
    if (d_out) k_bt_scatter_tm<<<M, thr, 0, s>>>((const bf16*)c->bt_xn, d_out + (size_t)t0 * d, rows, S, d);
  }
  ZB_CUDA(c->err, launch(c, k_set_pos, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, rows, pos0 + S, d_lengths, S));
  ZB_CUDA(c->err, cudaGetLastError());
  return ZB_OK;
}

void drop_graph(zb_ctx* c) {
  if (c->graph) { cudaGraphExecDestroy(c->graph); c->graph = nullptr; }
  c->graph_nodes = 0;
}

void free_kv(zb_ctx* c) {
  drop_graph(c);
  void* ptrs[] = {c->kv, c->page_table, c->x, c->q, c->y, c->h, c->logits, c->part_o, c->part_ml, c->counters,
                  c->tokens, c->remaining, c->stopping, c->out_len, c->rope_pos};
  for (void* p : ptrs) if (p) cudaFree(p);
  c->kv = nullptr; c->page_table = nullptr; c->x = c->q = c->y = c->h = nullptr; c->logits = c->part_o = c->part_ml = nullptr;
  c->counters = nullptr; c->tokens = nullptr; c->remaining = c->stopping = c->out_len = c->rope_pos = nullptr;
  void* mptrs[] = {c->m_conv_state, c->m_ssm_state, c->m_xbc, c->m_zx, c->m_y, c->m_ynorm};
  for (void* q : mptrs) if (q) cudaFree(q);
  c->m_conv_state = c->m_ssm_state = c->m_xbc = nullptr; c->m_zx = c->m_y = c->m_ynorm = nullptr;
  if (c->bt_xn) cudaFree(c->bt_xn);
  if (c->bt_partial) cudaFree(c->bt_partial);
  c->bt_xn = nullptr; c->bt_partial = nullptr; c->bt_ready = false;
  c->rows = 0;
}

}  // namespace

// =============================================================================================
extern "C" {

int zb_abi_version(void) { return ZB_ABI_VERSION; }

const char* zb_last_error(const zb_ctx* ctx) { return ctx ? ctx->err.msg.c_str() : g_create_error.c_str(); }

static int ctx_create_impl(const zb_config* cfg, const zb_hybrid* hyb, int device, zb_ctx** out);
int zb_ctx_create(const zb_config* cfg, int device, zb_ctx** out) { return ctx_create_impl(cfg, nullptr, device, out); }
int zb_ctx_create_hybrid(const zb_config* cfg, const zb_hybrid* hyb, int device, zb_ctx** out) {
  if (!hyb) { g_create_error = "null hybrid config"; return ZB_ERR_ARG; }
  return ctx_create_impl(cfg, hyb, device, out);
}

static int ctx_create_impl(const zb_config* cfg, const zb_hybrid* hyb, int device, zb_ctx** out) {
  ZbErr tmp;
  auto bad = [&](int code, const std::string& m) { g_create_error = m; return code; };
  if (!cfg || !out) return bad(ZB_ERR_ARG, "null argument");
  if (hyb) {
    if (hyb->d_state != 128 || hyb->d_conv < 2 || hyb->d_conv > 8 || hyb->ngroups != 1 || hyb->headdim <= 0 || hyb->expand <= 0)
      return bad(ZB_ERR_ARG, "hybrid: d_state must be 128, ngroups 1, d_conv in [2,8]");
    if ((hyb->expand * cfg->d_model) % hyb->headdim != 0 || (hyb->expand * cfg->d_model) % 256 != 0)
      return bad(ZB_ERR_ARG, "hybrid: d_inner must be a multiple of headdim and of 256");
    if (hyb->expand * cfg->d_model > 8192) return bad(ZB_ERR_ARG, "hybrid: d_inner above 8192 is not supported (k_m2_gated_norm keeps a row in registers)");
    if (hyb->n_attn_layers < 0 || hyb->n_attn_layers > 64) return bad(ZB_ERR_ARG, "hybrid: at most 64 attention layers");
    const int dinp = 2 * hyb->expand * cfg->d_model + 2 * hyb->d_state + hyb->expand * cfg->d_model / hyb->headdim;
    if (dinp % 8 != 0) return bad(ZB_ERR_ARG, "hybrid: in_proj rows must be a multiple of 8");
  }
  if (cfg->d_model <= 0 || cfg->d_model % 256 != 0 || cfg->d_model > GEMV_MAX_LN_DIM)
    return bad(ZB_ERR_ARG, "d_model must be a multiple of 256 and <= 2048");
  if (cfg->num_heads <= 0 || cfg->d_model != cfg->num_heads * ZB_DH) return bad(ZB_ERR_ARG, "head_dim must be 128");
  if (cfg->num_heads_kv <= 0 || cfg->num_heads % cfg->num_heads_kv != 0) return bad(ZB_ERR_ARG, "num_heads % num_heads_kv != 0");
  const int G = cfg->num_heads / cfg->num_heads_kv;
  if (!(G == 1 || G == 2 || G == 4 || G == 8)) return bad(ZB_ERR_ARG, "GQA group size must be 1, 2, 4 or 8");
  if (cfg->ffn <= 0 || cfg->ffn % 256 != 0) return bad(ZB_ERR_ARG, "ffn must be a multiple of 256");
  if (cfg->n_codebooks <= 0 || cfg->n_codebooks > 16 || cfg->vocab % 2 != 0 || cfg->vocab > SMP_MAXV)
    return bad(ZB_ERR_ARG, "n_codebooks in [1,16], vocab even and <= 2048");
  if (gemv_smem(8, cfg->ffn) > 227 * 1024) return bad(ZB_ERR_ARG, "ffn too large for the staged-activation GEMV");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= device)
    return bad(ZB_ERR_CUDA, std::string("no CUDA device ") + std::to_string(device) + ": " + cudaGetErrorString(e) +
                                " (libzonos_b200 has no CPU fallback)");
  if ((e = cudaSetDevice(device)) != cudaSuccess) return bad(ZB_ERR_CUDA, cudaGetErrorString(e));
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  if (prop.major != 10) return bad(ZB_ERR_CUDA, "libzonos_b200 is built for sm_100a only");

  zb_ctx* c = new zb_ctx();
  c->cfg = *cfg;
  c->device = device;
  const int d = cfg->d_model, F = cfg->ffn, nl = cfg->n_layer;
  const int qkv = (cfg->num_heads + 2 * cfg->num_heads_kv) * ZB_DH;
  // ---- carve the weight arena (layer-major so the stream order == address order)
  size_t off = 0;
  auto carve = [&](const std::string& name, int64_t numel, int dtype, int64_t alt = 0) {
    Slot s;
    s.numel = numel; s.dtype = dtype; s.alt_numel = alt;
    s.ptr = reinterpret_cast<void*>(off);
    off = align_up(off + (size_t)numel * (dtype == ZB_F32 ? 4 : 2), 256);
    c->slots[name] = s;
  };
  c->is_mamba.assign(nl, 0);
  c->kv_slot.assign(nl, 0);
  if (hyb) {
    c->hybrid = true;
    c->hyb = *hyb;
    c->m_dinner = hyb->expand * d;
    c->m_nheads = c->m_dinner / hyb->headdim;
    c->m_convdim = c->m_dinner + 2 * hyb->ngroups * hyb->d_state;
    c->m_dinproj = 2 * c->m_dinner + 2 * hyb->ngroups * hyb->d_state + c->m_nheads;
    for (int l = 0; l < nl; ++l) c->is_mamba[l] = 1;
    for (int i = 0; i < hyb->n_attn_layers; ++i)
      if (hyb->attn_layer_idx[i] >= 0 && hyb->attn_layer_idx[i] < nl) c->is_mamba[hyb->attn_layer_idx[i]] = 0;
  }
  for (int l = 0; l < nl; ++l) c->kv_slot[l] = c->is_mamba[l] ? c->n_mamba++ : c->n_attn++;
  for (int l = 0; l < nl; ++l) {
    const std::string p = "backbone.layers." + std::to_string(l);
    carve(p + ".norm.weight", d, ZB_BF16);
    carve(p + ".norm.bias", d, ZB_BF16);
    if (c->is_mamba[l]) {
      carve(p + ".mixer.in_proj.weight", (int64_t)c->m_dinproj * d, ZB_BF16);
      carve(p + ".mixer.conv1d.weight", (int64_t)c->m_convdim * hyb->d_conv, ZB_BF16);
      carve(p + ".mixer.conv1d.bias", c->m_convdim, ZB_BF16);
      carve(p + ".mixer.dt_bias", c->m_nheads, ZB_F32);
      carve(p + ".mixer.A_log", c->m_nheads, ZB_F32);
      carve(p + ".mixer.D", c->m_nheads, ZB_F32);
      carve(p + ".mixer.norm.weight", c->m_dinner, ZB_BF16);
      carve(p + ".mixer.out_proj.weight", (int64_t)d * c->m_dinner, ZB_BF16);
      continue;
    }
    carve(p + ".mixer.in_proj.weight", (int64_t)qkv * d, ZB_BF16);
    carve(p + ".mixer.out_proj.weight", (int64_t)d * cfg->num_heads * ZB_DH, ZB_BF16);
    carve(p + ".norm2.weight", d, ZB_BF16);
    carve(p + ".norm2.bias", d, ZB_BF16);
    carve(p + ".mlp.fc1.weight", (int64_t)2 * F * d, ZB_BF16);
    carve(p + ".mlp.fc2.weight", (int64_t)d * F, ZB_BF16);
  }
  carve("backbone.norm_f.weight", d, ZB_BF16);
  carve("backbone.norm_f.bias", d, ZB_BF16);
  for (int k = 0; k < cfg->n_codebooks; ++k) carve("heads." + std::to_string(k) + ".weight", (int64_t)cfg->vocab * d, ZB_BF16, (int64_t)(cfg->vocab - 1) * d);
  for (int k = 0; k < cfg->n_codebooks; ++k) carve("embeddings." + std::to_string(k) + ".weight", (int64_t)cfg->vocab * d, ZB_BF16);
  carve("rope.freqs_cis", (int64_t)cfg->rope_table_len * (ZB_DH / 2) * 2, ZB_F32);
  c->arena_bytes = off;
  if ((e = cudaMalloc(&c->arena, c->arena_bytes)) != cudaSuccess) {
    delete c;
    return bad(ZB_ERR_NOMEM, std::string("weight arena: ") + cudaGetErrorString(e));
  }
  cudaMemset(c->arena, 0, c->arena_bytes);
  for (auto& kv : c->slots) kv.second.ptr = c->arena + reinterpret_cast<size_t>(kv.second.ptr);
  auto W = [&](const std::string& n) { return reinterpret_cast<bf16*>(c->slots[n].ptr); };
  c->layers.resize(nl);
  c->mlayers.resize(nl);
  auto Wf = [&](const std::string& n) { return reinterpret_cast<float*>(c->slots[n].ptr); };
  for (int l = 0; l < nl; ++l) {
    const std::string p = "backbone.layers." + std::to_string(l);
    if (c->is_mamba[l]) {
      c->layers[l] = LayerW{};
      c->mlayers[l] = LayerM{W(p + ".norm.weight"), W(p + ".norm.bias"), W(p + ".mixer.in_proj.weight"), W(p + ".mixer.conv1d.weight"),
                             W(p + ".mixer.conv1d.bias"), W(p + ".mixer.norm.weight"), W(p + ".mixer.out_proj.weight"),
                             Wf(p + ".mixer.dt_bias"), Wf(p + ".mixer.A_log"), Wf(p + ".mixer.D")};
      continue;
    }
    c->layers[l] = LayerW{W(p + ".norm.weight"), W(p + ".norm.bias"), W(p + ".mixer.in_proj.weight"), W(p + ".mixer.out_proj.weight"),
                          W(p + ".norm2.weight"), W(p + ".norm2.bias"), W(p + ".mlp.fc1.weight"), W(p + ".mlp.fc2.weight")};
  }
  c->normf_w = W("backbone.norm_f.weight");
  c->normf_b = W("backbone.norm_f.bias");
  c->heads = W("heads.0.weight");        // heads are contiguous: [n_q][vocab][d] (vocab*d*2 is a multiple of 256)
  c->emb = W("embeddings.0.weight");
  c->rope = reinterpret_cast<float*>(c->slots["rope.freqs_cis"].ptr);
  if (((size_t)cfg->vocab * d * 2) % 256 != 0) { cudaFree(c->arena); delete c; return bad(ZB_ERR_ARG, "vocab*d_model*2 must be a multiple of 256"); }
  // algorithmic bytes streamed per decode step: layer weights + LN params + final LN + real head rows
  c->step_weight_bytes = (int64_t)c->n_attn * ((int64_t)qkv * d + (int64_t)d * d + (int64_t)2 * F * d + (int64_t)d * F + 4 * d) * 2 +
                         (int64_t)2 * d * 2 + (int64_t)cfg->n_codebooks * (cfg->vocab - 1) * d * 2;
  if (hyb)   // Mamba2 layer: LN + in_proj + conv + per-head scalars + gated-norm weight + out_proj
    c->step_weight_bytes += (int64_t)c->n_mamba * (((int64_t)c->m_dinproj * d + (int64_t)c->m_convdim * (hyb->d_conv + 1) + c->m_dinner +
                                                    (int64_t)d * c->m_dinner + 2 * d) * 2 + (int64_t)3 * c->m_nheads * 4);

  // opt-in to large dynamic shared memory for every GEMV instantiation.  The attribute is per FUNCTION (shared by all
  // contexts of the process), so it is always raised to the device maximum: a context for a small config created
  // after a large one must not shrink it.
  const int max_smem = (int)prop.sharedMemPerBlockOptin;
  for (int R = 2; R <= 8; R += 2)
    for (int epi = 0; epi < 5; ++epi) {
      if ((epi < 4 && (e = cudaFuncSetAttribute(gemv_kernel(R, epi), cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem)) != cudaSuccess) ||
          (e = cudaFuncSetAttribute(gemv_mma_kernel(R, epi), cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem)) != cudaSuccess) {
        cudaFree(c->arena); delete c;
        return bad(ZB_ERR_CUDA, std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
      }
    }
  c->n_sms = prop.multiProcessorCount;
  c->smem_optin = max_smem;
  for (int G2 = 1; G2 <= 8; G2 *= 2)
    if ((e = cudaFuncSetAttribute(attn_tma_kernel(G2, false), cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_TMA_SMEM)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(attn_tma_kernel(G2, true), cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_TMA_SMEM)) != cudaSuccess) {
      cudaFree(c->arena); delete c;
      return bad(ZB_ERR_CUDA, std::string("cudaFuncSetAttribute(attn): ") + cudaGetErrorString(e));
    }
  for (int G2 = 1; G2 <= 8; G2 *= 2)
    if ((e = cudaFuncSetAttribute(mega_kernel(G2, false, false), cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(mega_kernel(G2, false, true), cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(mega_kernel(G2, true, false), cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(mega_kernel(G2, true, true), cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem)) != cudaSuccess) {
      cudaFree(c->arena); delete c;
      return bad(ZB_ERR_CUDA, std::string("cudaFuncSetAttribute(mega): ") + cudaGetErrorString(e));
    }
  {
    // the persistent kernel addresses layer l as layer 0 + l * stride: check that the arena really is layer-periodic
    const LayerW& w0 = c->layers[0];
    c->mega_l0 = MegaLayer{w0.ln1_w, w0.ln1_b, w0.w_in, w0.w_out, w0.ln2_w, w0.ln2_b, w0.w_fc1, w0.w_fc2};
    c->mega_layer_stride = nl > 1 ? (long long)((const char*)c->layers[1].ln1_w - (const char*)w0.ln1_w) : 0;
    c->mega_ok = !c->hybrid;
    for (int l = 0; l < nl; ++l) {
      const LayerW& w = c->layers[l];
      const bf16* a[8] = {w.ln1_w, w.ln1_b, w.w_in, w.w_out, w.ln2_w, w.ln2_b, w.w_fc1, w.w_fc2};
      const bf16* b[8] = {w0.ln1_w, w0.ln1_b, w0.w_in, w0.w_out, w0.ln2_w, w0.ln2_b, w0.w_fc1, w0.w_fc2};
      for (int i = 0; i < 8; ++i)
        if ((const char*)a[i] != (const char*)b[i] + (long long)l * c->mega_layer_stride) c->mega_ok = false;
    }
    if ((e = cudaMalloc(&c->bar_counter, 256)) != cudaSuccess) {
      cudaFree(c->arena); delete c;
      return bad(ZB_ERR_NOMEM, std::string("barrier counter: ") + cudaGetErrorString(e));
    }
    // every epoch-tagged buffer of the persistent kernel in ONE allocation (zeroed at the start of every launch):
    // [x_ll 2 x 8 x d/2][qkv_ll 8 x (q + 2 kv)/2][y_ll 8 x q/2][h_ll 8 x F/2][att_ll 8 x Hkv x MAX_SPLIT x G (Dh + 2)] words
    const size_t w_x = (size_t)2 * 8 * (cfg->d_model / 2), w_qkv = (size_t)8 * ((cfg->num_heads + 2 * cfg->num_heads_kv) * ZB_DH / 2);
    const size_t w_y = (size_t)8 * (cfg->num_heads * ZB_DH / 2), w_h = (size_t)8 * (cfg->ffn / 2);
    const size_t w_att = (size_t)8 * cfg->num_heads_kv * MG_MAX_SPLIT * (cfg->num_heads / cfg->num_heads_kv) * (ZB_DH + 2);
    c->ll_bytes = (w_x + w_qkv + w_y + w_h + w_att) * sizeof(unsigned long long);
    if ((e = cudaMalloc(&c->x_ll, c->ll_bytes)) != cudaSuccess || (e = cudaMemset(c->x_ll, 0, c->ll_bytes)) != cudaSuccess) {
      cudaFree(c->arena); delete c;
      return bad(ZB_ERR_NOMEM, std::string("x_ll: ") + cudaGetErrorString(e));
    }
    c->qkv_ll = c->x_ll + w_x; c->y_ll = c->qkv_ll + w_qkv; c->h_ll = c->y_ll + w_y; c->att_ll = c->h_ll + w_h;
  }
  if ((e = cudaFuncSetAttribute(k_bt_gemm, cudaFuncAttributeMaxDynamicSharedMemorySize, BT_SMEM_BYTES)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(k_bt_gemm2, cudaFuncAttributeMaxDynamicSharedMemorySize, BT2_SMEM_BYTES)) != cudaSuccess) {
    cudaFree(c->arena); delete c;
    return bad(ZB_ERR_CUDA, std::string("cudaFuncSetAttribute(bt): ") + cudaGetErrorString(e));
  }
  {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
      c->encode = reinterpret_cast<zb_ctx::EncodeTiledFn>(fn);
  }
  if ((e = cudaMalloc(&c->st, sizeof(ZbDevState))) != cudaSuccess) { cudaFree(c->arena); delete c; return bad(ZB_ERR_NOMEM, cudaGetErrorString(e)); }
  cudaMemset(c->st, 0, sizeof(ZbDevState));
  cudaEventCreate(&c->ev0);
  cudaEventCreate(&c->ev1);
  for (auto& g : c->gen) {
    cudaEventCreate(&g.e0);
    cudaEventCreate(&g.e1);
    cudaEventCreateWithFlags(&g.e2, cudaEventDisableTiming);
    if (cudaMallocHost((void**)&g.h_st, sizeof(ZbDevState)) != cudaSuccess || cudaMallocHost((void**)&g.h_len, zb_ctx::kGenMaxB * 4) != cudaSuccess) {
      zb_ctx_destroy(c);
      return bad(ZB_ERR_NOMEM, "pinned host memory for the generation state snapshots");
    }
  }
  cudaStreamCreateWithFlags(&c->cap_stream, cudaStreamNonBlocking);
  *out = c;
  return ZB_OK;
}

int zb_ctx_destroy(zb_ctx* c) {
  if (!c) return ZB_OK;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  free_kv(c);
  if (c->arena) cudaFree(c->arena);
  if (c->st) cudaFree(c->st);
  if (c->bar_counter) cudaFree(c->bar_counter);
  if (c->x_ll) cudaFree(c->x_ll);
  if (c->pf_embed) cudaFree(c->pf_embed);
  for (void* q : {(void*)c->pf_x, (void*)c->pf_zx, (void*)c->pf_y, (void*)c->pf_yn, (void*)c->pf_xbc}) if (q) cudaFree(q);
  if (c->dbg) cudaFree(c->dbg);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  for (auto& g : c->gen) {
    if (g.e0) cudaEventDestroy(g.e0);
    if (g.e1) cudaEventDestroy(g.e1);
    if (g.e2) cudaEventDestroy(g.e2);
    if (g.h_st) cudaFreeHost(g.h_st);
    if (g.h_len) cudaFreeHost(g.h_len);
  }
  if (c->cap_stream) cudaStreamDestroy(c->cap_stream);
  delete c;
  return ZB_OK;
}

int zb_set_tensor(zb_ctx* c, const char* name, const void* d_ptr, int64_t numel, int dtype, void* stream) {
  if (!c || !name || !d_ptr) return ZB_ERR_ARG;
  auto it = c->slots.find(name);
  if (it == c->slots.end()) return c->err.fail(ZB_ERR_ARG, "unknown tensor '%s'", name);
  Slot& s = it->second;
  if (dtype != s.dtype) return c->err.fail(ZB_ERR_ARG, "tensor '%s': dtype %d, expected %d", name, dtype, s.dtype);
  if (numel != s.numel && !(s.alt_numel && numel == s.alt_numel))
    return c->err.fail(ZB_ERR_ARG, "tensor '%s': numel %lld, expected %lld", name, (long long)numel, (long long)s.numel);
  const size_t esz = s.dtype == ZB_F32 ? 4 : 2;
  cudaStream_t st = (cudaStream_t)stream;
  ZB_CUDA(c->err, cudaSetDevice(c->device));
  if (numel != s.numel) ZB_CUDA(c->err, cudaMemsetAsync(s.ptr, 0, (size_t)s.numel * esz, st));   // zero pad row (utils.py:12-27)
  ZB_CUDA(c->err, cudaMemcpyAsync(s.ptr, d_ptr, (size_t)numel * esz, cudaMemcpyDeviceToDevice, st));
  s.provided = true;
  return ZB_OK;
}

int zb_weights_finalize(zb_ctx* c, void* stream) {
  if (!c) return ZB_ERR_ARG;
  for (auto& kv : c->slots)
    if (!kv.second.provided) return c->err.fail(ZB_ERR_STATE, "tensor '%s' was not provided", kv.first.c_str());
  ZB_CUDA(c->err, cudaStreamSynchronize((cudaStream_t)stream));
  c->finalized = true;
  return ZB_OK;
}

int zb_kv_alloc(zb_ctx* c, int rows, int max_seqlen) {
  if (!c || rows <= 0 || max_seqlen <= 0) return ZB_ERR_ARG;
  ZB_CUDA(c->err, cudaSetDevice(c->device));
  if (max_seqlen > c->cfg.rope_table_len) return c->err.fail(ZB_ERR_ARG, "max_seqlen %d exceeds the RoPE table (%d)", max_seqlen, c->cfg.rope_table_len);
  const zb_config& g = c->cfg;
  const int pages_per_row = (max_seqlen + ZB_PAGE - 1) / ZB_PAGE;
  if (c->kv && c->rows == rows && c->max_pages == pages_per_row) { c->max_seqlen = max_seqlen; return ZB_OK; }
  ZB_CUDA(c->err, cudaDeviceSynchronize());
  free_kv(c);
  c->rows = rows; c->max_pages = pages_per_row; c->max_seqlen = max_seqlen;
  const size_t total_pages = (size_t)rows * pages_per_row;
  c->kv_layer_elems = total_pages * 2 * g.num_heads_kv * ZB_PAGE * ZB_DH;
  int ns = (4 * ZB_NUM_SMS) / (rows * g.num_heads_kv);
  ns = ns < 1 ? 1 : (ns > 32 ? 32 : ns);
  if (c->opt_nsplit > 0) ns = c->opt_nsplit > 32 ? 32 : c->opt_nsplit;
  c->nsplit = ns;
  const int G = g.num_heads / g.num_heads_kv;
  int rows_pad = (rows + 127) / 128 * 128;             // activation matrices are TMA operands of 128-row tiles (bt path)
  if (!c->hybrid) {
    // block prefill (bt_prefill) turns rows_pad / rows token positions into one pass over the weights: room for 64 positions,
    // at most 1 024 activation rows (8 M tiles per weight tile; batch 64: 8 positions per pass instead of 1)
    int want = rows * 64 > 1024 ? 1024 : rows * 64;
    want = (want + 127) / 128 * 128;
    if (want > rows_pad) rows_pad = want;
  }
#define ZB_ALLOC(ptr, bytes) ZB_CUDA(c->err, cudaMalloc((void**)&(ptr), (bytes)))
  const size_t n_kv_layers = c->n_attn > 0 ? c->n_attn : 1;
  ZB_ALLOC(c->kv, c->kv_layer_elems * n_kv_layers * sizeof(bf16));
  if (c->hybrid && c->n_mamba > 0) {
    const size_t cs = (size_t)c->n_mamba * rows * c->m_convdim * c->hyb.d_conv, ss = (size_t)c->n_mamba * rows * c->m_nheads * c->hyb.headdim * c->hyb.d_state;
    ZB_ALLOC(c->m_conv_state, cs * 4);
    ZB_ALLOC(c->m_ssm_state, ss * 4);
    ZB_ALLOC(c->m_xbc, (size_t)rows_pad * c->m_convdim * 4);
    ZB_ALLOC(c->m_zx, (size_t)rows_pad * c->m_dinproj * 2);
    ZB_ALLOC(c->m_y, (size_t)rows_pad * c->m_dinner * 2);
    ZB_ALLOC(c->m_ynorm, (size_t)rows_pad * c->m_dinner * 2);
    ZB_CUDA(c->err, cudaMemset(c->m_conv_state, 0, cs * 4));
    ZB_CUDA(c->err, cudaMemset(c->m_ssm_state, 0, ss * 4));
    ZB_CUDA(c->err, cudaMemset(c->m_ynorm, 0, (size_t)rows_pad * c->m_dinner * 2));
  }
  ZB_ALLOC(c->page_table, total_pages * sizeof(int));
  ZB_ALLOC(c->x, (size_t)rows_pad * g.d_model * 2);
  ZB_ALLOC(c->q, (size_t)rows_pad * g.num_heads * ZB_DH * 2);
  ZB_ALLOC(c->y, (size_t)rows_pad * g.num_heads * ZB_DH * 2);
  ZB_ALLOC(c->h, (size_t)rows_pad * g.ffn * 2);
  ZB_ALLOC(c->logits, (size_t)rows_pad * g.n_codebooks * g.vocab * 4);
  ZB_ALLOC(c->part_o, (size_t)rows * g.num_heads_kv * ns * G * ZB_DH * 4);
  ZB_ALLOC(c->part_ml, (size_t)rows * g.num_heads_kv * ns * G * 2 * 4);
  ZB_ALLOC(c->counters, (size_t)rows * g.num_heads_kv * 4);
  ZB_ALLOC(c->tokens, (size_t)rows_pad * g.n_codebooks * 8);
  ZB_ALLOC(c->remaining, (size_t)rows_pad * 4);
  ZB_ALLOC(c->stopping, (size_t)rows_pad * 4);
  ZB_ALLOC(c->out_len, (size_t)rows_pad * 4);
  ZB_ALLOC(c->rope_pos, (size_t)rows_pad * 4);
#undef ZB_ALLOC
  ZB_CUDA(c->err, cudaMemset(c->kv, 0, c->kv_layer_elems * n_kv_layers * sizeof(bf16)));
  ZB_CUDA(c->err, cudaMemset(c->counters, 0, (size_t)rows * g.num_heads_kv * 4));
  ZB_CUDA(c->err, cudaMemset(c->x, 0, (size_t)rows_pad * g.d_model * 2));
  ZB_CUDA(c->err, cudaMemset(c->y, 0, (size_t)rows_pad * g.num_heads * ZB_DH * 2));
  ZB_CUDA(c->err, cudaMemset(c->h, 0, (size_t)rows_pad * g.ffn * 2));
  {
    int rc_bt = bt_setup(c, rows_pad);
    if (rc_bt) return rc_bt;
  }
  ZB_CUDA(c->err, cudaMemset(c->rope_pos, 0, (size_t)rows_pad * 4));
  std::vector<int> pt(total_pages);
  for (size_t i = 0; i < total_pages; ++i) pt[i] = (int)i;          // identity paging: row r owns pages [r*P, (r+1)*P)
  ZB_CUDA(c->err, cudaMemcpy(c->page_table, pt.data(), total_pages * sizeof(int), cudaMemcpyHostToDevice));
  return ZB_OK;
}

int zb_kv_reset(zb_ctx* c, void* stream) {
  if (!c || !c->kv) return c ? c->err.fail(ZB_ERR_STATE, "zb_kv_alloc first") : ZB_ERR_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(c->err, cudaSetDevice(c->device));
  ZB_CUDA(c->err, cudaMemsetAsync(c->st, 0, sizeof(ZbDevState), s));
  ZB_CUDA(c->err, cudaMemsetAsync(c->rope_pos, 0, (size_t)c->rows * 4, s));
  if (c->hybrid && c->n_mamba > 0) {      // recurrent state starts from zero for every new sequence
    ZB_CUDA(c->err, cudaMemsetAsync(c->m_conv_state, 0, (size_t)c->n_mamba * c->rows * c->m_convdim * c->hyb.d_conv * 4, s));
    ZB_CUDA(c->err, cudaMemsetAsync(c->m_ssm_state, 0, (size_t)c->n_mamba * c->rows * c->m_nheads * c->hyb.headdim * c->hyb.d_state * 4, s));
  }
  return ZB_OK;
}

static int check_ready(zb_ctx* c) {
  if (!c) return ZB_ERR_ARG;
  if (!c->finalized) return c->err.fail(ZB_ERR_STATE, "weights not finalised");
  if (!c->kv) return c->err.fail(ZB_ERR_STATE, "zb_kv_alloc first");
  return ZB_OK;
}

int zb_backbone_forward(zb_ctx* c, const void* d_in, void* d_out, int rows, int S, int seqlen_offset,
                        const int32_t* d_lengths, void* stream) {
  int rc = check_ready(c);
  if (rc) return rc;
  if (rows != c->rows) return c->err.fail(ZB_ERR_ARG, "rows %d != allocated %d", rows, c->rows);
  if (seqlen_offset + S > c->max_seqlen) return c->err.fail(ZB_ERR_ARG, "sequence overflows the KV cache");
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(c->err, cudaSetDevice(c->device));
  const zb_config& g = c->cfg;
  if (S > 1 && c->opt_bt_prefill && c->bt_ready && !c->hybrid && rows <= c->bt_mpad)
    return bt_prefill(c, (const bf16*)d_in, (bf16*)d_out, S, seqlen_offset, (const int*)d_lengths, s);
  if (use_mega(c, rows)) {
    ZB_CUDA(c->err, launch(c, k_set_pos, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, rows, seqlen_offset, (const int*)d_lengths, 0));
    ZB_CUDA(c->err, launch_mega(c, MG_MODE_HIDDEN, S, rows, 0, nullptr, d_in, d_out, S, 0, s));   // advances pos / rope_pos by S
    return ZB_OK;
  }
  if (c->hybrid && S > 1 && c->opt_chunked_prefill && c->hyb.d_state == 128 && c->hyb.headdim == 64)
    return hybrid_prefill(c, (const bf16*)d_in, (bf16*)d_out, S, seqlen_offset, (const int*)d_lengths, s);
  const std::vector<Pass> passes = plain_passes(rows);
  int thr = g.d_model / 8 < 1024 ? g.d_model / 8 : 1024;
  for (int t = 0; t < S; ++t) {
    ZB_CUDA(c->err, launch(c, k_set_pos, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, rows, seqlen_offset + t, (const int*)d_lengths, t));
    ZB_CUDA(c->err, launch(c, k_load_hidden, dim3(rows), dim3(thr), 0, s, (const bf16*)d_in, c->x, g.d_model, S, t));
    if (use_bt(c, rows)) {
      ZB_CUDA(c->err, enqueue_bt_layers(c, s));                   // leaves norm_f(x) in bt_xn
      if (d_out) ZB_CUDA(c->err, launch(c, k_bt_store_hidden, dim3(rows), dim3(thr), 0, s, (const bf16*)c->bt_xn, (bf16*)d_out, g.d_model, S, t));
      continue;
    }
    ZB_CUDA(c->err, enqueue_layers(c, passes, s, false));
    if (d_out)
      ZB_CUDA(c->err, launch(c, k_final_ln, dim3(rows), dim3(256), 0, s, (const bf16*)c->x, (const bf16*)c->normf_w, (const bf16*)c->normf_b,
                             g.norm_eps, (bf16*)d_out, g.d_model, S, t));
  }
  // leave the device state positioned after the appended tokens (generation continues from here)
  ZB_CUDA(c->err, launch(c, k_set_pos, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, rows, seqlen_offset + S, (const int*)d_lengths, S));
  return ZB_OK;
}

int zb_wav_to_pcm16(const float* d_wav, int16_t* d_pcm, int64_t n, void* stream) {
  if (n < 0 || (n > 0 && (!d_wav || !d_pcm))) return ZB_ERR_ARG;
  if (n == 0) return ZB_OK;
  long long blocks = (n + 255) / 256;
  if (blocks > ZB_NUM_SMS * 8) blocks = ZB_NUM_SMS * 8;
  k_wav_to_pcm16<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_wav, (short*)d_pcm, (long long)n);
  return cudaGetLastError() == cudaSuccess ? ZB_OK : ZB_ERR_CUDA;
}

int zb_delay_apply(const int64_t* d_codes, int64_t* d_out, int B, int n_q, int T, int64_t mask_token, void* stream) {
  if (B < 0 || n_q <= 0 || T < 0) return ZB_ERR_ARG;
  const size_t n = (size_t)B * n_q * (T + n_q);
  if (n == 0) return ZB_OK;
  if (!d_out || (!d_codes && T > 0)) return ZB_ERR_ARG;
  int blocks = (int)((n + 255) / 256);
  if (blocks > ZB_NUM_SMS * 8) blocks = ZB_NUM_SMS * 8;
  k_delay_apply<<<blocks, 256, 0, (cudaStream_t)stream>>>((const long long*)d_codes, (long long*)d_out, B, n_q, T, (long long)mask_token);
  return cudaGetLastError() == cudaSuccess ? ZB_OK : ZB_ERR_CUDA;
}

int zb_delay_revert(const int64_t* d_delayed, int64_t* d_out, int B, int n_q, int L, void* stream) {
  if (B < 0 || n_q <= 0 || L < n_q) return ZB_ERR_ARG;
  const size_t n = (size_t)B * n_q * (L - n_q);
  if (n == 0) return ZB_OK;
  if (!d_delayed || !d_out) return ZB_ERR_ARG;
  int blocks = (int)((n + 255) / 256);
  if (blocks > ZB_NUM_SMS * 8) blocks = ZB_NUM_SMS * 8;
  k_delay_revert<<<blocks, 256, 0, (cudaStream_t)stream>>>((const long long*)d_delayed, (long long*)d_out, B, n_q, L);
  return cudaGetLastError() == cudaSuccess ? ZB_OK : ZB_ERR_CUDA;
}

int zb_sample(const float* d_logits, const int64_t* d_history, int hist_len, int64_t* d_tokens, float* d_probs_out,
              int B, int n_q, int V, const zb_sampling* sp, uint64_t seed, uint64_t step, void* stream) {
  if (!d_logits || !d_tokens || !sp || B <= 0 || n_q <= 0 || V <= 0 || V > SMP_MAXV) return ZB_ERR_ARG;
  SampleParams p;
  memset(&p, 0, sizeof(p));
  p.logits = d_logits; p.history = (const long long*)d_history; p.hist_stride = hist_len; p.hist_len = d_history ? hist_len : 0;
  p.tokens = (long long*)d_tokens; p.probs_out = d_probs_out; p.n_q = n_q; p.V = V; p.sp = *sp; p.seed = seed; p.step = step; p.st = nullptr;
  k_sample<<<dim3(n_q, B), SMP_THREADS, 0, (cudaStream_t)stream>>>(p);
  return cudaGetLastError() == cudaSuccess ? ZB_OK : ZB_ERR_CUDA;
}

static int upload_state(zb_ctx* c, const ZbDevState& h, cudaStream_t s) {
  // small pageable H2D copy: staged by the driver before returning, so `h` may live on the stack
  ZB_CUDA(c->err, cudaMemcpyAsync(c->st, &h, sizeof(h), cudaMemcpyHostToDevice, s));
  return ZB_OK;
}

int zb_decode_logits(zb_ctx* c, const int64_t* d_delayed, int B, int L, int col, float cfg_scale, int apply_eos_bias,
                     float* d_logits_out, void* stream) {
  int rc = check_ready(c);
  if (rc) return rc;
  if (2 * B != c->rows) return c->err.fail(ZB_ERR_ARG, "2*B=%d != allocated rows %d", 2 * B, c->rows);
  if (col < 0 || col >= L) return c->err.fail(ZB_ERR_ARG, "col out of range");
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(c->err, cudaSetDevice(c->device));
  const zb_config& g = c->cfg;
  // read back pos (tiny sync; this entry point is a parity/debug path, not the generation loop)
  ZbDevState h;
  ZB_CUDA(c->err, cudaMemcpyAsync(&h, c->st, sizeof(h), cudaMemcpyDeviceToHost, s));
  ZB_CUDA(c->err, cudaStreamSynchronize(s));
  if (h.pos + 1 > c->max_seqlen) return c->err.fail(ZB_ERR_ARG, "KV cache full");
  h.col = col; h.L = L; h.B = B; h.done = 0; h.cfg_scale = cfg_scale; h.apply_eos_bias = apply_eos_bias;
  if ((rc = upload_state(c, h, s))) return rc;
  c->B = B; c->L = L;
  if (use_mega(c, 2 * B)) {
    long long* keep = c->delayed;
    c->delayed = (long long*)d_delayed;
    cudaError_t em = launch_mega(c, MG_MODE_LOGITS, 1, 2 * B, B, d_logits_out, nullptr, nullptr, 0, 0, s);
    c->delayed = keep;
    ZB_CUDA(c->err, em);
    return ZB_OK;
  }
  const std::vector<Pass> passes = cfg_passes(B);
  int thr = g.d_model / 8 < 1024 ? g.d_model / 8 : 1024;
  ZB_CUDA(c->err, launch(c, k_embed, dim3(B), dim3(thr), 0, s, (const bf16*)c->emb, (const long long*)d_delayed, c->x,
                         g.d_model, g.n_codebooks, g.vocab, (const ZbDevState*)c->st));
  if (use_bt(c, 2 * B)) {
    ZB_CUDA(c->err, enqueue_bt_layers(c, s));
    ZB_CUDA(c->err, enqueue_bt_heads(c, d_logits_out, s));
  } else {
    ZB_CUDA(c->err, enqueue_layers(c, passes, s, true));
    ZB_CUDA(c->err, enqueue_heads(c, passes, d_logits_out, s));
  }
  ZB_CUDA(c->err, launch(c, k_advance, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, c->rows));
  return ZB_OK;
}

int zb_gen_begin(zb_ctx* c, const void* d_prefix, int B, int Lc, int64_t* d_delayed, int L, int prefix_cols,
                 float cfg_scale, const zb_sampling* sp, uint64_t seed, int compat_scatter, void* stream) {
  int rc = check_ready(c);
  if (rc) return rc;
  if (!d_prefix || !d_delayed || !sp) return c->err.fail(ZB_ERR_ARG, "null argument");
  if (2 * B != c->rows) return c->err.fail(ZB_ERR_ARG, "2*B=%d != allocated rows %d", 2 * B, c->rows);
  if (prefix_cols < 1 || prefix_cols >= L) return c->err.fail(ZB_ERR_ARG, "prefix_cols out of range");
  if (Lc + L > c->max_seqlen) return c->err.fail(ZB_ERR_ARG, "Lc + L = %d exceeds max_seqlen %d", Lc + L, c->max_seqlen);
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(c->err, cudaSetDevice(c->device));
  const zb_config& g = c->cfg;
  if (B > zb_ctx::kGenMaxB) return c->err.fail(ZB_ERR_ARG, "at most %d utterances per context", zb_ctx::kGenMaxB);
  c->B = B; c->L = L; c->delayed = (long long*)d_delayed;
  c->gen_head = c->gen_tail = 0; c->last_ms = -1.f;

  ZbDevState h;
  memset(&h, 0, sizeof(h));
  h.L = L; h.B = B; h.cfg_scale = cfg_scale; h.seed = seed; h.sp = *sp; h.compat_scatter = compat_scatter;
  if ((rc = upload_state(c, h, s))) return rc;
  ZB_CUDA(c->err, cudaMemsetAsync(c->rope_pos, 0, (size_t)c->rows * 4, s));
  ZB_CUDA(c->err, cudaMemsetAsync(c->stopping, 0, (size_t)B * 4, s));
  ZB_CUDA(c->err, cudaMemsetAsync(c->out_len, 0, (size_t)B * 4, s));
  std::vector<int> rem(B, L - prefix_cols);                           // max_steps, model.py:270-271
  ZB_CUDA(c->err, cudaMemcpyAsync(c->remaining, rem.data(), (size_t)B * 4, cudaMemcpyHostToDevice, s));
  ZB_CUDA(c->err, cudaStreamSynchronize(s));                          // `rem` is host-stack memory

  const bool mega = use_mega(c, 2 * B);
  const bool block_prefill = c->opt_bt_prefill && c->bt_ready && !c->hybrid && c->rows <= c->bt_mpad;
  if (block_prefill) {
    // prefill (model.py:181-196,253-264) as blocks of token rows on the batched tcgen05 path: the Lc conditioning tokens,
    // then the embedded known audio-prefix columns - every weight is streamed once per block of tokens, not once per token
    if (Lc > 1 || !mega) { if ((rc = bt_prefill(c, (const bf16*)d_prefix, nullptr, Lc, 0, nullptr, s))) return rc; }
    else ZB_CUDA(c->err, launch_mega(c, MG_MODE_HIDDEN, Lc, 2 * B, B, nullptr, d_prefix, nullptr, Lc, 0, s));
    const int P = prefix_cols - 1;
    if (P > 0) {
      const size_t need = (size_t)c->rows * P * g.d_model;
      if (need > c->pf_embed_elems) {
        ZB_CUDA(c->err, cudaStreamSynchronize(s));
        if (c->pf_embed) cudaFree(c->pf_embed);
        c->pf_embed = nullptr; c->pf_embed_elems = 0;
        ZB_CUDA(c->err, cudaMalloc((void**)&c->pf_embed, need * 2));
        c->pf_embed_elems = need;
      }
      int thr = g.d_model / 8 < 1024 ? g.d_model / 8 : 1024;
      k_embed_cols<<<dim3(P, B), thr, 0, s>>>((const bf16*)c->emb, (const long long*)c->delayed, c->pf_embed, g.d_model, g.n_codebooks, g.vocab, B, L, P);
      ZB_CUDA(c->err, cudaGetLastError());
      if ((rc = bt_prefill(c, c->pf_embed, nullptr, P, Lc, nullptr, s))) return rc;
    }
  } else if (mega) {
    // prefill through the persistent kernel: Lc conditioning tokens, then the known audio-prefix columns
    ZB_CUDA(c->err, launch_mega(c, MG_MODE_HIDDEN, Lc, 2 * B, B, nullptr, d_prefix, nullptr, Lc, 0, s));
    if (prefix_cols > 1) {
      h.pos = Lc; h.col = 0;
      if ((rc = upload_state(c, h, s))) return rc;
      ZB_CUDA(c->err, launch_mega(c, MG_MODE_FORCED, prefix_cols - 1, 2 * B, B, nullptr, nullptr, nullptr, 0, 0, s));
    }
  } else {
  // ---- prefill (model.py:181-196,253-264), token by token through the decode-step kernels: causal attention over
    // the prefix is exactly what a sequence of single-token steps computes.
    const std::vector<Pass> passes = cfg_passes(B);
    int thr = g.d_model / 8 < 1024 ? g.d_model / 8 : 1024;
    if (c->hybrid && Lc > 1 && c->opt_chunked_prefill && c->hyb.d_state == 128 && c->hyb.headdim == 64) {
      if ((rc = hybrid_prefill(c, (const bf16*)d_prefix, nullptr, Lc, 0, nullptr, s))) return rc;      // leaves pos = rope_pos = Lc
    } else {
      for (int t = 0; t < Lc; ++t) {
        ZB_CUDA(c->err, launch(c, k_load_hidden, dim3(2 * B), dim3(thr), 0, s, (const bf16*)d_prefix, c->x, g.d_model, Lc, t));
        if (use_bt(c, 2 * B)) ZB_CUDA(c->err, enqueue_bt_layers(c, s));
        else ZB_CUDA(c->err, enqueue_layers(c, passes, s, false));
        ZB_CUDA(c->err, launch(c, k_advance, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, c->rows));
      }
    }
    for (int j = 0; j + 1 < prefix_cols; ++j) {      // known audio-prefix columns: no sampling
      h.pos = Lc + j; h.col = j;
      if ((rc = upload_state(c, h, s))) return rc;
      ZB_CUDA(c->err, launch(c, k_embed, dim3(B), dim3(thr), 0, s, (const bf16*)c->emb, (const long long*)c->delayed, c->x,
                             g.d_model, g.n_codebooks, g.vocab, (const ZbDevState*)c->st));
      if (use_bt(c, 2 * B)) ZB_CUDA(c->err, enqueue_bt_layers(c, s));
      else ZB_CUDA(c->err, enqueue_layers(c, passes, s, false));
      ZB_CUDA(c->err, launch(c, k_advance, dim3(1), dim3(128), 0, s, c->st, c->rope_pos, c->rows));
    }
  }
  // last prefix column -> logits -> sample (no logit_bias, no repetition penalty, no EOS logic: model.py:255-260)
  h.pos = Lc + prefix_cols - 1; h.col = prefix_cols - 1;
  h.apply_eos_bias = 0; h.use_history = 0; h.eos_logic = 0;
  if ((rc = upload_state(c, h, s))) return rc;
  if (mega) ZB_CUDA(c->err, launch_mega(c, MG_MODE_DECODE, 1, 2 * B, B, nullptr, nullptr, nullptr, 0, 0, s));
  else ZB_CUDA(c->err, enqueue_decode_step(c, s));
  // switch to loop mode (model.py:266-301)
  h.pos = Lc + prefix_cols; h.col = prefix_cols; h.step = 1;
  h.apply_eos_bias = 1; h.use_history = 1; h.eos_logic = 1;
  if ((rc = upload_state(c, h, s))) return rc;
  return ZB_OK;
}

int zb_gen_steps(zb_ctx* c, int n_steps, void* stream) {
  int rc = check_ready(c);
  if (rc) return rc;
  if (!c->delayed) return c->err.fail(ZB_ERR_STATE, "zb_gen_begin first");
  if (c->gen_head - c->gen_tail >= zb_ctx::kGenSlots)
    return c->err.fail(ZB_ERR_STATE, "%d zb_gen_steps calls in flight: zb_gen_poll first", zb_ctx::kGenSlots);
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(c->err, cudaSetDevice(c->device));
  zb_ctx::GenSlot& g = c->gen[c->gen_head % zb_ctx::kGenSlots];
  ZB_CUDA(c->err, cudaEventRecord(g.e0, s));
  if (use_mega(c, 2 * c->B)) {
    ZB_CUDA(c->err, launch_mega(c, MG_MODE_DECODE, n_steps, 2 * c->B, c->B, nullptr, nullptr, nullptr, 0, 0, s));
  } else if (c->opt_graph) {
    if (c->graph && (c->graph_B != c->B || c->graph_L != c->L || c->graph_delayed != c->delayed)) drop_graph(c);
    if (!c->graph) {
      cudaGraph_t graph = nullptr;
      const int64_t before = c->launches;
      ZB_CUDA(c->err, cudaStreamBeginCapture(c->cap_stream, cudaStreamCaptureModeThreadLocal));
      cudaError_t e = enqueue_decode_step(c, c->cap_stream);
      cudaError_t e2 = cudaStreamEndCapture(c->cap_stream, &graph);
      if (e != cudaSuccess || e2 != cudaSuccess) {
        if (graph) cudaGraphDestroy(graph);
        return c->err.fail(ZB_ERR_CUDA, "graph capture failed: %s / %s", cudaGetErrorString(e), cudaGetErrorString(e2));
      }
      c->graph_nodes = (int)(c->launches - before);
      c->launches = before;
      cudaError_t e3 = cudaGraphInstantiate(&c->graph, graph, 0);
      cudaGraphDestroy(graph);
      if (e3 != cudaSuccess) { c->graph = nullptr; return c->err.fail(ZB_ERR_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e3)); }
      c->graph_B = c->B; c->graph_L = c->L; c->graph_delayed = c->delayed;
      ZB_CUDA(c->err, cudaEventRecord(g.e0, s));      // keep the capture / instantiation out of the timed span
    }
    for (int i = 0; i < n_steps; ++i) ZB_CUDA(c->err, cudaGraphLaunch(c->graph, s));
    c->launches += (int64_t)n_steps * c->graph_nodes;
  } else {
    for (int i = 0; i < n_steps; ++i) ZB_CUDA(c->err, enqueue_decode_step(c, s));
  }
  ZB_CUDA(c->err, cudaEventRecord(g.e1, s));
  // snapshot of the loop state after this call, into pinned memory: zb_gen_poll reads it once e2 has fired
  ZB_CUDA(c->err, cudaMemcpyAsync(g.h_st, c->st, sizeof(ZbDevState), cudaMemcpyDeviceToHost, s));
  ZB_CUDA(c->err, cudaMemcpyAsync(g.h_len, c->out_len, (size_t)c->B * 4, cudaMemcpyDeviceToHost, s));
  ZB_CUDA(c->err, cudaEventRecord(g.e2, s));
  c->gen_head += 1;
  return ZB_OK;
}

int zb_gen_poll(zb_ctx* c, int* done, int* col, int32_t* h_out_len, void* stream) {
  int rc = check_ready(c);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(c->err, cudaSetDevice(c->device));
  if (c->gen_head > c->gen_tail) {
    // state after the OLDEST zb_gen_steps call that has not been polled yet; later calls keep running meanwhile
    zb_ctx::GenSlot& g = c->gen[c->gen_tail % zb_ctx::kGenSlots];
    ZB_CUDA(c->err, cudaEventSynchronize(g.e2));
    float ms = -1.f;
    if (cudaEventElapsedTime(&ms, g.e0, g.e1) == cudaSuccess) c->last_ms = ms;
    c->gen_tail += 1;
    if (done) *done = g.h_st->done;
    if (col) *col = g.h_st->col;
    if (h_out_len) memcpy(h_out_len, g.h_len, (size_t)c->B * 4);
    return ZB_OK;
  }
  ZbDevState h;
  ZB_CUDA(c->err, cudaMemcpyAsync(&h, c->st, sizeof(h), cudaMemcpyDeviceToHost, s));
  if (h_out_len) ZB_CUDA(c->err, cudaMemcpyAsync(h_out_len, c->out_len, (size_t)c->B * 4, cudaMemcpyDeviceToHost, s));
  ZB_CUDA(c->err, cudaStreamSynchronize(s));
  if (done) *done = h.done;
  if (col) *col = h.col;
  return ZB_OK;
}

int zb_update_frame(zb_ctx* c, int64_t* d_delayed, int B, int L, int col, const int64_t* d_tokens, int32_t* d_remaining,
                    int32_t* d_stopping, int32_t* d_out_len, int eos_logic, int compat_scatter, void* stream) {
  // One application of the frame write-back / EOS wind-down kernel on caller-owned state (exact-integer parity tests
  // against model.py:283-301 without any model numerics in the way).
  if (!c || !d_delayed || !d_tokens || !d_remaining || !d_stopping || !d_out_len || B <= 0 || col < 0) return ZB_ERR_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(c->err, cudaSetDevice(c->device));
  ZbDevState h;
  memset(&h, 0, sizeof(h));
  h.col = col; h.L = L; h.B = B; h.eos_logic = eos_logic; h.compat_scatter = compat_scatter;
  int rc = upload_state(c, h, s);
  if (rc) return rc;
  UpdateParams up;
  memset(&up, 0, sizeof(up));
  up.st = c->st; up.delayed = (long long*)d_delayed; up.tokens = (const long long*)d_tokens; up.remaining = d_remaining;
  up.stopping = d_stopping; up.out_len = d_out_len; up.rope_pos = nullptr; up.rows = 0; up.n_q = c->cfg.n_codebooks;
  up.eos = c->cfg.eos_token_id; up.mask = c->cfg.masked_token_id;
  const int ub = B <= 8 ? 32 * B : 256;              // one warp per utterance
  k_update<<<1, ub, 0, s>>>(up);
  ZB_CUDA(c->err, cudaGetLastError());
  return ZB_OK;
}

// ---- PrefixConditioner building blocks: stateless, caller-owned tensors ------------------------------------
int zb_pc_fourier(const float* d_x, const void* d_weight, float min_val, float max_val, void* d_out, int n, int in_dim, int half, void* stream) {
  if (!d_x || !d_weight || !d_out || n < 0 || in_dim <= 0 || half <= 0) return ZB_ERR_ARG;
  if (n == 0) return ZB_OK;
  k_pc_fourier<<<n, 128, 0, (cudaStream_t)stream>>>(d_x, (const bf16*)d_weight, min_val, max_val, (bf16*)d_out, in_dim, half);
  return cudaGetLastError() == cudaSuccess ? ZB_OK : ZB_ERR_CUDA;
}
int zb_pc_embed(const int64_t* d_ids, const void* d_table, void* d_out, int n, int d, int64_t id_offset, int64_t table_rows, void* stream) {
  if (!d_ids || !d_table || !d_out || n < 0 || d <= 0 || d % 8 != 0 || table_rows <= 0) return ZB_ERR_ARG;
  if (n == 0) return ZB_OK;
  k_pc_embed<<<n, 128, 0, (cudaStream_t)stream>>>((const long long*)d_ids, (const bf16*)d_table, (bf16*)d_out, d, (long long)id_offset, (long long)table_rows);
  return cudaGetLastError() == cudaSuccess ? ZB_OK : ZB_ERR_CUDA;
}
int zb_pc_linear(const void* d_in, const void* d_w, const void* d_b, void* d_out, int n, int K, int N, int silu, void* stream) {
  if (!d_in || !d_w || !d_out || n < 0 || K <= 0 || K % 8 != 0 || N <= 0) return ZB_ERR_ARG;
  if (n == 0) return ZB_OK;
  k_pc_linear<<<(N + 7) / 8, 256, 0, (cudaStream_t)stream>>>((const bf16*)d_in, (const bf16*)d_w, (const bf16*)d_b, (bf16*)d_out, n, K, N, silu);
  return cudaGetLastError() == cudaSuccess ? ZB_OK : ZB_ERR_CUDA;
}
int zb_pc_layernorm(const void* d_in, const void* d_w, const void* d_b, float eps, void* d_out, int n, int d, void* stream) {
  if (!d_in || !d_w || !d_b || !d_out || n < 0 || d <= 0) return ZB_ERR_ARG;
  if (n == 0) return ZB_OK;
  k_pc_layernorm<<<n, 256, 0, (cudaStream_t)stream>>>((const bf16*)d_in, (const bf16*)d_w, (const bf16*)d_b, eps, (bf16*)d_out, d);
  return cudaGetLastError() == cudaSuccess ? ZB_OK : ZB_ERR_CUDA;
}

int64_t zb_weight_bytes(const zb_ctx* c) { return c ? c->step_weight_bytes : 0; }
int64_t zb_launch_count(const zb_ctx* c) { return c ? c->launches : 0; }

int zb_set_option(zb_ctx* c, const char* key, int value) {
  if (!c || !key) return ZB_ERR_ARG;
  const std::string k(key);
  if (k == "pdl") c->opt_pdl = value;
  else if (k == "l2_prefetch") c->opt_l2_prefetch_mb = value;
  else if (k == "graph") c->opt_graph = value;
  else if (k == "gemv") c->opt_gemv = value;
  else if (k == "mega") c->opt_mega = value;
  else if (k == "trace") c->opt_trace = value;
  else if (k == "trace_quiet") c->opt_trace_quiet = value;
  else if (k == "att_chunk") c->opt_att_chunk = value;
  else if (k == "ll") c->opt_ll = value;
  else if (k == "chain") c->opt_chain = value;
  else if (k == "bt_cluster") {
    c->opt_bt_cluster = (value == 2 || value == 4) ? value : 1;
    if (c->bt_ready) { int rc = bt_make_slices(c); if (rc) return rc; }
  }
  else if (k == "att_single") c->opt_att_single = value;
  else if (k == "ring") c->opt_ring = value;
  else if (k == "psleep") c->opt_psleep = value;
  else if (k == "evict") c->opt_evict = value;
  else if (k == "chunked_prefill") c->opt_chunked_prefill = value;
  else if (k == "bt_prefill") c->opt_bt_prefill = value;
  else if (k == "bt_min_rows") c->opt_bt_min_rows = value;
  else if (k == "attn_mma") c->opt_attn_mma = value;
  else if (k == "attn_tma") c->opt_attn_tma = value;
  else if (k == "bt_cta2") c->opt_bt_cta2 = value;
  else if (k == "nsplit") c->opt_nsplit = value;
  else return c->err.fail(ZB_ERR_ARG, "unknown option '%s'", key);
  drop_graph(c);
  return ZB_OK;
}

int zb_bench_kernel(zb_ctx* c, const char* name, int iters, float* ms_out, void* stream) {
  // Isolated timing of one kernel family: launch it `iters` times back to back, cycling over the layers so that
  // consecutive launches read different weights (26 x 67 MB >> L2), CUDA events on `stream`.  Outputs are scratch.
  int rc = check_ready(c);
  if (rc) return rc;
  if (!name || !ms_out || iters <= 0) return ZB_ERR_ARG;
  if (c->hybrid) return c->err.fail(ZB_ERR_ARG, "zb_bench_kernel covers the transformer layers only");
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(c->err, cudaSetDevice(c->device));
  const zb_config& g = c->cfg;
  const int d = g.d_model, H = g.num_heads, Hkv = g.num_heads_kv, F = g.ffn;
  const int q_size = H * ZB_DH, kv_size = Hkv * ZB_DH;
  const std::string k(name);
  const Pass ps = (c->rows % 2 == 0 && c->rows <= 8) ? Pass{c->rows, 0, c->rows / 2, c->rows / 2, c->rows} : Pass{2, 0, 1, 1, 2};
  const bool counting = c->counting;
  c->counting = false;
  cudaError_t e = cudaSuccess;
  {
    ZbDevState h;
    ZB_CUDA(c->err, cudaMemcpyAsync(&h, c->st, sizeof(h), cudaMemcpyDeviceToHost, s));
    ZB_CUDA(c->err, cudaStreamSynchronize(s));
    h.done = 0;
    if (h.pos >= c->max_seqlen) h.pos = c->max_seqlen - 1;
    if ((rc = upload_state(c, h, s))) return rc;
  }
  ZB_CUDA(c->err, cudaEventRecord(c->ev0, s));
  for (int i = 0; i < iters && e == cudaSuccess; ++i) {
    const int l = i % g.n_layer;
    const LayerW& w = c->layers[l];
    GemvParams p = base_params(c, ps);
    if (k == "fc1") {
      p.W = w.w_fc1; p.K = d; p.n_tasks = F; p.row0_mul = 1; p.row1_off = F; p.x = c->x; p.ln_w = w.ln2_w; p.ln_b = w.ln2_b;
      p.out = c->h; p.out_stride = F;
      e = launch_gemv(c, EPI_SWIGLU, ps, p, s);
    } else if (k == "fc2") {
      p.W = w.w_fc2; p.K = F; p.n_tasks = d / 2; p.row0_mul = 2; p.row1_off = 1; p.x = c->h; p.out = c->y; p.out_stride = d;
      e = launch_gemv(c, EPI_RESID, ps, p, s);
    } else if (k == "out_proj") {
      p.W = w.w_out; p.K = q_size; p.n_tasks = d / 2; p.row0_mul = 2; p.row1_off = 1; p.x = c->q; p.out = c->y; p.out_stride = d;
      e = launch_gemv(c, EPI_RESID, ps, p, s);
    } else if (k == "in_proj") {
      p.W = w.w_in; p.K = d; p.n_tasks = (q_size + 2 * kv_size) / 2; p.row0_mul = 2; p.row1_off = 1; p.x = c->x; p.ln_w = w.ln1_w; p.ln_b = w.ln1_b;
      p.out = c->q; p.out_stride = q_size; p.kv_layer = c->kv + (size_t)l * c->kv_layer_elems; p.page_table = c->page_table;
      p.max_pages = c->max_pages; p.Hkv = Hkv; p.q_size = q_size; p.kv_size = kv_size; p.rope = c->rope; p.rope_pos = c->rope_pos;
      e = launch_gemv(c, EPI_QKV, ps, p, s);
    } else if (k == "attn") {
      AttnParams a;
      memset(&a, 0, sizeof(a));
      a.q = c->q; a.kv_layer = c->kv + (size_t)l * c->kv_layer_elems; a.page_table = c->page_table; a.max_pages = c->max_pages;
      a.H = H; a.Hkv = Hkv; a.y = c->y; a.part_o = c->part_o; a.part_ml = c->part_ml; a.counters = c->counters;
      a.st = c->st; a.scale_log2 = (1.0f / sqrtf((float)ZB_DH)) * 1.4426950408889634f; a.total_pages = c->rows * c->max_pages;
      e = launch(c, attn_kernel(H / Hkv, c->opt_attn_mma != 0), dim3(c->nsplit, Hkv, c->rows), dim3(ATT_THREADS), 0, s, a);
    } else {
      c->counting = counting;
      return c->err.fail(ZB_ERR_ARG, "unknown kernel '%s'", name);
    }
  }
  c->counting = counting;
  ZB_CUDA(c->err, e);
  ZB_CUDA(c->err, cudaEventRecord(c->ev1, s));
  ZB_CUDA(c->err, cudaEventSynchronize(c->ev1));
  ZB_CUDA(c->err, cudaEventElapsedTime(ms_out, c->ev0, c->ev1));
  return ZB_OK;
}

float zb_last_steps_ms(zb_ctx* c) { return c ? c->last_ms : -1.f; }

int zb_last_trace(zb_ctx* c, uint64_t* out, int max_slots) {
  if (!c || !out || max_slots <= 0) return 0;
  int n = 0;
  for (; n < (int)c->last_trace.size() && n < max_slots && c->last_trace[n] != 0; ++n) out[n] = c->last_trace[n];
  return n;
}

}  // extern "C"
