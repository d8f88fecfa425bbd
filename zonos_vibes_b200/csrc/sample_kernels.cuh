// Fused sampling (zonos/sampling.py:117-182), frame write-back + EOS wind-down (zonos/model.py:283-301) and the
// delay-pattern index kernels (zonos/codebook_pattern.py:5-12).  All latency-bound: one CTA per (utterance, codebook).
#pragma once
#include "zb_common.cuh"

#define SMP_THREADS 256
#define SMP_MAXV 2048            // bitonic sort width (vocab 1026 -> padded)
#define SMP_PER ((SMP_MAXV + SMP_THREADS - 1) / SMP_THREADS)

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011) — own counter-based RNG; key = seed, counter = (elem block, stream, step).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll 1
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0; key.y += W1;
  }
  return ctr;
}
// uniform in (0, 1]
__device__ __forceinline__ float u01(uint32_t r) { return ((float)(r >> 8) + 1.0f) * (1.0f / 16777216.0f); }

// Block reductions over SMP_THREADS threads.  Out of line and shared by all call sites: this code runs once per decode
// step inside the persistent kernel, where it is instruction-cache cold every time (~0.15 us per 128-byte line of
// straight-line code, profiles/r02_summary.md) - compactness is what makes it fast.  The summation order (own elements
// ascending, xor butterfly inside the warp, warps in order) is fixed, so results are reproducible.
template <class Sync>
__device__ __noinline__ float block_reduce(float v, float* red, const bool is_max) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = is_max ? warp_max(v) : warp_sum(v);
  Sync::sync();
  if (lane == 0) red[warp] = v;
  Sync::sync();
  float t = is_max ? -INFINITY : 0.f;
#pragma unroll 1
  for (int w = 0; w < SMP_THREADS / 32; ++w) t = is_max ? fmaxf(t, red[w]) : t + red[w];
  return t;
}
template <class Sync>
__device__ __forceinline__ float block_reduce_sum(float v, float* red) { return block_reduce<Sync>(v, red, false); }
template <class Sync>
__device__ __forceinline__ float block_reduce_max(float v, float* red) { return block_reduce<Sync>(v, red, true); }
// argmax with first-index tie break (torch.argmax)
template <class Sync>
__device__ __noinline__ int block_argmax(float v, int idx, float* red, int* redi) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll 1
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
  }
  Sync::sync();
  if (lane == 0) { red[warp] = v; redi[warp] = idx; }
  Sync::sync();
  float bv = red[0]; int bi = redi[0];
#pragma unroll 1
  for (int w = 1; w < SMP_THREADS / 32; ++w)
    if (red[w] > bv || (red[w] == bv && redi[w] < bi)) { bv = red[w]; bi = redi[w]; }
  return bi;
}

struct SampleParams {
  const float* logits;          // [B][n_q][V]
  const long long* history;     // [B][n_q][hist_stride]; columns [hist_lo, hist_hi) are the repetition window source
  int hist_stride;
  int hist_len;                 // number of valid leading columns (generated_tokens.shape[-1]); <=0: no penalty
  long long* tokens;            // [B][n_q]
  float* probs_out;             // optional [B][n_q][V]
  int n_q, V;
  zb_sampling sp;
  unsigned long long seed, step;
  const ZbDevState* st;         // non-NULL: sampling params / history bounds / seed / step come from device state
  unsigned long long* dbg;      // optional (traced persistent kernel only): %globaltimer stamps of thread 0
};
#define SMP_STAMP()                                                                             \
  do {                                                                                          \
    if (dbg_q != nullptr && threadIdx.x == 0) {                                                 \
      unsigned long long t_;                                                                    \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                    \
      *dbg_q++ = t_;                                                                            \
    }                                                                                           \
  } while (0)

// Shared-memory scratch of one sampling block (SMP_THREADS threads): 3 * SMP_MAXV * 4 + 80 bytes.
struct SampleSmem {
  float* s_p;      // [SMP_MAXV]
  float* s_key;    // [SMP_MAXV]
  int* s_idx;      // [SMP_MAXV]
  float* red;      // [SMP_THREADS / 32]
  int* redi;       // [SMP_THREADS / 32]
  float* s_kth;    // [1]
};
#define SMP_SMEM_BYTES (3 * SMP_MAXV * 4 + 2 * (SMP_THREADS / 32) * 4 + 16)
__device__ __forceinline__ SampleSmem sample_smem_carve(unsigned char* base) {
  SampleSmem m;
  m.s_p = reinterpret_cast<float*>(base);
  m.s_key = m.s_p + SMP_MAXV;
  m.s_idx = reinterpret_cast<int*>(m.s_key + SMP_MAXV);
  m.red = reinterpret_cast<float*>(m.s_idx + SMP_MAXV);
  m.redi = reinterpret_cast<int*>(m.red + SMP_THREADS / 32);
  m.s_kth = reinterpret_cast<float*>(m.redi + SMP_THREADS / 32);
  return m;
}

// One sampling block = threads 0..SMP_THREADS-1 of the CTA, all of which must call this (block-uniform control flow);
// Sync::sync() is the barrier over exactly those threads.  (k, b) = (codebook, utterance).
template <class Sync>
__device__ void sample_body(const SampleParams& p, const int k, const int b, const SampleSmem& sm) {
  float* const s_p = sm.s_p;
  float* const s_key = sm.s_key;
  int* const s_idx = sm.s_idx;
  float* const red = sm.red;
  int* const redi = sm.redi;

  unsigned long long* dbg_q = p.dbg;
  SMP_STAMP();                                   // s0: entered
  zb_sampling sp = p.sp;
  unsigned long long seed = p.seed, step = p.step;
  int hist_len = p.hist_len;
  if (p.st) {
    if (p.st->done) return;
    sp = p.st->sp; seed = p.st->seed; step = (unsigned long long)p.st->step;
    hist_len = p.st->use_history ? p.st->col + 1 : 0;      // generated_tokens = delayed[..., :offset], offset = col+1
  }
  const int tid = threadIdx.x, V = p.V;
  const float* lg = p.logits + ((size_t)b * p.n_q + k) * V;
  // The working vector lives in shared memory (s_p) and every pass is a rolled loop over this thread's elements
  // i = tid, tid + SMP_THREADS, ... (ascending, as before: the per-thread accumulation order is unchanged).

  // ---- repetition penalty (sampling.py:99-114): factor = penalty^(#occurrences in the window), ids clamped to V-1;
  // factor built by repeated multiplication, exactly like scatter_reduce(prod) over ones.
  int win = 0;
  const long long* h = nullptr;
  if (sp.repetition_penalty != 1.0f && p.history != nullptr && hist_len > 0) {
    win = min(sp.repetition_penalty_window, hist_len);
    h = p.history + ((size_t)b * p.n_q + k) * p.hist_stride + hist_len - win;
#pragma unroll 1
    for (int w = tid; w < min(win, SMP_MAXV); w += SMP_THREADS) {       // window tokens staged once (s_idx is free until the sort)
      const long long tok = __ldcg(h + w);
      s_idx[w] = (int)(tok > V - 1 ? V - 1 : tok);
    }
  }
  Sync::sync();
#pragma unroll 1
  for (int i = tid; i < V; i += SMP_THREADS) {
    const float raw = __ldcg(lg + i);
    float f = 1.0f;
#pragma unroll 1
    for (int w = 0; w < win; ++w) {
      long long tok = w < SMP_MAXV ? (long long)s_idx[w] : __ldcg(h + w);
      if (tok > V - 1) tok = V - 1;
      if ((long long)i == tok) f *= sp.repetition_penalty;
    }
    s_p[i] = win > 0 ? (raw <= 0.f ? raw * f : raw / f) : raw;
  }
  Sync::sync();                                  // s_idx is reused below
  SMP_STAMP();                                   // s1: logits (+ history) loaded, penalty applied

  int token;
  if (sp.temperature > 0.f) {
    // ---- softmax(logits / T)
    float mx = -INFINITY;
#pragma unroll 1
    for (int i = tid; i < V; i += SMP_THREADS) { const float v = s_p[i] / sp.temperature; s_p[i] = v; mx = fmaxf(mx, v); }
    mx = block_reduce_max<Sync>(mx, red);
    float sum = 0.f;
#pragma unroll 1
    for (int i = tid; i < V; i += SMP_THREADS) { const float e = expf(s_p[i] - mx); s_p[i] = e; sum += e; }
    sum = block_reduce_sum<Sync>(sum, red);
#pragma unroll 1
    for (int i = tid; i < V; i += SMP_THREADS) s_p[i] = s_p[i] / sum;
    SMP_STAMP();                                 // s2: softmax done

    // ---- unified sampler (sampling.py:29-43); s_key holds log p
    if (sp.linear > 0.f) {
      float ent = 0.f;
#pragma unroll 1
      for (int i = tid; i < V; i += SMP_THREADS) { const float x = s_p[i], lp = logf(fmaxf(x, 1e-20f)); s_key[i] = lp; ent += x * lp; }
      ent = -block_reduce_sum<Sync>(ent, red);
      float mx2 = -INFINITY;
#pragma unroll 1
      for (int i = tid; i < V; i += SMP_THREADS) {
        const float lp = s_key[i];
        const float raw = lp * (sp.linear + ent * sp.conf) - lp * lp * sp.quad;
        s_p[i] = raw;
        mx2 = fmaxf(mx2, raw);
      }
      mx2 = block_reduce_max<Sync>(mx2, red);
      float sum2 = 0.f;
#pragma unroll 1
      for (int i = tid; i < V; i += SMP_THREADS) { const float e = expf(s_p[i] - mx2); s_p[i] = e; sum2 += e; }
      sum2 = block_reduce_sum<Sync>(sum2, red);
#pragma unroll 1
      for (int i = tid; i < V; i += SMP_THREADS) s_p[i] = s_p[i] / sum2;
    }

    // ---- top-p / top-k need the descending order: bitonic sort of (prob, index) in shared memory
    if (sp.top_p > 0.f || sp.top_k > 0) {
#pragma unroll 1
      for (int i = tid; i < SMP_MAXV; i += SMP_THREADS) {
        s_key[i] = i < V ? s_p[i] : -1.f;          // padding sorts last
        s_idx[i] = i;
        if (i >= V) s_p[i] = 0.f;
      }
      Sync::sync();
#pragma unroll 1
      for (int size = 2; size <= SMP_MAXV; size <<= 1) {
#pragma unroll 1
        for (int strd = size >> 1; strd > 0; strd >>= 1) {
#pragma unroll 1
          for (int t = tid; t < SMP_MAXV / 2; t += SMP_THREADS) {
            const int lo = 2 * t - (t & (strd - 1));
            const int hi = lo + strd;
            const bool desc = ((lo & size) == 0);
            const float a = s_key[lo], c = s_key[hi];
            const int ia = s_idx[lo], ic = s_idx[hi];
            // descending by key, ascending index among equal keys (stable like torch's sort on ties)
            const bool a_first = (a > c) || (a == c && ia < ic);
            if (a_first != desc) { s_key[lo] = c; s_key[hi] = a; s_idx[lo] = ic; s_idx[hi] = ia; }
          }
          Sync::sync();
        }
      }
      if (sp.top_p > 0.f) {
        // inclusive cumsum in sorted order, double accumulator rounded to fp32 per element (ATen CPU cumsum),
        // mask = cumsum - p_sorted > top_p  (sampling.py:73-75)
        if (tid == 0) {
          double c = 0.0;
#pragma unroll 1
          for (int r = 0; r < V; ++r) {
            c += (double)s_key[r];
            const float incl = (float)c;
            const bool drop = (incl - s_key[r]) > sp.top_p;
            if (drop) s_p[s_idx[r]] = 0.f;
          }
        }
        Sync::sync();
        float sum3 = 0.f;
#pragma unroll 1
        for (int i = tid; i < V; i += SMP_THREADS) sum3 += s_p[i];
        sum3 = block_reduce_sum<Sync>(sum3, red);
#pragma unroll 1
        for (int i = tid; i < V; i += SMP_THREADS) s_p[i] = s_p[i] / sum3;
        Sync::sync();
      }
      if (sp.top_k > 0) {
        // pivot = k-th largest CURRENT probability; order is unchanged by the monotone steps above
        if (tid == 0) *sm.s_kth = s_p[s_idx[min(sp.top_k, V) - 1]];
        Sync::sync();
        const float kth = *sm.s_kth;
        float sum4 = 0.f;
#pragma unroll 1
        for (int i = tid; i < V; i += SMP_THREADS) { float x = s_p[i]; if (x < kth) x = 0.f; s_p[i] = x; sum4 += x; }
        sum4 = block_reduce_sum<Sync>(sum4, red);
#pragma unroll 1
        for (int i = tid; i < V; i += SMP_THREADS) s_p[i] = s_p[i] / sum4;
      }
    }
    SMP_STAMP();                                 // s3: unified / top-p / top-k done
    // ---- min-p (sampling.py:82-96)
    if (sp.min_p > 0.f) {
      float top = 0.f;
#pragma unroll 1
      for (int i = tid; i < V; i += SMP_THREADS) top = fmaxf(top, s_p[i]);
      top = block_reduce_max<Sync>(top, red);
      const float thr = sp.min_p * top;
      float sum5 = 0.f;
#pragma unroll 1
      for (int i = tid; i < V; i += SMP_THREADS) { float x = s_p[i]; if (x < thr) x = 0.f; s_p[i] = x; sum5 += x; }
      sum5 = block_reduce_sum<Sync>(sum5, red);
#pragma unroll 1
      for (int i = tid; i < V; i += SMP_THREADS) s_p[i] = s_p[i] / sum5;
    }
    SMP_STAMP();                                 // s4: min-p done
    // ---- exponential race (sampling.py:19-21): argmax_i p_i / q_i, q_i ~ Exp(1) = -log(u_i)
    float best = -1.f; int besti = 0x7fffffff;
#pragma unroll 1
    for (int i = tid; i < V; i += SMP_THREADS) {
      const float x = s_p[i];
      if (p.probs_out) p.probs_out[((size_t)b * p.n_q + k) * V + i] = x;
      const uint4 r = philox4x32_10(make_uint4((uint32_t)(i >> 2), (uint32_t)(b * p.n_q + k), (uint32_t)step, (uint32_t)(step >> 32)),
                                    make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
      const uint32_t bits = (i & 3) == 0 ? r.x : (i & 3) == 1 ? r.y : (i & 3) == 2 ? r.z : r.w;
      const float qe = -logf(u01(bits));
      const float sc = x / fmaxf(qe, 1e-30f);
      if (sc > best || (sc == best && i < besti)) { best = sc; besti = i; }
    }
    token = block_argmax<Sync>(best, besti, red, redi);
  } else {
    // greedy: argmax of the (penalised) logits, first index on ties (sampling.py:180)
    float best = -INFINITY; int besti = 0x7fffffff;
#pragma unroll 1
    for (int i = tid; i < V; i += SMP_THREADS) {
      const float x = s_p[i];
      if (p.probs_out) p.probs_out[((size_t)b * p.n_q + k) * V + i] = x;
      if (x > best || (x == best && i < besti) || besti == 0x7fffffff) { best = x; besti = i; }
    }
    token = block_argmax<Sync>(best, besti, red, redi);
  }
  if (tid == 0) p.tokens[(size_t)b * p.n_q + k] = token;
  SMP_STAMP();                                   // s5: token drawn and stored
}

// grid = (n_q, B), block = SMP_THREADS
__global__ void __launch_bounds__(SMP_THREADS) k_sample(const SampleParams p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  __shared__ __align__(16) unsigned char s_raw[SMP_SMEM_BYTES];
  sample_body<SyncCta>(p, blockIdx.x, blockIdx.y, sample_smem_carve(s_raw));
}

// ---------------------------------------------------------------------------------------------
// Frame write-back + EOS wind-down + counters (model.py:283-301).  One thread per utterance does the 9-element
// sequential logic; thread 0 then advances the shared step state.  grid = 1, block >= B.
// ---------------------------------------------------------------------------------------------
struct UpdateParams {
  ZbDevState* st;
  long long* delayed;       // [B][n_q][L]
  const long long* tokens;  // [B][n_q]
  int* remaining;           // [B]
  int* stopping;            // [B]
  int* out_len;             // [B]: valid output frames once finished (col_end - n_q)
  int* rope_pos;            // [rows]
  int rows;
  int n_q;
  int eos, mask;
};

// One block; threads 0..nthreads-1 all call it (nthreads a multiple of 32).  `s_alive` is one int of shared memory.
// One WARP per utterance: lane k owns codebook k, so the nine cells of the written column are read in one round trip and
// masked_scatter_'s sequential routing (source element j goes to the j-th unknown cell) is a ballot +
// popcount instead of nine dependent load-test-store rounds.  Compact on purpose: this runs once per step, cold.
template <class Sync>
__device__ void update_body(const UpdateParams& p, int* s_alive_p, const int nthreads) {
  ZbDevState* st = p.st;
  if (st->done) return;
  int& s_alive = *s_alive_p;
  if (threadIdx.x == 0) s_alive = 0;
  Sync::sync();
  const int B = st->B, L = st->L, col = st->col, wcol = col + 1, nq = p.n_q;
  const int eos_logic = st->eos_logic, compat = st->compat_scatter;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = nthreads >> 5;
#pragma unroll 1
  for (int b = warp; b < B; b += nwarps) {
    int rem = p.remaining[b];
    if (rem <= 0) continue;                           // finished utterances are frozen (independent batch-1 runs)
    long long tok = lane < nq ? __ldcg(p.tokens + (size_t)b * nq + lane) : 0;
    if (eos_logic) {
      const long long tok0 = __shfl_sync(0xffffffffu, tok, 0);
      int stop = p.stopping[b];
      if (tok0 == p.eos) { rem = min(rem, nq); stop = 1; if (lane == 0) p.stopping[b] = 1; }
      if (stop) {
        const int idx = min(nq - rem, nq - 1);
        if (lane < idx) tok = p.mask;
        if (lane == idx) tok = p.eos;
      }
    }
    if (wcol < L) {
      long long* cell = p.delayed + ((size_t)b * nq + lane) * L + wcol;
      const bool unk = lane < nq && *cell == -1;
      const unsigned m = __ballot_sync(0xffffffffu, unk);
      const int j = __popc(m & ((1u << lane) - 1u));
      const long long routed = __shfl_sync(0xffffffffu, tok, j);
      if (unk) *cell = compat ? routed : tok;
    }
    if (eos_logic) {
      rem -= 1;
      if (lane == 0) {
        p.remaining[b] = rem;
        if (rem <= 0) p.out_len[b] = wcol - nq;      // out_codes[..., :offset - 9], offset = wcol
      }
    }
    if (rem > 0 && lane == 0) atomicOr(&s_alive, 1);
  }
  Sync::sync();
  if (threadIdx.x == 0) {
    st->col = wcol;
    st->pos += 1;
    st->step += 1;
    if (eos_logic && !s_alive) st->done = 1;
  }
#pragma unroll 1
  for (int r = threadIdx.x; r < p.rows; r += nthreads) p.rope_pos[r] += 1;
}

__global__ void k_update(const UpdateParams p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  __shared__ int s_alive;
  update_body<SyncCta>(p, &s_alive, (int)blockDim.x);
}

// ---------------------------------------------------------------------------------------------
// Delay pattern (codebook_pattern.py:5-12) — pure index kernels, int64, bit-exact.
// ---------------------------------------------------------------------------------------------
__global__ void k_delay_apply(const long long* __restrict__ codes, long long* __restrict__ out, int B, int nq, int T, long long mask) {
  const int L = T + nq;
  const size_t n = (size_t)B * nq * L;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % L);
    const int k = (int)((i / L) % nq);
    const size_t b = i / ((size_t)L * nq);
    int src = t - (k + 1);
    if (src < 0) src += L;                                  // roll by k+1 over the padded length
    out[i] = src < T ? codes[(b * nq + k) * T + src] : mask;
  }
}
__global__ void k_delay_revert(const long long* __restrict__ delayed, long long* __restrict__ out, int B, int nq, int L) {
  const int T = L - nq;
  const size_t n = (size_t)B * nq * T;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % T);
    const int k = (int)((i / T) % nq);
    const size_t b = i / ((size_t)T * nq);
    out[i] = delayed[(b * nq + k) * L + t + k + 1];
  }
}
