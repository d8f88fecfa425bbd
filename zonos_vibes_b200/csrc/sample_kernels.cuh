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
#pragma unroll
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

template <class Sync>
__device__ __forceinline__ float block_reduce_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  Sync::sync();
  if (lane == 0) red[warp] = v;
  Sync::sync();
  float t = 0.f;
  for (int w = 0; w < SMP_THREADS / 32; ++w) t += red[w];
  return t;
}
template <class Sync>
__device__ __forceinline__ float block_reduce_max(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_max(v);
  Sync::sync();
  if (lane == 0) red[warp] = v;
  Sync::sync();
  float t = -INFINITY;
  for (int w = 0; w < SMP_THREADS / 32; ++w) t = fmaxf(t, red[w]);
  return t;
}
// argmax with first-index tie break (torch.argmax)
template <class Sync>
__device__ __forceinline__ int block_argmax(float v, int idx, float* red, int* redi) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
  }
  Sync::sync();
  if (lane == 0) { red[warp] = v; redi[warp] = idx; }
  Sync::sync();
  float bv = red[0]; int bi = redi[0];
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
};

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

  // ---- repetition penalty (sampling.py:99-114): factor = penalty^(#occurrences in the window), ids clamped to V-1
  float x[SMP_PER];
#pragma unroll
  for (int j = 0; j < SMP_PER; ++j) {
    const int i = tid + j * SMP_THREADS;
    x[j] = i < V ? __ldcg(lg + i) : -INFINITY;
  }
  // factor = penalty^count by repeated multiplication, exactly like scatter_reduce(prod) over ones.
  if (sp.repetition_penalty != 1.0f && p.history != nullptr && hist_len > 0) {
    const int win = min(sp.repetition_penalty_window, hist_len);
    const long long* h = p.history + ((size_t)b * p.n_q + k) * p.hist_stride;
#pragma unroll
    for (int j = 0; j < SMP_PER; ++j) {
      const int i = tid + j * SMP_THREADS;
      if (i < V) {
        float f = 1.0f;
        for (int w = 0; w < win; ++w) {
          long long tok = __ldcg(h + hist_len - win + w);
          if (tok > V - 1) tok = V - 1;
          if ((long long)i == tok) f *= sp.repetition_penalty;
        }
        const float raw = __ldcg(lg + i);
        x[j] = raw <= 0.f ? raw * f : raw / f;
      }
    }
  }

  int token;
  if (sp.temperature > 0.f) {
    // ---- softmax(logits / T)
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < SMP_PER; ++j) { x[j] = x[j] / sp.temperature; mx = fmaxf(mx, x[j]); }
    mx = block_reduce_max<Sync>(mx, red);
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < SMP_PER; ++j) { x[j] = (tid + j * SMP_THREADS < V) ? expf(x[j] - mx) : 0.f; sum += x[j]; }
    sum = block_reduce_sum<Sync>(sum, red);
#pragma unroll
    for (int j = 0; j < SMP_PER; ++j) x[j] = x[j] / sum;

    // ---- unified sampler (sampling.py:29-43)
    if (sp.linear > 0.f) {
      float ent = 0.f;
      float lp[SMP_PER];
#pragma unroll
      for (int j = 0; j < SMP_PER; ++j) {
        lp[j] = logf(fmaxf(x[j], 1e-20f));
        if (tid + j * SMP_THREADS < V) ent += x[j] * lp[j];
      }
      ent = -block_reduce_sum<Sync>(ent, red);
      float mx2 = -INFINITY;
#pragma unroll
      for (int j = 0; j < SMP_PER; ++j) {
        const float raw = lp[j] * (sp.linear + ent * sp.conf) - lp[j] * lp[j] * sp.quad;
        x[j] = (tid + j * SMP_THREADS < V) ? raw : -INFINITY;
        mx2 = fmaxf(mx2, x[j]);
      }
      mx2 = block_reduce_max<Sync>(mx2, red);
      float sum2 = 0.f;
#pragma unroll
      for (int j = 0; j < SMP_PER; ++j) { x[j] = (tid + j * SMP_THREADS < V) ? expf(x[j] - mx2) : 0.f; sum2 += x[j]; }
      sum2 = block_reduce_sum<Sync>(sum2, red);
#pragma unroll
      for (int j = 0; j < SMP_PER; ++j) x[j] = x[j] / sum2;
    }

    // ---- top-p / top-k need the descending order: bitonic sort of (prob, index) in shared memory
    if (sp.top_p > 0.f || sp.top_k > 0) {
#pragma unroll
      for (int j = 0; j < SMP_PER; ++j) {
        const int i = tid + j * SMP_THREADS;
        s_key[i] = i < V ? x[j] : -1.f;          // padding sorts last
        s_idx[i] = i;
        s_p[i] = i < V ? x[j] : 0.f;
      }
      Sync::sync();
      for (int size = 2; size <= SMP_MAXV; size <<= 1) {
        for (int strd = size >> 1; strd > 0; strd >>= 1) {
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
          for (int r = 0; r < V; ++r) {
            c += (double)s_key[r];
            const float incl = (float)c;
            const bool drop = (incl - s_key[r]) > sp.top_p;
            if (drop) s_p[s_idx[r]] = 0.f;
          }
        }
        Sync::sync();
        float sum3 = 0.f;
#pragma unroll
        for (int j = 0; j < SMP_PER; ++j) { const int i = tid + j * SMP_THREADS; x[j] = i < V ? s_p[i] : 0.f; sum3 += x[j]; }
        sum3 = block_reduce_sum<Sync>(sum3, red);
#pragma unroll
        for (int j = 0; j < SMP_PER; ++j) { x[j] = x[j] / sum3; const int i = tid + j * SMP_THREADS; if (i < V) s_p[i] = x[j]; }
        Sync::sync();
      }
      if (sp.top_k > 0) {
        // pivot = k-th largest CURRENT probability; order is unchanged by the monotone steps above
        if (tid == 0) *sm.s_kth = s_p[s_idx[min(sp.top_k, V) - 1]];
        Sync::sync();
        const float kth = *sm.s_kth;
        float sum4 = 0.f;
#pragma unroll
        for (int j = 0; j < SMP_PER; ++j) { if (x[j] < kth) x[j] = 0.f; sum4 += x[j]; }
        sum4 = block_reduce_sum<Sync>(sum4, red);
#pragma unroll
        for (int j = 0; j < SMP_PER; ++j) x[j] = x[j] / sum4;
      }
    }
    // ---- min-p (sampling.py:82-96)
    if (sp.min_p > 0.f) {
      float top = 0.f;
#pragma unroll
      for (int j = 0; j < SMP_PER; ++j) top = fmaxf(top, x[j]);
      top = block_reduce_max<Sync>(top, red);
      const float thr = sp.min_p * top;
      float sum5 = 0.f;
#pragma unroll
      for (int j = 0; j < SMP_PER; ++j) { if (x[j] < thr) x[j] = 0.f; sum5 += x[j]; }
      sum5 = block_reduce_sum<Sync>(sum5, red);
#pragma unroll
      for (int j = 0; j < SMP_PER; ++j) x[j] = x[j] / sum5;
    }
    if (p.probs_out) {
#pragma unroll
      for (int j = 0; j < SMP_PER; ++j) {
        const int i = tid + j * SMP_THREADS;
        if (i < V) p.probs_out[((size_t)b * p.n_q + k) * V + i] = x[j];
      }
    }
    // ---- exponential race (sampling.py:19-21): argmax_i p_i / q_i, q_i ~ Exp(1) = -log(u_i)
    float best = -1.f; int besti = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < SMP_PER; ++j) {
      const int i = tid + j * SMP_THREADS;
      if (i < V) {
        const uint4 r = philox4x32_10(make_uint4((uint32_t)(i >> 2), (uint32_t)(b * p.n_q + k), (uint32_t)step, (uint32_t)(step >> 32)),
                                      make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
        const uint32_t bits = (i & 3) == 0 ? r.x : (i & 3) == 1 ? r.y : (i & 3) == 2 ? r.z : r.w;
        const float qe = -logf(u01(bits));
        const float sc = x[j] / fmaxf(qe, 1e-30f);
        if (sc > best || (sc == best && i < besti)) { best = sc; besti = i; }
      }
    }
    token = block_argmax<Sync>(best, besti, red, redi);
  } else {
    // greedy: argmax of the (penalised) logits, first index on ties (sampling.py:180)
    float best = -INFINITY; int besti = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < SMP_PER; ++j) {
      const int i = tid + j * SMP_THREADS;
      if (i < V && (x[j] > best || (x[j] == best && i < besti) || besti == 0x7fffffff)) { best = x[j]; besti = i; }
    }
    token = block_argmax<Sync>(best, besti, red, redi);
    if (p.probs_out) {
#pragma unroll
      for (int j = 0; j < SMP_PER; ++j) {
        const int i = tid + j * SMP_THREADS;
        if (i < V) p.probs_out[((size_t)b * p.n_q + k) * V + i] = x[j];
      }
    }
  }
  if (tid == 0) p.tokens[(size_t)b * p.n_q + k] = token;
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

// One block; threads 0..nthreads-1 all call it.  `s_alive` is one int of shared memory.
template <class Sync>
__device__ void update_body(const UpdateParams& p, int* s_alive_p, const int nthreads) {
  ZbDevState* st = p.st;
  if (st->done) return;
  int& s_alive = *s_alive_p;
  if (threadIdx.x == 0) s_alive = 0;
  Sync::sync();
  const int B = st->B, L = st->L, col = st->col, wcol = col + 1, nq = p.n_q;
  for (int b = threadIdx.x; b < B; b += nthreads) {
    int rem = p.remaining[b];
    if (rem > 0) {                                    // finished utterances are frozen (independent batch-1 runs)
      long long tok[16];
      for (int k = 0; k < nq; ++k) tok[k] = __ldcg(p.tokens + (size_t)b * nq + k);
      if (st->eos_logic) {
        if (tok[0] == p.eos) { rem = min(rem, nq); p.stopping[b] = 1; }
        if (p.stopping[b]) {
          const int idx = min(nq - rem, nq - 1);
          for (int k = 0; k < idx; ++k) tok[k] = p.mask;
          tok[idx] = p.eos;
        }
      }
      if (wcol < L) {
        int j = 0;
        for (int k = 0; k < nq; ++k) {
          long long* cell = p.delayed + ((size_t)b * nq + k) * L + wcol;
          if (*cell == -1) { *cell = st->compat_scatter ? tok[j] : tok[k]; ++j; }
        }
      }
      if (st->eos_logic) {
        rem -= 1;
        p.remaining[b] = rem;
        if (rem <= 0) p.out_len[b] = wcol - nq;      // out_codes[..., :offset - 9], offset = wcol
      }
      if (rem > 0) atomicOr(&s_alive, 1);
    }
  }
  Sync::sync();
  if (threadIdx.x == 0) {
    st->col = wcol;
    st->pos += 1;
    st->step += 1;
    if (st->eos_logic && !s_alive) st->done = 1;
  }
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
