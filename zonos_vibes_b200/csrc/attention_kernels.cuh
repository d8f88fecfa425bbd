// Split-KV GQA decode attention over the paged bf16 KV cache (reference: F.scaled_dot_product_attention with
// enable_gqa, zonos/backbone/_torch.py:131-136; S=1 so no causal mask, all ctx = pos+1 cached tokens attended).
//
// HBM-bound: 2 * ctx * Hkv * Dh * 2 B per row per layer.  grid = (nsplit, Hkv, rows); each CTA owns a token range
// of one (row, kv-head), all G = H/Hkv query heads share the K/V it streams.  A half-warp owns one token (16 lanes
// x 16 B = one 256 B K or V row, coalesced), a warp two, and every lane keeps 4 token-pairs (8 x 128-bit loads) in
// flight.  Softmax is the online (running max / sum) form in fp32 with warp-shuffle reductions; partials of the
// splits are merged by the last CTA to finish for that (row, kv-head) (ticket counter), so one launch does it all.
#pragma once
#include "zb_common.cuh"

#define ATT_THREADS 128
#define ATT_WARPS (ATT_THREADS / 32)
#define ATT_UB 4       // token-pair iterations in flight per warp

struct AttnParams {
  const bf16* q;            // [rows][H*Dh] post-RoPE
  const bf16* kv_layer;     // [page][2][Hkv][ZB_PAGE][ZB_DH]
  const int* page_table;    // [rows][max_pages]
  int max_pages;
  int H, Hkv;
  bf16* y;                  // [rows][H*Dh]
  float* part_o;            // [rows][Hkv][nsplit][G][Dh]
  float* part_ml;           // [rows][Hkv][nsplit][G][2]
  int* counters;            // [rows][Hkv], zero before launch; reset by the merging CTA
  const ZbDevState* st;
  float scale_log2;         // (1/sqrt(Dh)) * log2(e)
};

__device__ __forceinline__ void att_unpack8(const uint4& raw, float (&v)[8]) {
  v[0] = bf16lo(raw.x); v[1] = bf16hi(raw.x); v[2] = bf16lo(raw.y); v[3] = bf16hi(raw.y);
  v[4] = bf16lo(raw.z); v[5] = bf16hi(raw.z); v[6] = bf16lo(raw.w); v[7] = bf16hi(raw.w);
}

// Four tokens per half-warp at once: the scores of all of them are reduced together (16 independent shuffle chains
// instead of 4 dependent rounds of online softmax), then ONE max / rescale / accumulate step.
template <int G>
__device__ __forceinline__ void att_tokens4(const uint4 (&kraw)[4], const uint4 (&vraw)[4], const bool (&valid)[4], const float (&q)[G][8],
                                               float (&m)[G], float (&l)[G], float (&acc)[G][8], float scale_log2) {
  float s[4][G];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float kf[8];
    att_unpack8(kraw[k], kf);
#pragma unroll
    for (int g = 0; g < G; ++g) {
      float d = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) d = fmaf(q[g][j], kf[j], d);
      s[k][g] = d;
    }
  }
#pragma unroll
  for (int o = 8; o > 0; o >>= 1)
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int g = 0; g < G; ++g) s[k][g] += __shfl_xor_sync(0xffffffffu, s[k][g], o);
  if (!(valid[0] || valid[1] || valid[2] || valid[3])) return;     // uniform per half-warp; no shuffles below
  float vf[4][8];
#pragma unroll
  for (int k = 0; k < 4; ++k) att_unpack8(vraw[k], vf[k]);
#pragma unroll
  for (int g = 0; g < G; ++g) {
    float mn = m[g];
#pragma unroll
    for (int k = 0; k < 4; ++k) { s[k][g] *= scale_log2; if (valid[k]) mn = fmaxf(mn, s[k][g]); }
    const float corr = exp2f(m[g] - mn);           // m = -inf before the first token -> 0 (mn is finite here)
    float pr[4], ps = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) { pr[k] = valid[k] ? exp2f(s[k][g] - mn) : 0.f; ps += pr[k]; }
    l[g] = l[g] * corr + ps;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float a = acc[g][j] * corr;
#pragma unroll
      for (int k = 0; k < 4; ++k) a = fmaf(pr[k], vf[k][j], a);
      acc[g][j] = a;
    }
    m[g] = mn;
  }
}

template <int G>
__global__ void __launch_bounds__(ATT_THREADS) k_attn_decode(const AttnParams p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;

  const int split = blockIdx.x, nsplit = gridDim.x, kvh = blockIdx.y, row = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int half = lane >> 4, l16 = lane & 15;
  const int ctx = p.st->pos + 1;
  int chunk = (ctx + nsplit - 1) / nsplit;
  chunk = (chunk + 7) & ~7;
  const int t0 = split * chunk;
  const int t1 = min(ctx, t0 + chunk);

  // q for the G heads of this kv head: lane keeps dims [l16*8, l16*8+8)
  float q[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const uint4 raw = *reinterpret_cast<const uint4*>(p.q + (size_t)row * p.H * ZB_DH + (size_t)(kvh * G + g) * ZB_DH + l16 * 8);
    q[g][0] = bf16lo(raw.x); q[g][1] = bf16hi(raw.x); q[g][2] = bf16lo(raw.y); q[g][3] = bf16hi(raw.y);
    q[g][4] = bf16lo(raw.z); q[g][5] = bf16hi(raw.z); q[g][6] = bf16lo(raw.w); q[g][7] = bf16hi(raw.w);
  }
  float m[G], l[G], acc[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    m[g] = -INFINITY; l[g] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[g][j] = 0.f;
  }

  const int* pt = p.page_table + (size_t)row * p.max_pages;
  const size_t head_stride = (size_t)ZB_PAGE * ZB_DH;             // elements between kv heads in a page
  const size_t kv_stride = (size_t)p.Hkv * head_stride;           // K -> V inside a page
  const size_t page_stride = 2 * kv_stride;

  // warp w handles token pairs (t0 + 2*(w + ATT_WARPS*i)); batches of ATT_UB pairs, double-buffered: the loads of the
  // next batch are in flight while the current one is reduced (16 x 128-bit loads per lane outstanding).
  auto load_batch = [&](int base, uint4 (&kr)[ATT_UB], uint4 (&vr)[ATT_UB], bool (&vl)[ATT_UB]) {
#pragma unroll
    for (int i = 0; i < ATT_UB; ++i) {
      const int t = base + 2 * ATT_WARPS * i + half;
      vl[i] = t < t1;
      kr[i] = make_uint4(0, 0, 0, 0);
      vr[i] = make_uint4(0, 0, 0, 0);
      if (vl[i]) {
        const int page = pt[t / ZB_PAGE];
        const bf16* kp = p.kv_layer + (size_t)page * page_stride + (size_t)kvh * head_stride + (size_t)(t % ZB_PAGE) * ZB_DH + l16 * 8;
        kr[i] = ldg_stream(kp);
        vr[i] = ldg_stream(kp + kv_stride);
      }
    }
  };
  const int bstep = 2 * ATT_WARPS * ATT_UB;
  uint4 ka[ATT_UB], va[ATT_UB], kb[ATT_UB], vb[ATT_UB];
  bool la[ATT_UB], lb[ATT_UB];
  int base = t0 + 2 * warp;
  if (base < t1) load_batch(base, ka, va, la);
  while (base < t1) {
    if (base + bstep < t1) load_batch(base + bstep, kb, vb, lb);
    att_tokens4<G>(ka, va, la, q, m, l, acc, p.scale_log2);
    base += bstep;
    if (base < t1) {
      if (base + bstep < t1) load_batch(base + bstep, ka, va, la);
      att_tokens4<G>(kb, vb, lb, q, m, l, acc, p.scale_log2);
      base += bstep;
    }
  }

  // merge the two half-warps (they hold disjoint token subsets, same dims)
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const float mo = __shfl_xor_sync(0xffffffffu, m[g], 16);
    const float lo = __shfl_xor_sync(0xffffffffu, l[g], 16);
    const float mn = fmaxf(m[g], mo);
    const float cs = (m[g] == -INFINITY) ? 0.f : exp2f(m[g] - mn);
    const float co = (mo == -INFINITY) ? 0.f : exp2f(mo - mn);
    l[g] = l[g] * cs + lo * co;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float ao = __shfl_xor_sync(0xffffffffu, acc[g][j], 16);
      acc[g][j] = acc[g][j] * cs + ao * co;
    }
    m[g] = mn;
  }

  // merge warps through shared memory
  __shared__ float s_o[ATT_WARPS][G][ZB_DH];
  __shared__ float s_m[ATT_WARPS][G], s_l[ATT_WARPS][G];
  __shared__ int s_last;
  if (half == 0) {
#pragma unroll
    for (int g = 0; g < G; ++g) {
#pragma unroll
      for (int j = 0; j < 8; ++j) s_o[warp][g][l16 * 8 + j] = acc[g][j];
      if (l16 == 0) { s_m[warp][g] = m[g]; s_l[warp][g] = l[g]; }
    }
  }
  __syncthreads();
  // thread `tid` owns dim tid (ATT_THREADS == ZB_DH)
  const size_t pbase = (((size_t)row * p.Hkv + kvh) * nsplit + split) * G;
#pragma unroll
  for (int g = 0; g < G; ++g) {
    float mn = -INFINITY;
#pragma unroll
    for (int w = 0; w < ATT_WARPS; ++w) mn = fmaxf(mn, s_m[w][g]);
    float lt = 0.f, ot = 0.f;
#pragma unroll
    for (int w = 0; w < ATT_WARPS; ++w) {
      const float c = (s_m[w][g] == -INFINITY) ? 0.f : exp2f(s_m[w][g] - mn);
      lt += s_l[w][g] * c;
      ot += s_o[w][g][tid] * c;
    }
    if (nsplit == 1) {                               // the only split: this IS the attention output (no partials, no ticket)
      p.y[(size_t)row * p.H * ZB_DH + (size_t)(kvh * G + g) * ZB_DH + tid] = __float2bfloat16_rn(ot / lt);
    } else {
      p.part_o[(pbase + g) * ZB_DH + tid] = ot;
      if (tid == 0) { p.part_ml[(pbase + g) * 2 + 0] = mn; p.part_ml[(pbase + g) * 2 + 1] = lt; }
    }
  }
  if (nsplit == 1) return;
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const int ticket = atomicAdd(&p.counters[row * p.Hkv + kvh], 1);
    s_last = (ticket == nsplit - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // ---- merge of the split partials by the last CTA.  Only the first `nact` splits saw tokens.  (m, l) of all splits are
  // staged in shared memory with one parallel load; the partial outputs are then read 8 splits at a time so the L2
  // round trips overlap instead of serialising (the first version of this loop cost 26 us per launch, r01 launch list).
  __shared__ float s_ml[32 * 8 * 2];
  const int nact = min(nsplit, (ctx + chunk - 1) / chunk);
  const size_t rbase = ((size_t)row * p.Hkv + kvh) * nsplit * G;
  for (int i = tid; i < nact * G * 2; i += ATT_THREADS) s_ml[i] = __ldcg(&p.part_ml[rbase * 2 + i]);
  __syncthreads();
#pragma unroll
  for (int g = 0; g < G; ++g) {
    float mn = -INFINITY;
    for (int s2 = 0; s2 < nact; ++s2) mn = fmaxf(mn, s_ml[(s2 * G + g) * 2]);
    float lt = 0.f, ot = 0.f;
    for (int s0 = 0; s0 < nact; s0 += 8) {
      float o[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) o[u] = (s0 + u < nact) ? __ldcg(&p.part_o[(rbase + (size_t)(s0 + u) * G + g) * ZB_DH + tid]) : 0.f;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (s0 + u < nact) {
          const float ms = s_ml[((s0 + u) * G + g) * 2];
          const float c = (ms == -INFINITY) ? 0.f : exp2f(ms - mn);
          lt += s_ml[((s0 + u) * G + g) * 2 + 1] * c;
          ot += o[u] * c;
        }
      }
    }
    p.y[(size_t)row * p.H * ZB_DH + (size_t)(kvh * G + g) * ZB_DH + tid] = __float2bfloat16_rn(ot / lt);
  }
  if (tid == 0) p.counters[row * p.Hkv + kvh] = 0;
}
