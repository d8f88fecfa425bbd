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
  int total_pages;          // pages of one layer's cache (checked build: bound of every page-table entry)
  int vrows;                // 0: row = cache row, ctx = pos + 1.  > 0 (block prefill): activation row a is token a / vrows of cache
                            // row a % vrows and attends positions [0, pos + a / vrows] - causal attention over the block
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

// Cross-warp merge of the per-warp online-softmax states in shared memory, then (nsplit > 1) the ticketed merge of the
// split partials by the last CTA of the (row, kv-head).  Shared by the SIMT and the mma.sync kernels.
// `split` of `nparts` ticket holders writes partial slot `split` of the (row, kv head)'s `stride` slots; the last ticket merges
// the first `nact` slots.
template <int G>
__device__ __forceinline__ void att_finish_parts(const AttnParams& p, int row, int kvh, int split, int stride, int nparts, int nact,
                                                 float (*s_o)[G][ZB_DH], float (*s_m)[G], float (*s_l)[G]) {
  __shared__ int s_last;
  const int tid = threadIdx.x;
  const int nsplit = nparts;
  __syncthreads();
  // thread `tid` owns dim tid (ATT_THREADS == ZB_DH)
  const size_t pbase = (((size_t)row * p.Hkv + kvh) * stride + split) * G;
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
  const size_t rbase = ((size_t)row * p.Hkv + kvh) * stride * G;
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
template <int G>
__device__ __forceinline__ void att_finish(const AttnParams& p, int row, int kvh, int split, int nsplit, int ctx, int chunk,
                                           float (*s_o)[G][ZB_DH], float (*s_m)[G], float (*s_l)[G]) {
  att_finish_parts<G>(p, row, kvh, split, nsplit, nsplit, min(nsplit, (ctx + chunk - 1) / chunk), s_o, s_m, s_l);
}

template <int G>
__global__ void __launch_bounds__(ATT_THREADS) k_attn_decode(const AttnParams p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;

  const int split = blockIdx.x, nsplit = gridDim.x, kvh = blockIdx.y, row = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int half = lane >> 4, l16 = lane & 15;
  const int crow = p.vrows ? row % p.vrows : row;
  const int ctx = p.st->pos + 1 + (p.vrows ? row / p.vrows : 0);
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

  const int* pt = p.page_table + (size_t)crow * p.max_pages;
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
        ZB_ASSERT(t / ZB_PAGE < p.max_pages);
        const int page = pt[t / ZB_PAGE];
        ZB_ASSERT(page >= 0 && page < p.total_pages);
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
  if (half == 0) {
#pragma unroll
    for (int g = 0; g < G; ++g) {
#pragma unroll
      for (int j = 0; j < 8; ++j) s_o[warp][g][l16 * 8 + j] = acc[g][j];
      if (l16 == 0) { s_m[warp][g] = m[g]; s_l[warp][g] = l[g]; }
    }
  }
  att_finish<G>(p, row, kvh, split, nsplit, ctx, chunk, s_o, s_m, s_l);
}

// =====================================================================================================================
// mma.sync variant.  The SIMT kernel above spends ~65 warp instructions per cached token (bf16 unpack, 1024 FMAs, 16
// shuffle chains), which makes long-context decode issue-bound at ~40 % of the KV stream's HBM time once 128 rows are in
// flight (B = 64).  Here a warp owns 16 tokens at a time and both contractions run on the tensor cores straight from
// the registers the coalesced 128-bit global loads land in — no shared-memory staging, no ldmatrix:
//   S[head, tok]   = Q K^T     m16n8k16, A = Q (rows = the G <= 8 heads of the kv head), B = K.  The contraction index
//                              is a free permutation of the 128 dims, so lane (g, t) feeds the 8 consecutive dims of
//                              its 16-byte load of token g as the (k, k+1 | k+8, k+9) pairs of two k-steps and Q is
//                              loaded once with the same permutation.
//   O^T[dim, head] = V^T P^T   A = V^T: lane (g, t) loads dims [16g, 16g+16) of tokens {2t, 2t+1, 2t+8, 2t+9}; one
//                              PRMT pairs the two tokens of a dim into an A register (the dim order of O is a free
//                              permutation too); B = P^T is exactly the layout S's accumulators already have.
// 24 HMMA + 32 PRMT + a 4-lane softmax reduction per 16 tokens: ~8 warp instructions per token.  P is rounded to bf16
// before the PV product (as flash-attention / SDPA's fused kernels do); m, l and O stay fp32.
__device__ __forceinline__ void att_mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t att_pack_bf16(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}

struct AttFrag {
  uint4 k[2][4];      // token g (+8): dims 32s + 8t .. +8
  uint4 v[4][2];      // tokens 2t, 2t+1, 2t+8, 2t+9: dims 16g .. +16
};

// per-warp running state: lanes 4h..4h+3 hold (m, l) of head h; o[j] = O^T tile j (dims 16g + 2j, 16g + 2j + 1; heads 2t, 2t+1)
template <int G>
__device__ __forceinline__ void att_group16(const AttFrag& f, int tb, int t1, const uint32_t (&qa)[8][2], float& m, float& l,
                                            float (&o)[8][4], float scale_log2, int lane) {
  const int t = lane & 3;
  float c0[4] = {0.f, 0.f, 0.f, 0.f}, c1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    att_mma(c0, qa[2 * s][0], 0u, qa[2 * s][1], 0u, f.k[0][s].x, f.k[0][s].y);
    att_mma(c1, qa[2 * s][0], 0u, qa[2 * s][1], 0u, f.k[1][s].x, f.k[1][s].y);
    att_mma(c0, qa[2 * s + 1][0], 0u, qa[2 * s + 1][1], 0u, f.k[0][s].z, f.k[0][s].w);
    att_mma(c1, qa[2 * s + 1][0], 0u, qa[2 * s + 1][1], 0u, f.k[1][s].z, f.k[1][s].w);
  }
  // scores of head g for tokens tb + {2t, 2t+1, 8+2t, 9+2t}
  const int ta = tb + 2 * t;
  float s00 = ta < t1 ? c0[0] * scale_log2 : -INFINITY, s01 = ta + 1 < t1 ? c0[1] * scale_log2 : -INFINITY;
  float s10 = ta + 8 < t1 ? c1[0] * scale_log2 : -INFINITY, s11 = ta + 9 < t1 ? c1[1] * scale_log2 : -INFINITY;
  float mx = fmaxf(fmaxf(s00, s01), fmaxf(s10, s11));
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
  const float mn = fmaxf(m, mx);                        // finite: token tb itself is valid
  const float corr = exp2f(m - mn);                     // m = -inf before the first group -> 0
  const float p00 = exp2f(s00 - mn), p01 = exp2f(s01 - mn), p10 = exp2f(s10 - mn), p11 = exp2f(s11 - mn);
  float ps = (p00 + p01) + (p10 + p11);
  ps += __shfl_xor_sync(0xffffffffu, ps, 1);
  ps += __shfl_xor_sync(0xffffffffu, ps, 2);
  l = l * corr + ps;
  m = mn;
  const uint32_t b0 = att_pack_bf16(p00, p01), b1 = att_pack_bf16(p10, p11);
  // O^T columns of this lane are heads 2t and 2t+1: fetch their rescale factors from the lanes that own those heads
  const float ca = __shfl_sync(0xffffffffu, corr, 8 * t), cb = __shfl_sync(0xffffffffu, corr, 8 * t + 4);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    o[j][0] *= ca; o[j][1] *= cb; o[j][2] *= ca; o[j][3] *= cb;
    const uint32_t w0 = j < 4 ? (&f.v[0][0].x)[j] : (&f.v[0][1].x)[j - 4], w1 = j < 4 ? (&f.v[1][0].x)[j] : (&f.v[1][1].x)[j - 4];
    const uint32_t w2 = j < 4 ? (&f.v[2][0].x)[j] : (&f.v[2][1].x)[j - 4], w3 = j < 4 ? (&f.v[3][0].x)[j] : (&f.v[3][1].x)[j - 4];
    att_mma(o[j], __byte_perm(w0, w1, 0x5410), __byte_perm(w0, w1, 0x7632), __byte_perm(w2, w3, 0x5410), __byte_perm(w2, w3, 0x7632), b0, b1);
  }
}

template <int G>
__global__ void __launch_bounds__(ATT_THREADS) k_attn_decode_mma(const AttnParams p) {
  static_assert(G <= 8, "the heads of one kv head are the n8 dimension of the PV product");
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;

  const int split = blockIdx.x, nsplit = gridDim.x, kvh = blockIdx.y, row = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int crow = p.vrows ? row % p.vrows : row;
  const int ctx = p.st->pos + 1 + (p.vrows ? row / p.vrows : 0);
  int chunk = (ctx + nsplit - 1) / nsplit;
  chunk = (chunk + 15) & ~15;                                     // 16-token groups never straddle a page (ZB_PAGE % 16 == 0)
  const int t0 = split * chunk;
  const int t1 = min(ctx, t0 + chunk);

  uint32_t qa[8][2];
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    uint4 w = make_uint4(0, 0, 0, 0);
    if (g < G) w = *reinterpret_cast<const uint4*>(p.q + (size_t)row * p.H * ZB_DH + (size_t)(kvh * G + g) * ZB_DH + 32 * s + 8 * t);
    qa[2 * s][0] = w.x; qa[2 * s][1] = w.y; qa[2 * s + 1][0] = w.z; qa[2 * s + 1][1] = w.w;
  }
  float m = -INFINITY, l = 0.f, o[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) { o[j][0] = 0.f; o[j][1] = 0.f; o[j][2] = 0.f; o[j][3] = 0.f; }

  const int* pt = p.page_table + (size_t)crow * p.max_pages;
  const size_t head_stride = (size_t)ZB_PAGE * ZB_DH;
  const size_t kv_stride = (size_t)p.Hkv * head_stride;
  const size_t page_stride = 2 * kv_stride;
  auto load_group = [&](int tb, AttFrag& f) {
    ZB_ASSERT(tb / ZB_PAGE < p.max_pages);
    const int page = pt[tb / ZB_PAGE];
    ZB_ASSERT(page >= 0 && page < p.total_pages);
    const bf16* kbase = p.kv_layer + (size_t)page * page_stride + (size_t)kvh * head_stride + (size_t)(tb % ZB_PAGE) * ZB_DH;
    const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int n = 0; n < 2; ++n) {
      const bool ok = tb + 8 * n + g < t1;
#pragma unroll
      for (int s = 0; s < 4; ++s) f.k[n][s] = ok ? ldg_stream(kbase + (size_t)(8 * n + g) * ZB_DH + 32 * s + 8 * t) : z;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int tok = 2 * t + (u & 1) + 8 * (u >> 1);
      const bool ok = tb + tok < t1;
#pragma unroll
      for (int h = 0; h < 2; ++h) f.v[u][h] = ok ? ldg_stream(kbase + kv_stride + (size_t)tok * ZB_DH + 16 * g + 8 * h) : z;
    }
  };
  // warp w handles groups t0 + 16*(w + ATT_WARPS*i); the next group's 16 loads are in flight while this one is reduced
  const int gstep = 16 * ATT_WARPS;
  AttFrag fa, fb;
  int tb = t0 + 16 * warp;
  if (tb < t1) load_group(tb, fa);
  while (tb < t1) {
    if (tb + gstep < t1) load_group(tb + gstep, fb);
    att_group16<G>(fa, tb, t1, qa, m, l, o, p.scale_log2, lane);
    tb += gstep;
    if (tb < t1) {
      if (tb + gstep < t1) load_group(tb + gstep, fa);
      att_group16<G>(fb, tb, t1, qa, m, l, o, p.scale_log2, lane);
      tb += gstep;
    }
  }

  __shared__ float s_o[ATT_WARPS][G][ZB_DH];
  __shared__ float s_m[ATT_WARPS][G], s_l[ATT_WARPS][G];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int d = 16 * g + 2 * j;
    if (2 * t < G) { s_o[warp][2 * t][d] = o[j][0]; s_o[warp][2 * t][d + 1] = o[j][2]; }
    if (2 * t + 1 < G) { s_o[warp][2 * t + 1][d] = o[j][1]; s_o[warp][2 * t + 1][d + 1] = o[j][3]; }
  }
  if (t == 0 && g < G) { s_m[warp][g] = m; s_l[warp][g] = l; }
  att_finish<G>(p, row, kvh, split, nsplit, ctx, chunk, s_o, s_m, s_l);
}

// =====================================================================================================================
// k_attn_decode_tma: the same split-KV attention with K / V brought in by the TMA engine (cp.async.bulk, 64 tokens =
// 16 KB of K + 16 KB of V per stage, ring of ATT_TMA_STAGES stages with full mbarriers) instead of per-lane ld.global.
// Why: scripts/ring_bench.cu measures ~50 GB/s per SM for bulk copies against ~38 GB/s per SM for the ld.global path; at
// batch 64 the ld.global kernel sits at 5.45 TB/s = 36.8 GB/s per SM (profiles/r02_summary.md), i.e. on that ceiling.
// One warp = one 16-token group of a stage (4 warps = the 4 groups); fragments come from shared memory in the mma.sync
// layout of att_group16.  A stage never straddles more than one page boundary (ZB_PAGE = 128, stages of 64 tokens): at
// most two copies per tensor per stage.
// =====================================================================================================================
#define ATT_TMA_TOK 64
#define ATT_TMA_STAGE_BYTES (2 * ATT_TMA_TOK * ZB_DH * 2)       // 32 KB
#define ATT_TMA_STAGES 3    // 3 stages = 96 KB = 2 CTAs per SM.  Measured at batch 64, ctx 1 300 (profiles/r02_attn_tma_ab.txt): 2 stages
                            // (3 CTAs per SM) 3 675 us per decode step, 4 stages (1 CTA per SM) 3 987, 3 stages 3 152-3 230
#define ATT_TMA_SMEM (ATT_TMA_STAGES * ATT_TMA_STAGE_BYTES + 64)

__device__ __forceinline__ uint32_t att_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool att_mbar_try(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ uint4 att_lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}

template <int G, bool EARLY>
__global__ void __launch_bounds__(ATT_THREADS) k_attn_decode_tma(const AttnParams p) {
  extern __shared__ __align__(128) unsigned char att_ring[];
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;

  const int split = blockIdx.x, nsplit = gridDim.x, kvh = blockIdx.y, row = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int crow = p.vrows ? row % p.vrows : row;
  const int ctx = p.st->pos + 1 + (p.vrows ? row / p.vrows : 0);
  int chunk = (ctx + nsplit - 1) / nsplit;
  chunk = (chunk + 15) & ~15;
  const int t0 = split * chunk;
  const int t1 = min(ctx, t0 + chunk);
  const int n_st = t1 > t0 ? (t1 - t0 + ATT_TMA_TOK - 1) / ATT_TMA_TOK : 0;

  const uint32_t ring = att_smem_u32(att_ring);
  const uint32_t bar0 = ring + ATT_TMA_STAGES * ATT_TMA_STAGE_BYTES;
  const int* pt = p.page_table + (size_t)crow * p.max_pages;
  const size_t head_stride = (size_t)ZB_PAGE * ZB_DH;
  const size_t kv_stride = (size_t)p.Hkv * head_stride;
  const size_t page_stride = 2 * kv_stride;

  // thread 0 is the producer: one stage = K rows [0, 16 KB) and V rows [16 KB, 32 KB) of tokens [ts, ts + n)
  auto issue = [&](int si) {
    const int slot = si % ATT_TMA_STAGES;
    const int ts = t0 + si * ATT_TMA_TOK;
    const int n = min(ATT_TMA_TOK, t1 - ts);
    const uint32_t bar = bar0 + 8u * slot, dst = ring + (uint32_t)slot * ATT_TMA_STAGE_BYTES;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)(n * ZB_DH * 2 * 2)) : "memory");
    const int n1 = min(n, ZB_PAGE - ts % ZB_PAGE);                 // tokens before the page boundary
    for (int run = 0; run < 2; ++run) {
      const int tb = run == 0 ? ts : ts + n1, nn = run == 0 ? n1 : n - n1;
      if (nn <= 0) continue;
      ZB_ASSERT(tb / ZB_PAGE < p.max_pages && pt[tb / ZB_PAGE] >= 0 && pt[tb / ZB_PAGE] < p.total_pages);
      ZB_ASSERT(tb % ZB_PAGE + nn <= ZB_PAGE && (tb - ts) + nn <= ATT_TMA_TOK && slot < ATT_TMA_STAGES);
      const bf16* src = p.kv_layer + (size_t)pt[tb / ZB_PAGE] * page_stride + (size_t)kvh * head_stride + (size_t)(tb % ZB_PAGE) * ZB_DH;
      const uint32_t off = (uint32_t)(tb - ts) * ZB_DH * 2;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst + off), "l"(src),
                   "r"((uint32_t)(nn * ZB_DH * 2)), "r"(bar) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst + ATT_TMA_TOK * ZB_DH * 2 + off),
                   "l"(src + kv_stride), "r"((uint32_t)(nn * ZB_DH * 2)), "r"(bar) : "memory");
    }
  };
  if (tid == 0) {
    for (int s = 0; s < ATT_TMA_STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar0 + 8u * s), "r"(1u) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async;" ::: "memory");             // this step's K / V rows were written by ordinary stores of the previous kernel
    for (int si = 0; si < min(n_st, ATT_TMA_STAGES); ++si) issue(si);
  }

  uint32_t qa[8][2];
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    uint4 w = make_uint4(0, 0, 0, 0);
    if (g < G) w = *reinterpret_cast<const uint4*>(p.q + (size_t)row * p.H * ZB_DH + (size_t)(kvh * G + g) * ZB_DH + 32 * s + 8 * t);
    qa[2 * s][0] = w.x; qa[2 * s][1] = w.y; qa[2 * s + 1][0] = w.z; qa[2 * s + 1][1] = w.w;
  }
  float m = -INFINITY, l = 0.f, o[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) { o[j][0] = 0.f; o[j][1] = 0.f; o[j][2] = 0.f; o[j][3] = 0.f; }
  __syncthreads();                                                  // barriers initialised before anyone polls them

  const uint4 z4 = make_uint4(0, 0, 0, 0);
  for (int si = 0; si < n_st; ++si) {
    const int slot = si % ATT_TMA_STAGES;
    const int ts = t0 + si * ATT_TMA_TOK;
    const int n = min(ATT_TMA_TOK, t1 - ts);
    if (lane == 0) { while (!att_mbar_try(bar0 + 8u * slot, (uint32_t)((si / ATT_TMA_STAGES) & 1))) {} }
    __syncwarp();
    const int tl = 16 * warp;                                       // this warp's 16-token group of the stage
    AttFrag f;
    if (tl < n) {
      const uint32_t sp = ring + (uint32_t)slot * ATT_TMA_STAGE_BYTES;
#pragma unroll
      for (int nn = 0; nn < 2; ++nn) {
        const int tok = tl + 8 * nn + g;
#pragma unroll
        for (int s = 0; s < 4; ++s) f.k[nn][s] = tok < n ? att_lds128(sp + tok * (ZB_DH * 2) + (32 * s + 8 * t) * 2) : z4;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int tok = tl + 2 * t + (u & 1) + 8 * (u >> 1);
#pragma unroll
        for (int h = 0; h < 2; ++h) f.v[u][h] = tok < n ? att_lds128(sp + ATT_TMA_TOK * ZB_DH * 2 + tok * (ZB_DH * 2) + (16 * g + 8 * h) * 2) : z4;
      }
    }
    // EARLY: the slot is handed back as soon as the fragments are in registers; otherwise after the group's math
    if (!EARLY && tl < n) att_group16<G>(f, ts + tl, t1, qa, m, l, o, p.scale_log2, lane);
    __syncthreads();                                                // every warp has its fragments out of the slot
    if (tid == 0 && si + ATT_TMA_STAGES < n_st) issue(si + ATT_TMA_STAGES);
    if (EARLY && tl < n) att_group16<G>(f, ts + tl, t1, qa, m, l, o, p.scale_log2, lane);
  }

  // every stage has been consumed (the loop ends on a __syncthreads): the ring is free to hold the per-warp states
  float (*s_o)[G][ZB_DH] = reinterpret_cast<float (*)[G][ZB_DH]>(att_ring);
  float (*s_m)[G] = reinterpret_cast<float (*)[G]>(att_ring + sizeof(float) * ATT_WARPS * G * ZB_DH);
  float (*s_l)[G] = s_m + ATT_WARPS;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int d = 16 * g + 2 * j;
    if (2 * t < G) { s_o[warp][2 * t][d] = o[j][0]; s_o[warp][2 * t][d + 1] = o[j][2]; }
    if (2 * t + 1 < G) { s_o[warp][2 * t + 1][d] = o[j][1]; s_o[warp][2 * t + 1][d + 1] = o[j][3]; }
  }
  if (t == 0 && g < G) { s_m[warp][g] = m; s_l[warp][g] = l; }
  att_finish<G>(p, row, kvh, split, nsplit, ctx, chunk, s_o, s_m, s_l);
}
