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

  // warp w handles token pairs (t0 + 2*(w + ATT_WARPS*i)) ; batches of ATT_UB pairs
  for (int base = t0 + 2 * warp; base < t1; base += 2 * ATT_WARPS * ATT_UB) {
    uint4 kraw[ATT_UB], vraw[ATT_UB];
    bool valid[ATT_UB];
#pragma unroll
    for (int i = 0; i < ATT_UB; ++i) {
      const int t = base + 2 * ATT_WARPS * i + half;
      valid[i] = t < t1;
      kraw[i] = make_uint4(0, 0, 0, 0);
      vraw[i] = make_uint4(0, 0, 0, 0);
      if (valid[i]) {
        const int page = pt[t / ZB_PAGE];
        const bf16* kp = p.kv_layer + (size_t)page * page_stride + (size_t)kvh * head_stride + (size_t)(t % ZB_PAGE) * ZB_DH + l16 * 8;
        kraw[i] = ldg_stream(kp);
        vraw[i] = ldg_stream(kp + kv_stride);
      }
    }
#pragma unroll
    for (int i = 0; i < ATT_UB; ++i) {
      const float kf[8] = {bf16lo(kraw[i].x), bf16hi(kraw[i].x), bf16lo(kraw[i].y), bf16hi(kraw[i].y),
                           bf16lo(kraw[i].z), bf16hi(kraw[i].z), bf16lo(kraw[i].w), bf16hi(kraw[i].w)};
      const float vf[8] = {bf16lo(vraw[i].x), bf16hi(vraw[i].x), bf16lo(vraw[i].y), bf16hi(vraw[i].y),
                           bf16lo(vraw[i].z), bf16hi(vraw[i].z), bf16lo(vraw[i].w), bf16hi(vraw[i].w)};
#pragma unroll
      for (int g = 0; g < G; ++g) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) s = fmaf(q[g][j], kf[j], s);
        // sum over the 16 lanes of this half-warp
        s += __shfl_xor_sync(0xffffffffu, s, 8);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        if (valid[i]) {
          s *= p.scale_log2;
          const float mn = fmaxf(m[g], s);
          const float corr = exp2f(m[g] - mn);       // m = -inf on first token -> 0
          const float pr = exp2f(s - mn);
          l[g] = l[g] * corr + pr;
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[g][j] = fmaf(pr, vf[j], acc[g][j] * corr);
          m[g] = mn;
        }
      }
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
    p.part_o[(pbase + g) * ZB_DH + tid] = ot;
    if (tid == 0) { p.part_ml[(pbase + g) * 2 + 0] = mn; p.part_ml[(pbase + g) * 2 + 1] = lt; }
  }
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
