// Backbone decode-step kernels (transformer): weight-streaming GEMV family, split-KV GQA decode attention,
// embedding gather-sum.  Reference arithmetic: zonos/backbone/_torch.py:83-152, zonos/model.py:97-116.
//
// Roofline: every kernel here is HBM-bound: one decode step streams each weight exactly once
// (3.2 GB at the full config) plus the KV cache.  Design rules followed:
//   * 128-bit, fully coalesced, L1-bypassing loads (one weight row = contiguous 2K bytes; a warp reads
//     512 B per instruction), 8-16 loads in flight per lane through a register double buffer;
//   * two 256-thread CTAs per SM (grid = 296) with (row-pair) tasks dealt round-robin so per-SM bytes differ
//     by at most one task; CTA-granular exit lets the next kernel (PDL) start streaming on freed SM halves;
//   * activations (a few KB) are staged once per CTA in shared memory as bf16, LayerNorm fused into the staging;
//   * every kernel prefetches its first weights BEFORE `griddepcontrol.wait`, so with programmatic dependent
//     launch the HBM stream does not drain at kernel boundaries; optionally it also L2-prefetches the head of
//     the next kernel's weights.
#pragma once
#include "zb_common.cuh"

enum { EPI_QKV = 0, EPI_RESID = 1, EPI_SWIGLU = 2, EPI_HEADS = 3, EPI_PLAIN = 4 };   // PLAIN: out = bf16(W x) (MMA kernel only)

#define GEMV_THREADS 256
#define GEMV_CTAS_PER_SM 2
#define GEMV_U 4                 // 128-bit chunks per row per pipeline unit (x2 rows per task)
#define GEMV_MAX_LN_DIM 2048     // LayerNorm staging keeps one uint4 (8 elements) per thread per row

struct GemvParams {
  const bf16* W;                 // [rows][K] row-major bf16
  int K;                         // reduction length, multiple of 256
  int n_tasks;                   // tasks of two weight rows each
  int row0_mul;                  // row0 = task * row0_mul
  int row1_off;                  // row1 = row0 + row1_off
  const bf16* x;                 // activations [rows][K]
  // template row r maps to activation row (r < R/2 ? base0 + r : base1 + r - R/2), valid iff < lim0 / lim1.
  // CFG pass over utterances [b0, b0+R/2): base0=b0, lim0=B, base1=B+b0, lim1=2B (cond/uncond stay together).
  int base0, base1, lim0, lim1;
  const bf16* ln_w;              // LayerNorm affine (NULL = no LayerNorm, x is used as is)
  const bf16* ln_b;
  float eps;
  const ZbDevState* st;
  // --- epilogue targets
  bf16* out;                     // RESID: residual stream [R][N] (in place); SWIGLU: h [R][F]; QKV: q [R][q_size]
  int out_stride;                // row stride of `out`
  // QKV
  bf16* kv_layer;                // this layer's pages: [page][2][Hkv][ZB_PAGE][ZB_DH]
  const int* page_table;         // [R][max_pages]
  int max_pages;
  int Hkv;
  int q_size;                    // H*Dh
  int kv_size;                   // Hkv*Dh
  const float* rope;             // [table_len][Dh/2][2] fp32 (cos, sin)
  const int* rope_pos;           // [R]
  // HEADS
  float* logits;                 // [B][n_q][V]
  int n_q;
  int V;
  int eos;                       // logit_bias index (model.py:266-267)
  int n_rows;                    // total weight rows (MMA kernel: guards the ragged last 8-row task)
  // optional L2 prefetch of (the head of) the next kernel's weights: up to two byte ranges
  const char* pf0;
  const char* pf1;
  unsigned pf_bytes;             // per range, multiple of 16*gridDim.x
};

template <int R>
__device__ __forceinline__ int gemv_row(const GemvParams& p, int r, bool& valid) {
  const int a = r < R / 2 ? p.base0 + r : p.base1 + (r - R / 2);
  valid = a < (r < R / 2 ? p.lim0 : p.lim1);
  return a;
}

template <int R>
__device__ __forceinline__ void gemv_stage_activations(const GemvParams& p, bf16* xs, float* red, int ld) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  const int K = p.K;
  if (p.ln_w == nullptr) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      bool ok;
      const int a = gemv_row<R>(p, r, ok);
      for (int e = tid * 8; e < K; e += blockDim.x * 8)
        *reinterpret_cast<uint4*>(xs + (size_t)r * ld + e) =
            ok ? *reinterpret_cast<const uint4*>(p.x + (size_t)a * K + e) : make_uint4(0, 0, 0, 0);
    }
    return;
  }
  // LayerNorm (F.layer_norm semantics: biased variance, eps inside the sqrt, fp32 statistics,
  // result rounded to bf16 like the reference's bf16 module output).  K <= GEMV_MAX_LN_DIM.
  const int e = tid * 8;
  const bool active = e < K;
  float v[R][8];
  float s[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    uint4 raw = make_uint4(0, 0, 0, 0);
    bool ok;
    const int a = gemv_row<R>(p, r, ok);
    if (active && ok) raw = *reinterpret_cast<const uint4*>(p.x + (size_t)a * K + e);
    v[r][0] = bf16lo(raw.x); v[r][1] = bf16hi(raw.x); v[r][2] = bf16lo(raw.y); v[r][3] = bf16hi(raw.y);
    v[r][4] = bf16lo(raw.z); v[r][5] = bf16hi(raw.z); v[r][6] = bf16lo(raw.w); v[r][7] = bf16hi(raw.w);
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += v[r][j];
    s[r] = warp_sum(t);
  }
  if (lane == 0) {
#pragma unroll
    for (int r = 0; r < R; ++r) red[warp * R + r] = s[r];
  }
  __syncthreads();
  float mean[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    float t = 0.f;
    for (int w = 0; w < nwarp; ++w) t += red[w * R + r];
    mean[r] = t / (float)K;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < R; ++r) {
    float t = 0.f;
    if (active) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { float d = v[r][j] - mean[r]; t += d * d; }
    }
    s[r] = warp_sum(t);
  }
  if (lane == 0) {
#pragma unroll
    for (int r = 0; r < R; ++r) red[warp * R + r] = s[r];
  }
  __syncthreads();
  if (active) {
    uint4 wr = *reinterpret_cast<const uint4*>(p.ln_w + e);
    uint4 br = *reinterpret_cast<const uint4*>(p.ln_b + e);
    float g[8] = {bf16lo(wr.x), bf16hi(wr.x), bf16lo(wr.y), bf16hi(wr.y), bf16lo(wr.z), bf16hi(wr.z), bf16lo(wr.w), bf16hi(wr.w)};
    float b[8] = {bf16lo(br.x), bf16hi(br.x), bf16lo(br.y), bf16hi(br.y), bf16lo(br.z), bf16hi(br.z), bf16lo(br.w), bf16hi(br.w)};
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float t = 0.f;
      for (int w = 0; w < nwarp; ++w) t += red[w * R + r];
      const float rstd = rsqrtf(t / (float)K + p.eps);
      __nv_bfloat162 o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float a0 = (v[r][2 * j] - mean[r]) * rstd * g[2 * j] + b[2 * j];
        float a1 = (v[r][2 * j + 1] - mean[r]) * rstd * g[2 * j + 1] + b[2 * j + 1];
        o[j] = __floats2bfloat162_rn(a0, a1);
      }
      *reinterpret_cast<uint4*>(xs + (size_t)r * ld + e) = *reinterpret_cast<uint4*>(o);
    }
  }
}

template <int R, int EPI>
__device__ __forceinline__ void gemv_epilogue(const GemvParams& p, int task, int lane, const float (&a0)[R], const float (&a1)[R]) {
  const int n0 = task * p.row0_mul;
  if constexpr (EPI == EPI_RESID) {
    // x <- bf16(x + bf16(W a))   (_torch.py:100-101, bf16 residual stream)
#pragma unroll
    for (int r = 0; r < R; ++r) {
      bool ok;
      const int a = gemv_row<R>(p, r, ok);
      if (lane == r && ok) {
        __nv_bfloat162* px = reinterpret_cast<__nv_bfloat162*>(p.out + (size_t)a * p.out_stride + n0);
        float2 old = __bfloat1622float2(*px);
        *px = __floats2bfloat162_rn(old.x + bf16_round(a0[r]), old.y + bf16_round(a1[r]));
      }
    }
  } else if constexpr (EPI == EPI_SWIGLU) {
    // h = bf16( bf16(y) * bf16(silu(bf16(gate))) )   (_torch.py:150-152; row n0 = value, row n0+F = gate)
#pragma unroll
    for (int r = 0; r < R; ++r) {
      bool ok;
      const int a = gemv_row<R>(p, r, ok);
      if (lane == r && ok) {
        const float y = bf16_round(a0[r]);
        const float g = bf16_round(a1[r]);
        const float s = bf16_round(g / (1.0f + expf(-g)));
        p.out[(size_t)a * p.out_stride + n0] = __float2bfloat16_rn(y * s);
      }
    }
  } else if constexpr (EPI == EPI_QKV) {
    // fused-QKV rows: [0,q) -> RoPE -> q buffer; [q, q+kv) -> RoPE -> K page; rest -> V page.
    // RoPE: adjacent pair rotation in fp32 without FMA contraction (_torch.py:18-30).
    const int pos = p.st->pos;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      bool ok;
      const int a = gemv_row<R>(p, r, ok);
      if (lane == r && ok) {
        const float v0 = bf16_round(a0[r]);
        const float v1 = bf16_round(a1[r]);
        if (n0 < p.q_size + p.kv_size) {
          const int i = (n0 % ZB_DH) >> 1;
          const float2 cs = *reinterpret_cast<const float2*>(p.rope + ((size_t)p.rope_pos[a] * (ZB_DH / 2) + i) * 2);
          const float o0 = __fsub_rn(__fmul_rn(v0, cs.x), __fmul_rn(v1, cs.y));
          const float o1 = __fadd_rn(__fmul_rn(v1, cs.x), __fmul_rn(v0, cs.y));
          const __nv_bfloat162 o = __floats2bfloat162_rn(o0, o1);
          if (n0 < p.q_size) {
            *reinterpret_cast<__nv_bfloat162*>(p.out + (size_t)a * p.out_stride + n0) = o;
          } else {
            const int m = n0 - p.q_size;
            const int page = p.page_table[a * p.max_pages + pos / ZB_PAGE];
            bf16* dst = p.kv_layer + ((((size_t)page * 2 + 0) * p.Hkv + m / ZB_DH) * ZB_PAGE + pos % ZB_PAGE) * ZB_DH + m % ZB_DH;
            *reinterpret_cast<__nv_bfloat162*>(dst) = o;
          }
        } else {
          const int m = n0 - p.q_size - p.kv_size;
          const int page = p.page_table[a * p.max_pages + pos / ZB_PAGE];
          bf16* dst = p.kv_layer + ((((size_t)page * 2 + 1) * p.Hkv + m / ZB_DH) * ZB_PAGE + pos % ZB_PAGE) * ZB_DH + m % ZB_DH;
          *reinterpret_cast<__nv_bfloat162*>(dst) = __floats2bfloat162_rn(v0, v1);
        }
      }
    }
  } else {  // EPI_HEADS: rows of the 9 stacked heads; CFG mix + pad mask + eos bias (model.py:111-115,266-267,280)
    const int k = n0 / p.V, n = n0 % p.V;
    const float scale = p.st->cfg_scale;
    const int eos_bias = p.st->apply_eos_bias;
#pragma unroll
    for (int b = 0; b < R / 2; ++b) {
      if (lane == b && p.base0 + b < p.lim0) {
        float out[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const float c = bf16_round(j == 0 ? a0[b] : a1[b]);
          const float u = bf16_round(j == 0 ? a0[b + R / 2] : a1[b + R / 2]);
          float l = __fadd_rn(u, __fmul_rn(__fsub_rn(c, u), scale));
          const int nn = n + j;
          if (nn >= 1025) l = -INFINITY;
          if (eos_bias && k >= 1 && nn == p.eos) l = -INFINITY;
          out[j] = l;
        }
        *reinterpret_cast<float2*>(p.logits + ((size_t)(p.base0 + b) * p.n_q + k) * p.V + n) = make_float2(out[0], out[1]);
      }
    }
  }
}

// One launch = one weight matrix.  grid = 2*#SMs, block = 256, dyn smem = R*K*2 + 8*R*4 bytes.
template <int R, int EPI>
__global__ void __launch_bounds__(GEMV_THREADS, GEMV_CTAS_PER_SM) k_gemv(const GemvParams p) {
  extern __shared__ __align__(16) unsigned char zb_smem[];
  bf16* xs = reinterpret_cast<bf16*>(zb_smem);
  float* red = reinterpret_cast<float*>(zb_smem + (size_t)R * p.K * sizeof(bf16));

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  const int K = p.K;
  const int nchunk = K >> 8;                               // uint4 chunks per lane per row
  const int upt = (nchunk + GEMV_U - 1) / GEMV_U;          // pipeline units per task
  const int first = blockIdx.x + gridDim.x * warp;         // task ids: first + i*stride
  const int stride = gridDim.x * nwarp;
  const int my_tasks = first < p.n_tasks ? (p.n_tasks - first + stride - 1) / stride : 0;
  const int nunits = my_tasks * upt;

  pdl_launch_dependents();

  uint4 wa0[GEMV_U], wa1[GEMV_U], wb0[GEMV_U], wb1[GEMV_U];
  auto load_unit = [&](uint4(&d0)[GEMV_U], uint4(&d1)[GEMV_U], int u) {
    const int task = first + (u / upt) * stride;
    const int g = u % upt;
    const size_t row0 = (size_t)task * p.row0_mul;
    const bf16* r0 = p.W + row0 * K;
    const bf16* r1 = p.W + (row0 + p.row1_off) * K;
#pragma unroll
    for (int j = 0; j < GEMV_U; ++j) {
      const int c = g * GEMV_U + j;
      if (c < nchunk) {
        const int k = (c * 32 + lane) * 8;
        d0[j] = ldg_stream(r0 + k);
        d1[j] = ldg_stream(r1 + k);
      }
    }
  };

  // Weights do not depend on the previous kernel: get the first unit in flight before waiting on it.
  if (nunits > 0) load_unit(wa0, wa1, 0);
  if (p.pf_bytes && tid < 2) {
    const char* base = tid == 0 ? p.pf0 : p.pf1;
    if (base) {
      const unsigned per = p.pf_bytes / gridDim.x;
      const char* q = base + (size_t)blockIdx.x * per;
      for (unsigned off = 0; off < per; off += 32768u) l2_prefetch_bulk(q + off, min(32768u, per - off));
    }
  }

  pdl_wait_prior_grid();
  if (p.st->done) return;

  gemv_stage_activations<R>(p, xs, red, K);
  __syncthreads();

  float acc0[R], acc1[R];
#pragma unroll
  for (int r = 0; r < R; ++r) { acc0[r] = 0.f; acc1[r] = 0.f; }

  auto compute_unit = [&](const uint4(&d0)[GEMV_U], const uint4(&d1)[GEMV_U], int u) {
    const int g = u % upt;
#pragma unroll
    for (int j = 0; j < GEMV_U; ++j) {
      const int c = g * GEMV_U + j;
      if (c < nchunk) {
        const int k = (c * 32 + lane) * 8;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const uint4 xv = *reinterpret_cast<const uint4*>(xs + (size_t)r * K + k);
          acc0[r] = dot8(d0[j], xv, acc0[r]);
          acc1[r] = dot8(d1[j], xv, acc1[r]);
        }
      }
    }
    if (g == upt - 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) { acc0[r] = warp_sum(acc0[r]); acc1[r] = warp_sum(acc1[r]); }
      gemv_epilogue<R, EPI>(p, first + (u / upt) * stride, lane, acc0, acc1);
#pragma unroll
      for (int r = 0; r < R; ++r) { acc0[r] = 0.f; acc1[r] = 0.f; }
    }
  };

  for (int u = 0; u < nunits; u += 2) {
    if (u + 1 < nunits) load_unit(wb0, wb1, u + 1);
    compute_unit(wa0, wa1, u);
    if (u + 1 < nunits) {
      if (u + 2 < nunits) load_unit(wa0, wa1, u + 2);
      compute_unit(wb0, wb1, u + 1);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Embedding gather-sum (model.py:97-98): x[b] = ((E0[c0] + E1[c1]) + ...) sequential bf16 adds; rows b and B+b
// get the same vector (model.py:141-142).  grid = B, block = d/8 threads (one uint4 each), capped 1024.
// ---------------------------------------------------------------------------------------------
__global__ void k_embed(const bf16* __restrict__ emb /*[n_q][vocab_rows][d]*/, const long long* __restrict__ delayed,
                        bf16* __restrict__ x, int d, int n_q, int emb_rows, const ZbDevState* st) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (st->done) return;
  const int b = blockIdx.x, B = st->B, L = st->L, col = st->col;
  for (int e = threadIdx.x * 8; e < d; e += blockDim.x * 8) {
    float acc[8];
    for (int k = 0; k < n_q; ++k) {
      long long id = delayed[((size_t)b * n_q + k) * L + col];
      id = id < 0 ? 0 : (id >= emb_rows ? emb_rows - 1 : id);
      ZB_ASSERT(id >= 0 && id < emb_rows);
      const uint4 raw = *reinterpret_cast<const uint4*>(emb + ((size_t)k * emb_rows + id) * d + e);
      const float v[8] = {bf16lo(raw.x), bf16hi(raw.x), bf16lo(raw.y), bf16hi(raw.y), bf16lo(raw.z), bf16hi(raw.z), bf16lo(raw.w), bf16hi(raw.w)};
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = k == 0 ? v[j] : bf16_round(acc[j] + v[j]);
    }
    __nv_bfloat162 o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = __floats2bfloat162_rn(acc[2 * j], acc[2 * j + 1]);
    *reinterpret_cast<uint4*>(x + (size_t)b * d + e) = *reinterpret_cast<uint4*>(o);
    *reinterpret_cast<uint4*>(x + (size_t)(b + B) * d + e) = *reinterpret_cast<uint4*>(o);
  }
}

// Embedding of P known delayed-code columns at once (the audio prefix of model.py:246-255): out[b][j] = out[B + b][j] =
// embed_codes(delayed[b, :, j]), out is [2B][P][d] like a prefix_conditioning block.  grid = (P, B).
__global__ void k_embed_cols(const bf16* __restrict__ emb, const long long* __restrict__ delayed, bf16* __restrict__ out, int d, int n_q,
                             int emb_rows, int B, int L, int P) {
  const int j = blockIdx.x, b = blockIdx.y;
  for (int e = threadIdx.x * 8; e < d; e += blockDim.x * 8) {
    float acc[8];
    for (int k = 0; k < n_q; ++k) {
      long long id = delayed[((size_t)b * n_q + k) * L + j];
      id = id < 0 ? 0 : (id >= emb_rows ? emb_rows - 1 : id);
      ZB_ASSERT(id >= 0 && id < emb_rows);
      const uint4 raw = *reinterpret_cast<const uint4*>(emb + ((size_t)k * emb_rows + id) * d + e);
      const float v[8] = {bf16lo(raw.x), bf16hi(raw.x), bf16lo(raw.y), bf16hi(raw.y), bf16lo(raw.z), bf16hi(raw.z), bf16lo(raw.w), bf16hi(raw.w)};
#pragma unroll
      for (int q = 0; q < 8; ++q) acc[q] = k == 0 ? v[q] : bf16_round(acc[q] + v[q]);
    }
    __nv_bfloat162 o[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) o[q] = __floats2bfloat162_rn(acc[2 * q], acc[2 * q + 1]);
    *reinterpret_cast<uint4*>(out + ((size_t)b * P + j) * d + e) = *reinterpret_cast<uint4*>(o);
    *reinterpret_cast<uint4*>(out + ((size_t)(b + B) * P + j) * d + e) = *reinterpret_cast<uint4*>(o);
  }
}

// x[r] <- src[r, s, :]  (prefill / plugin forward: feed given hidden states)
__global__ void k_load_hidden(const bf16* __restrict__ src, bf16* __restrict__ x, int d, int S, int s) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  const int r = blockIdx.x;
  for (int e = threadIdx.x * 8; e < d; e += blockDim.x * 8)
    *reinterpret_cast<uint4*>(x + (size_t)r * d + e) = *reinterpret_cast<const uint4*>(src + ((size_t)r * S + s) * d + e);
}

// plain LayerNorm of the residual stream into dst[r, s, :] (norm_f for the plugin-level forward)
__global__ void k_final_ln(const bf16* __restrict__ x, const bf16* __restrict__ w, const bf16* __restrict__ b, float eps,
                           bf16* __restrict__ dst, int d, int S, int s) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  __shared__ float red[32];
  const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  float sum = 0.f;
  for (int e = tid; e < d; e += blockDim.x) sum += __bfloat162float(x[(size_t)r * d + e]);
  sum = warp_sum(sum);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  float mean = 0.f;
  for (int w2 = 0; w2 < nwarp; ++w2) mean += red[w2];
  mean /= (float)d;
  __syncthreads();
  float var = 0.f;
  for (int e = tid; e < d; e += blockDim.x) { float t = __bfloat162float(x[(size_t)r * d + e]) - mean; var += t * t; }
  var = warp_sum(var);
  if (lane == 0) red[warp] = var;
  __syncthreads();
  var = 0.f;
  for (int w2 = 0; w2 < nwarp; ++w2) var += red[w2];
  const float rstd = rsqrtf(var / (float)d + eps);
  for (int e = tid; e < d; e += blockDim.x) {
    const float v = (__bfloat162float(x[(size_t)r * d + e]) - mean) * rstd * __bfloat162float(w[e]) + __bfloat162float(b[e]);
    dst[((size_t)r * S + s) * d + e] = __float2bfloat16_rn(v);
  }
}

// host-driven position bookkeeping for prefill / plugin forward
__global__ void k_set_pos(ZbDevState* st, int* rope_pos, int R, int pos, const int* lengths, int add) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (threadIdx.x == 0) { st->pos = pos; st->done = 0; }    // a plugin-level forward after a finished generate must not be skipped
  for (int r = threadIdx.x; r < R; r += blockDim.x) rope_pos[r] = lengths ? lengths[r] + add : pos;
}
__global__ void k_advance(ZbDevState* st, int* rope_pos, int R) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (st->done) return;
  if (threadIdx.x == 0) st->pos += 1;
  for (int r = threadIdx.x; r < R; r += blockDim.x) rope_pos[r] += 1;
}

// server.py:142-147: np.clip(wav, -1, 1) * 32767 -> astype(int16) (C cast: truncation toward zero)
__global__ void k_wav_to_pcm16(const float* __restrict__ wav, short* __restrict__ pcm, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float v = wav[i];
    v = v != v ? 0.f : fminf(fmaxf(v, -1.0f), 1.0f);
    pcm[i] = (short)(int)(__fmul_rn(v, 32767.0f));
  }
}
