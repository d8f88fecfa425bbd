// Mamba2 decode-step kernels for the hybrid backbone (zonos/backbone/_mamba_ssm.py:8-57; arithmetic of
// mamba_ssm.modules.mamba2.Mamba2.step as restated by transformers/models/mamba2/modeling_mamba2.py:419-502,103-118).
//
// One Mamba2 layer per token =
//   k_gemv_mma<R, EPI_PLAIN>   LayerNorm (fused into the activation staging) + in_proj -> zxbcdt [rows][2*d_inner + 2*G*N + H]
//   k_m2_conv                  rolling depthwise conv state (k = d_conv) + bias + SiLU on xBC
//   k_m2_ssm                   dt = softplus(dt + dt_bias); state = state * exp(dt*A) + dt * B (x) x; y = state . C + D * x
//   k_m2_gated_norm            rmsnorm(y * silu(z)) * w
//   k_gemv_mma<R, EPI_RESID>   out_proj + residual
// All of it is HBM-bound: 51.7 MB of weights per layer (in_proj 34.9 + out_proj 16.8) against 2 MB of fp32 state per
// row.  The states are kept in fp32 (mamba-ssm allocates them in the model dtype; fp32 is the numerically safer
// superset and costs 4 % of the layer's bytes).  The prefill steps the same kernels token by token — the chunked SSD
// scan computes the same recurrence.
#pragma once
#include "zb_common.cuh"

struct M2Params {
  const bf16* zx;            // [rows][d_in_proj] bf16: z | xBC | dt
  int d_in_proj, d_inner, conv_dim, d_conv, nheads, headdim, d_state, ngroups;
  const bf16 *conv_w, *conv_b;     // [conv_dim][d_conv], [conv_dim]
  const float *dt_bias, *A_log, *D;   // [nheads]
  const bf16* norm_w;        // [d_inner]
  float* conv_state;         // [rows][conv_dim][d_conv]
  float* ssm_state;          // [rows][nheads][headdim][d_state]
  float* xbc;                // [rows][conv_dim]  conv + SiLU output (bf16-rounded values)
  bf16* y;                   // [rows][d_inner]
  bf16* ynorm;               // [rows][d_inner]
  float eps;
  const ZbDevState* st;
};

__device__ __forceinline__ float m2_silu(float v) { return v / (1.0f + expf(-v)); }

// grid = (ceil(conv_dim / 256), rows), block = 256
__global__ void __launch_bounds__(256) k_m2_conv(const M2Params p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;
  const int ch = blockIdx.x * 256 + threadIdx.x, row = blockIdx.y;
  if (ch >= p.conv_dim) return;
  float* s = p.conv_state + ((size_t)row * p.conv_dim + ch) * p.d_conv;
  const float xin = __bfloat162float(p.zx[(size_t)row * p.d_in_proj + p.d_inner + ch]);
  float acc = __bfloat162float(p.conv_b[ch]);
  for (int k = 0; k < p.d_conv; ++k) {
    const float v = k + 1 < p.d_conv ? s[k + 1] : xin;        // roll by one, newest sample last
    s[k] = v;
    acc = fmaf(v, __bfloat162float(p.conv_w[(size_t)ch * p.d_conv + k]), acc);
  }
  p.xbc[(size_t)row * p.conv_dim + ch] = bf16_round(m2_silu(acc));
}

// grid = (nheads, rows), block = 256 (8 warps; one warp per state row p, lanes over d_state in float4)
__global__ void __launch_bounds__(256) k_m2_ssm(const M2Params p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;
  const int h = blockIdx.x, row = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = h / (p.nheads / p.ngroups);
  const float A = -expf(p.A_log[h]);
  const float dt_raw = __bfloat162float(p.zx[(size_t)row * p.d_in_proj + p.d_inner + p.conv_dim + h]) + p.dt_bias[h];
  const float dt = dt_raw > 20.f ? dt_raw : log1pf(expf(dt_raw));                 // F.softplus (threshold 20)
  const float dA = expf(dt * A);
  const float* xr = p.xbc + (size_t)row * p.conv_dim;
  const float* Bv = xr + p.d_inner + (size_t)g * p.d_state;
  const float* Cv = Bv + (size_t)p.ngroups * p.d_state;
  const float Dh = p.D[h];
  float* S = p.ssm_state + (((size_t)row * p.nheads + h) * p.headdim) * p.d_state;
  // d_state == 128 (checked at context creation): lane l holds columns [4 l, 4 l + 4) of B, C and of every state row.
  // The state rows of this warp go EIGHT AT A TIME: all eight 512-byte loads of a warp are in flight before the first
  // multiply (one load per iteration left the kernel at 2.9 TB/s: 38 us per layer at 32 rows, the largest item of the
  // batched hybrid step).
  const int n = lane * 4;
  const float4 b = *reinterpret_cast<const float4*>(Bv + n);
  const float4 c = *reinterpret_cast<const float4*>(Cv + n);
  for (int p0 = warp; p0 < p.headdim; p0 += 64) {
    float4 sv[8];
    float xv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int pp = p0 + 8 * i;
      if (pp < p.headdim) {
        sv[i] = *reinterpret_cast<const float4*>(S + (size_t)pp * p.d_state + n);
        xv[i] = xr[h * p.headdim + pp];
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int pp = p0 + 8 * i;
      if (pp < p.headdim) {
        const float dtx = dt * xv[i];
        float4 t = sv[i];
        t.x = fmaf(t.x, dA, dtx * b.x); t.y = fmaf(t.y, dA, dtx * b.y); t.z = fmaf(t.z, dA, dtx * b.z); t.w = fmaf(t.w, dA, dtx * b.w);
        *reinterpret_cast<float4*>(S + (size_t)pp * p.d_state + n) = t;
        const float y = warp_sum(t.x * c.x + t.y * c.y + t.z * c.z + t.w * c.w);
        if (lane == 0) p.y[(size_t)row * p.d_inner + h * p.headdim + pp] = __float2bfloat16_rn(fmaf(Dh, xv[i], y));
      }
    }
  }
}

// grid = rows, block = 256.  RMSNormGated with norm_before_gate = False, one group (ngroups == 1).
__global__ void __launch_bounds__(256) k_m2_gated_norm(const M2Params p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;
  __shared__ float red[8];
  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // 8 values per 16-byte load, up to 4 such groups per thread kept in registers (d_inner <= 8192): y * silu(z) is computed
  // ONCE (the scalar two-pass version took 16 us per layer for 0.5 MB: 32 dependent round trips per thread)
  float v[4][8];
  float ss = 0.f;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int i = (u * 256 + tid) * 8;
    if (i < p.d_inner) {
      const uint4 zr = *reinterpret_cast<const uint4*>(p.zx + (size_t)row * p.d_in_proj + i);
      const uint4 yr = *reinterpret_cast<const uint4*>(p.y + (size_t)row * p.d_inner + i);
      const unsigned zw[4] = {zr.x, zr.y, zr.z, zr.w}, yw[4] = {yr.x, yr.y, yr.z, yr.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[u][2 * j] = bf16lo(yw[j]) * m2_silu(bf16lo(zw[j]));
        v[u][2 * j + 1] = bf16hi(yw[j]) * m2_silu(bf16hi(zw[j]));
        ss = fmaf(v[u][2 * j], v[u][2 * j], ss);
        ss = fmaf(v[u][2 * j + 1], v[u][2 * j + 1], ss);
      }
    }
  }
  ss = warp_sum(ss);
  if (lane == 0) red[warp] = ss;
  __syncthreads();
  float tot = 0.f;
  for (int w = 0; w < 8; ++w) tot += red[w];
  const float r = rsqrtf(tot / (float)p.d_inner + p.eps);
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int i = (u * 256 + tid) * 8;
    if (i < p.d_inner) {
      const uint4 wr = *reinterpret_cast<const uint4*>(p.norm_w + i);
      const unsigned ww[4] = {wr.x, wr.y, wr.z, wr.w};
      __nv_bfloat162 o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) o[j] = __floats2bfloat162_rn(v[u][2 * j] * r * bf16lo(ww[j]), v[u][2 * j + 1] * r * bf16hi(ww[j]));
      *reinterpret_cast<uint4*>(p.ynorm + (size_t)row * p.d_inner + i) = *reinterpret_cast<uint4*>(o);
    }
  }
}

// =====================================================================================================================
// Prefill: the SEQUENCE form of the Mamba2 layer (mamba_ssm Mamba2.forward's chunked SSD scan, restated in
// transformers/models/mamba2/modeling_mamba2.py:503-577).  Activations are token-major: row a = t * rows + r.
// =====================================================================================================================
struct M2SeqParams {
  const bf16* zx;            // [S * rows][d_in_proj] bf16: z | xBC | dt
  int rows, S;
  int d_in_proj, d_inner, conv_dim, d_conv, nheads, headdim, d_state, ngroups;
  const bf16 *conv_w, *conv_b;
  const float *dt_bias, *A_log, *D;
  float* conv_state;         // [rows][conv_dim][d_conv]   (read at t = 0, left as after token S - 1)
  float* ssm_state;          // [rows][nheads][headdim][d_state]
  float* xbc;                // [S * rows][conv_dim]
  bf16* y;                   // [S * rows][d_inner]
};

// Depthwise causal conv over the sequence, rolling state carried in registers.  grid = (ceil(conv_dim / 256), rows).
// Same accumulation order as k_m2_conv, so stepping and the sequence form give bit-identical xBC.
__global__ void __launch_bounds__(256) k_m2_conv_seq(const M2SeqParams p) {
  const int ch = blockIdx.x * 256 + threadIdx.x, row = blockIdx.y;
  if (ch >= p.conv_dim) return;
  float s[8], w[8];
  float* sp = p.conv_state + ((size_t)row * p.conv_dim + ch) * p.d_conv;
  for (int k = 0; k < 8; ++k) { s[k] = k < p.d_conv ? sp[k] : 0.f; w[k] = k < p.d_conv ? __bfloat162float(p.conv_w[(size_t)ch * p.d_conv + k]) : 0.f; }
  const float bias = __bfloat162float(p.conv_b[ch]);
  for (int t = 0; t < p.S; ++t) {
    const size_t a = (size_t)t * p.rows + row;
    const float xin = __bfloat162float(p.zx[a * p.d_in_proj + p.d_inner + ch]);
    float acc = bias;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (k < p.d_conv) {
        const float v = k + 1 < p.d_conv ? s[k + 1] : xin;
        s[k] = v;
        acc = fmaf(v, w[k], acc);
      }
    }
    p.xbc[a * p.conv_dim + ch] = bf16_round(m2_silu(acc));
  }
  for (int k = 0; k < 8; ++k) if (k < p.d_conv) sp[k] = s[k];
}

// Chunked SSD scan.  One CTA per (head, row) walks the sequence in chunks of M2_L tokens; per chunk, with
// a_t = dt_t * A (log decay), cs_t = sum_{u <= t} a_u, X_t = dt_t * x_t:
//   G[t][s]  = (C_t . B_s) * exp(cs_t - cs_s)            for s <= t     (intra-chunk "attention" matrix, L x L, K = d_state)
//   Y[t]     = sum_{s <= t} G[t][s] X_s  +  exp(cs_t) * (C_t . S0)  +  D x_t
//   S_end    = exp(cs_L) * S0 + sum_s exp(cs_L - cs_s) X_s (x) B_s   (chunk state, headdim x d_state, K = L)
// i.e. three small GEMMs per chunk instead of S dependent rank-1 updates, and the state passes from chunk to chunk in
// shared memory (the inter-chunk scan is the chunk loop itself: one CTA owns the whole sequence of its head).  All of
// it in fp32 (the step kernel's precision; the products are far too small - 1.4 MFLOP per chunk - for the tensor
// pipe to matter: a 600-token prefill of 41 layers is 40 GFLOP).  d_state == 128, headdim == 64 (checked by the host).
#define M2_L 64
#define M2_LDN 132          // row stride (floats) of the [.][d_state] tiles: float4-aligned, rows 4 banks apart
#define M2_LDP 68           // row stride of the [.][headdim] tiles
#define M2_SSD_SMEM ((2 * M2_L * M2_LDN + 64 * M2_LDN + 2 * M2_L * M2_LDP + M2_L * (M2_L + 4) + 4 * M2_L) * 4)
__global__ void __launch_bounds__(256) k_m2_ssd_chunked(const M2SeqParams p) {
  extern __shared__ __align__(16) float m2s[];
  float* sB = m2s;                              // [L][LDN]
  float* sC = sB + M2_L * M2_LDN;               // [L][LDN]
  float* sS = sC + M2_L * M2_LDN;               // [64][LDN]   state S[p][n]
  float* sX = sS + 64 * M2_LDN;                 // [L][LDP]    dt_t * x_t[p]
  float* sXr = sX + M2_L * M2_LDP;              // [L][LDP]    x_t[p]
  float* sG = sXr + M2_L * M2_LDP;              // [L][L + 4]
  float* s_cs = sG + M2_L * (M2_L + 4);         // [L] inclusive cumulative log decay
  float* s_dt = s_cs + M2_L;                    // [L]
  const int h = blockIdx.x, row = blockIdx.y, tid = threadIdx.x;
  const int g = h / (p.nheads / p.ngroups);
  const float A = -expf(p.A_log[h]), Dh = p.D[h], dtb = p.dt_bias[h];
  float* Sg = p.ssm_state + (((size_t)row * p.nheads + h) * 64) * 128;
  for (int i = tid; i < 64 * 32; i += 256) {
    const int pp = i >> 5, n4 = (i & 31) * 4;
    *reinterpret_cast<float4*>(sS + pp * M2_LDN + n4) = *reinterpret_cast<const float4*>(Sg + pp * 128 + n4);
  }
  for (int t0 = 0; t0 < p.S; t0 += M2_L) {
    const int Lc = min(M2_L, p.S - t0);
    __syncthreads();                            // previous chunk's readers of sX / sB / sC / s_cs are done
    // ---- (a) operands of this chunk
    if (tid < M2_L) {
      float dt = 0.f;
      if (tid < Lc) {
        const size_t a = (size_t)(t0 + tid) * p.rows + row;
        const float raw = __bfloat162float(p.zx[a * p.d_in_proj + p.d_inner + p.conv_dim + h]) + dtb;
        dt = raw > 20.f ? raw : log1pf(expf(raw));                 // F.softplus (threshold 20)
      }
      s_dt[tid] = dt;
    }
    for (int i = tid; i < M2_L * 32; i += 256) {
      const int t = i >> 5, n4 = (i & 31) * 4;
      float4 b = make_float4(0.f, 0.f, 0.f, 0.f), c = b;
      if (t < Lc) {
        const float* xr = p.xbc + ((size_t)(t0 + t) * p.rows + row) * p.conv_dim;
        b = *reinterpret_cast<const float4*>(xr + p.d_inner + (size_t)g * 128 + n4);
        c = *reinterpret_cast<const float4*>(xr + p.d_inner + (size_t)p.ngroups * 128 + (size_t)g * 128 + n4);
      }
      *reinterpret_cast<float4*>(sB + t * M2_LDN + n4) = b;
      *reinterpret_cast<float4*>(sC + t * M2_LDN + n4) = c;
    }
    __syncthreads();
    if (tid == 0) {                             // inclusive scan of the log decays (64 adds)
      float c = 0.f;
      for (int t = 0; t < M2_L; ++t) { c += s_dt[t] * A; s_cs[t] = c; }
    }
    for (int i = tid; i < M2_L * 16; i += 256) {
      const int t = i >> 4, p4 = (i & 15) * 4;
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (t < Lc) x = *reinterpret_cast<const float4*>(p.xbc + ((size_t)(t0 + t) * p.rows + row) * p.conv_dim + h * 64 + p4);
      const float dt = s_dt[t];
      *reinterpret_cast<float4*>(sXr + t * M2_LDP + p4) = x;
      *reinterpret_cast<float4*>(sX + t * M2_LDP + p4) = make_float4(x.x * dt, x.y * dt, x.z * dt, x.w * dt);
    }
    __syncthreads();
    // ---- (b) G = (C B^T) .* decay, lower triangle; thread = 4 x 4 tile (ti, si)
    {
      const int ti = tid >> 4, si = tid & 15;
      float acc[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
      if (si <= ti) {
        for (int n = 0; n < 128; n += 4) {
          float4 c[4], b[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) { c[i] = *reinterpret_cast<const float4*>(sC + (4 * ti + i) * M2_LDN + n); b[i] = *reinterpret_cast<const float4*>(sB + (4 * si + i) * M2_LDN + n); }
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] += c[i].x * b[j].x + c[i].y * b[j].y + c[i].z * b[j].z + c[i].w * b[j].w;
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int t = 4 * ti + i, s = 4 * si + j;
          sG[t * (M2_L + 4) + s] = s <= t ? acc[i][j] * expf(s_cs[t] - s_cs[s]) : 0.f;
        }
    }
    __syncthreads();
    // ---- (c) Y = G X + exp(cs) .* (C S0^T) + D x; thread = 4 x 4 tile (ti, pi)
    {
      const int ti = tid >> 4, pi = tid & 15;
      float acc[4][4], acs[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) { acc[i][j] = 0.f; acs[i][j] = 0.f; }
      const int smax = min(Lc, 4 * ti + 4);
      for (int s = 0; s < smax; ++s) {
        const float4 x = *reinterpret_cast<const float4*>(sX + s * M2_LDP + 4 * pi);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float gv = sG[(4 * ti + i) * (M2_L + 4) + s];
          acc[i][0] += gv * x.x; acc[i][1] += gv * x.y; acc[i][2] += gv * x.z; acc[i][3] += gv * x.w;
        }
      }
      for (int n = 0; n < 128; n += 4) {
        float4 c[4], sv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { c[i] = *reinterpret_cast<const float4*>(sC + (4 * ti + i) * M2_LDN + n); sv[i] = *reinterpret_cast<const float4*>(sS + (4 * pi + i) * M2_LDN + n); }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acs[i][j] += c[i].x * sv[j].x + c[i].y * sv[j].y + c[i].z * sv[j].z + c[i].w * sv[j].w;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int t = 4 * ti + i;
        if (t < Lc) {
          const float e = expf(s_cs[t]);
          const float4 xr = *reinterpret_cast<const float4*>(sXr + t * M2_LDP + 4 * pi);
          const float y0 = acc[i][0] + e * acs[i][0] + Dh * xr.x, y1 = acc[i][1] + e * acs[i][1] + Dh * xr.y;
          const float y2 = acc[i][2] + e * acs[i][2] + Dh * xr.z, y3 = acc[i][3] + e * acs[i][3] + Dh * xr.w;
          __nv_bfloat162* dst = reinterpret_cast<__nv_bfloat162*>(p.y + ((size_t)(t0 + t) * p.rows + row) * p.d_inner + h * 64 + 4 * pi);
          dst[0] = __floats2bfloat162_rn(y0, y1);
          dst[1] = __floats2bfloat162_rn(y2, y3);
        }
      }
    }
    __syncthreads();                            // every reader of S0 is done
    // ---- (d) S = exp(cs_last) S0 + sum_s exp(cs_last - cs_s) X_s (x) B_s; thread = 4 (p) x 8 (n) tile
    {
      const int pi = tid >> 4, ni = tid & 15;
      const float cl = s_cs[Lc - 1];
      float acc[4][8];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
      for (int s = 0; s < Lc; ++s) {
        const float w = expf(cl - s_cs[s]);
        const float4 x = *reinterpret_cast<const float4*>(sX + s * M2_LDP + 4 * pi);
        const float4 b0 = *reinterpret_cast<const float4*>(sB + s * M2_LDN + 8 * ni), b1 = *reinterpret_cast<const float4*>(sB + s * M2_LDN + 8 * ni + 4);
        const float xv[4] = {x.x * w, x.y * w, x.z * w, x.w * w};
        const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(xv[i], bv[j], acc[i][j]);
      }
      const float e = expf(cl);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float* d = sS + (4 * pi + i) * M2_LDN + 8 * ni + j;
          *d = fmaf(e, *d, acc[i][j]);
        }
    }
  }
  __syncthreads();
  for (int i = tid; i < 64 * 32; i += 256) {
    const int pp = i >> 5, n4 = (i & 31) * 4;
    *reinterpret_cast<float4*>(Sg + pp * 128 + n4) = *reinterpret_cast<const float4*>(sS + pp * M2_LDN + n4);
  }
}

// token-major gather / final LayerNorm scatter for the layer-major prefill: X[t * rows + r] <-> hidden[r][t]
__global__ void k_m2_gather_tm(const bf16* __restrict__ in, bf16* __restrict__ X, int rows, int S, int d) {
  const int a = blockIdx.x, t = a / rows, r = a % rows;
  for (int e = threadIdx.x * 8; e < d; e += blockDim.x * 8)
    *reinterpret_cast<uint4*>(X + (size_t)a * d + e) = *reinterpret_cast<const uint4*>(in + ((size_t)r * S + t) * d + e);
}
__global__ void k_bt_scatter_tm(const bf16* __restrict__ X, bf16* __restrict__ out, int rows, int S, int d) {
  const int a = blockIdx.x, t = a / rows, r = a % rows;
  for (int e = threadIdx.x * 8; e < d; e += blockDim.x * 8)
    *reinterpret_cast<uint4*>(out + ((size_t)r * S + t) * d + e) = *reinterpret_cast<const uint4*>(X + (size_t)a * d + e);
}
