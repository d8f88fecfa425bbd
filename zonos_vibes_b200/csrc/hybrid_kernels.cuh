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
  for (int pp = warp; pp < p.headdim; pp += 8) {
    const float x = xr[h * p.headdim + pp];
    const float dtx = dt * x;
    float y = 0.f;
    for (int n = lane * 4; n < p.d_state; n += 128) {
      float4 s = *reinterpret_cast<float4*>(S + (size_t)pp * p.d_state + n);
      const float4 b = *reinterpret_cast<const float4*>(Bv + n);
      const float4 c = *reinterpret_cast<const float4*>(Cv + n);
      s.x = fmaf(s.x, dA, dtx * b.x); s.y = fmaf(s.y, dA, dtx * b.y); s.z = fmaf(s.z, dA, dtx * b.z); s.w = fmaf(s.w, dA, dtx * b.w);
      *reinterpret_cast<float4*>(S + (size_t)pp * p.d_state + n) = s;
      y += s.x * c.x + s.y * c.y + s.z * c.z + s.w * c.w;
    }
    y = warp_sum(y);
    if (lane == 0) p.y[(size_t)row * p.d_inner + h * p.headdim + pp] = __float2bfloat16_rn(fmaf(Dh, x, y));
  }
}

// grid = rows, block = 256.  RMSNormGated with norm_before_gate = False, one group (ngroups == 1).
__global__ void __launch_bounds__(256) k_m2_gated_norm(const M2Params p) {
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  if (p.st->done) return;
  __shared__ float red[8];
  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float ss = 0.f;
  for (int i = tid; i < p.d_inner; i += 256) {
    const float z = __bfloat162float(p.zx[(size_t)row * p.d_in_proj + i]);
    const float v = __bfloat162float(p.y[(size_t)row * p.d_inner + i]) * m2_silu(z);
    ss += v * v;
  }
  ss = warp_sum(ss);
  if (lane == 0) red[warp] = ss;
  __syncthreads();
  float tot = 0.f;
  for (int w = 0; w < 8; ++w) tot += red[w];
  const float r = rsqrtf(tot / (float)p.d_inner + p.eps);
  for (int i = tid; i < p.d_inner; i += 256) {
    const float z = __bfloat162float(p.zx[(size_t)row * p.d_in_proj + i]);
    const float v = __bfloat162float(p.y[(size_t)row * p.d_inner + i]) * m2_silu(z);
    p.ynorm[(size_t)row * p.d_inner + i] = __float2bfloat16_rn(v * r * __bfloat162float(p.norm_w[i]));
  }
}
