// DAC 44 kHz decoder — fp32 SIMT kernels (impl=1).  These are the high-precision on-device cross-check for the
// tcgen05 path (dac_tc.cuh) and the fallback while a tensor-core layer is being brought up; same arithmetic as
// transformers/models/dac/modeling_dac.py:85-99,173-207,234-262,345-369,405-439.
//
// Layout: activations are TIME-MAJOR [T][C] (channels contiguous) so that a conv tap is a row shift and a warp's
// channel accesses coalesce; weights are repacked [tap][Cin][Cout].
#pragma once
#include <cuda_fp16.h>

#include "zb_common.cuh"

__device__ __forceinline__ float snake_f(float x, float alpha) {
  const float s = sinf(alpha * x);
  return x + (1.0f / (alpha + 1e-9f)) * (s * s);
}

// table[k][code][c] = bias_k[c] + sum_j W_k[c][j] * codebook_k[code][j]      (RVQ out_proj folded into the codebook)
__global__ void k_dac_fold(const float* __restrict__ codebook /*[1024][8]*/, const float* __restrict__ w /*[C][8]*/,
                           const float* __restrict__ bias /*[C]*/, float* __restrict__ table /*[1024][C]*/, int C, int ncode, int cdim) {
  const int code = blockIdx.x;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float acc = 0.f;
    for (int j = 0; j < cdim; ++j) acc = fmaf(w[c * cdim + j], codebook[code * cdim + j], acc);
    table[(size_t)code * C + c] = acc + bias[c];
  }
}

// z[b][t][c] = sum_k table[k][codes[b][k][t]][c]        codes int64 [nb][n_q][T], out fp32 and/or fp16 [nb][T][C]; grid = nb * T
__global__ void k_dac_latent(const long long* __restrict__ codes, const float* __restrict__ table, float* __restrict__ z32,
                             __half* __restrict__ z16, int T, int C, int nq, int ncode) {
  const int t = blockIdx.x % T;
  codes += (size_t)(blockIdx.x / T) * nq * T;
  if (z32) z32 += (size_t)(blockIdx.x / T) * T * C;
  if (z16) z16 += (size_t)(blockIdx.x / T) * T * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float acc = 0.f;
    for (int k = 0; k < nq; ++k) {
      long long code = codes[(size_t)k * T + t];
      code = code < 0 ? 0 : (code >= ncode ? ncode - 1 : code);
      acc += table[((size_t)k * ncode + code) * C + c];
    }
    if (z32) z32[(size_t)t * C + c] = acc;
    if (z16) z16[(size_t)t * C + c] = __float2half_rn(acc);
  }
}

// [Cout][Cin][K] (Conv1d) or [Cin][Cout][K] (ConvTranspose1d, transposed=1) -> fp32 [K][Cin][Cout] and fp16 [K][Cout][Cin]
__global__ void k_dac_repack(const float* __restrict__ w, float* __restrict__ w_kio, __half* __restrict__ w_koi, int Cout, int Cin, int K, int transposed) {
  const size_t n = (size_t)Cout * Cin * K;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int k = (int)(i % K);
    int co, ci;
    if (transposed) { co = (int)((i / K) % Cout); ci = (int)(i / ((size_t)K * Cout)); }
    else { ci = (int)((i / K) % Cin); co = (int)(i / ((size_t)K * Cin)); }
    const float v = w[i];
    w_kio[((size_t)k * Cin + ci) * Cout + co] = v;
    w_koi[((size_t)k * Cout + co) * Cin + ci] = __float2half_rn(v);
  }
}

#define DSC_TT 32      // output time steps per block
#define DSC_TC 64      // output channels per block
#define DSC_CI 16      // input-channel chunk
#define DSC_MAXROWS 96 // DSC_TT + 6*9 halo = 86 rows

struct ConvSimtParams {
  const float* x;        // [T_in][Cin]
  const float* w;        // [K][Cin][Cout]
  const float* bias;     // [Cout]
  const float* alpha;    // [Cin] snake on the input, or NULL
  const float* residual; // [T_out][Cout] or NULL
  float* out;            // [T_out][Cout]
  int T_in, T_out, Cin, Cout, K, dil, pad, stride;   // stride > 1: transposed convolution with K == 2*stride
  int tanh_out;
};

// Conv1d, "same" zero padding, dilation; grid = (ceil(T/32), ceil(Cout/64)), block = 256.
__global__ void __launch_bounds__(256) k_dac_conv_simt(const ConvSimtParams p) {
  __shared__ __align__(16) float xs[DSC_MAXROWS][DSC_CI];
  __shared__ float ws[7][DSC_CI][DSC_TC];
  const int tid = threadIdx.x, co_l = tid & 63, tq = tid >> 6;
  const int t0 = blockIdx.x * DSC_TT, co0 = blockIdx.y * DSC_TC;
  const int co = co0 + co_l;
  const int rows = DSC_TT + (p.K - 1) * p.dil;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int c0 = 0; c0 < p.Cin; c0 += DSC_CI) {
    __syncthreads();
    for (int i = tid; i < rows * DSC_CI; i += 256) {
      const int r = i / DSC_CI, ci = i % DSC_CI;
      const int t = t0 - p.pad + r;
      float v = 0.f;
      if (t >= 0 && t < p.T_in && c0 + ci < p.Cin) {
        v = p.x[(size_t)t * p.Cin + c0 + ci];
        if (p.alpha) v = snake_f(v, p.alpha[c0 + ci]);
      }
      xs[r][ci] = v;
    }
    for (int i = tid; i < p.K * DSC_CI * DSC_TC; i += 256) {
      const int c = i % DSC_TC, ci = (i / DSC_TC) % DSC_CI, j = i / (DSC_TC * DSC_CI);
      ws[j][ci][c] = (c0 + ci < p.Cin && co0 + c < p.Cout) ? p.w[((size_t)j * p.Cin + c0 + ci) * p.Cout + co0 + c] : 0.f;
    }
    __syncthreads();
    for (int j = 0; j < p.K; ++j) {
#pragma unroll
      for (int c4 = 0; c4 < DSC_CI / 4; ++c4) {
        const float w0 = ws[j][c4 * 4 + 0][co_l], w1 = ws[j][c4 * 4 + 1][co_l], w2 = ws[j][c4 * 4 + 2][co_l], w3 = ws[j][c4 * 4 + 3][co_l];
#pragma unroll
        for (int tt = 0; tt < 8; ++tt) {
          const float4 xv = *reinterpret_cast<const float4*>(&xs[tq * 8 + tt + j * p.dil][c4 * 4]);
          acc[tt] = fmaf(xv.x, w0, acc[tt]); acc[tt] = fmaf(xv.y, w1, acc[tt]);
          acc[tt] = fmaf(xv.z, w2, acc[tt]); acc[tt] = fmaf(xv.w, w3, acc[tt]);
        }
      }
    }
  }
  if (co < p.Cout) {
    const float b = p.bias[co];
#pragma unroll
    for (int tt = 0; tt < 8; ++tt) {
      const int t = t0 + tq * 8 + tt;
      if (t < p.T_out) {
        float v = acc[tt] + b;
        if (p.residual) v += p.residual[(size_t)t * p.Cout + co];
        if (p.tanh_out) v = tanhf(v);
        p.out[(size_t)t * p.Cout + co] = v;
      }
    }
  }
}

// ConvTranspose1d with K = 2*stride, padding = ceil(stride/2) (modeling_dac.py:241-247): T_out = T_in*stride.
// out[t][co] = b[co] + sum_ci x[q][ci] w[phi][ci][co] + x[q-1][ci] w[phi+s][ci][co],  q = (t+pad)/s, phi = (t+pad)%s.
// grid = (ceil(T_out/32), ceil(Cout/64)), block = 256; snake on the input fused.
#define DSC_CIT 8
__global__ void __launch_bounds__(256) k_dac_convt_simt(const ConvSimtParams p) {
  __shared__ float xs[DSC_TT + 2][DSC_CIT];
  __shared__ float ws[16][DSC_CIT][DSC_TC];
  const int tid = threadIdx.x, co_l = tid & 63, tq = tid >> 6;
  const int t0 = blockIdx.x * DSC_TT, co0 = blockIdx.y * DSC_TC;
  const int co = co0 + co_l;
  const int s = p.stride;
  const int q0 = (t0 + p.pad) / s - 1;                       // first input row needed
  const int nrows = (t0 + DSC_TT - 1 + p.pad) / s - q0 + 1;  // <= DSC_TT/s + 2
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int c0 = 0; c0 < p.Cin; c0 += DSC_CIT) {
    __syncthreads();
    for (int i = tid; i < nrows * DSC_CIT; i += 256) {
      const int r = i / DSC_CIT, ci = i % DSC_CIT;
      const int t = q0 + r;
      float v = 0.f;
      if (t >= 0 && t < p.T_in && c0 + ci < p.Cin) {
        v = p.x[(size_t)t * p.Cin + c0 + ci];
        if (p.alpha) v = snake_f(v, p.alpha[c0 + ci]);
      }
      xs[r][ci] = v;
    }
    for (int i = tid; i < p.K * DSC_CIT * DSC_TC; i += 256) {
      const int c = i % DSC_TC, ci = (i / DSC_TC) % DSC_CIT, j = i / (DSC_TC * DSC_CIT);
      ws[j][ci][c] = (c0 + ci < p.Cin && co0 + c < p.Cout) ? p.w[((size_t)j * p.Cin + c0 + ci) * p.Cout + co0 + c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int tt = 0; tt < 8; ++tt) {
      const int t = t0 + tq * 8 + tt;
      const int q = (t + p.pad) / s, phi = (t + p.pad) % s;
      const int r = q - q0;
#pragma unroll
      for (int ci = 0; ci < DSC_CIT; ++ci) {
        acc[tt] = fmaf(xs[r][ci], ws[phi][ci][co_l], acc[tt]);
        acc[tt] = fmaf(xs[r - 1][ci], ws[phi + s][ci][co_l], acc[tt]);
      }
    }
  }
  if (co < p.Cout) {
    const float b = p.bias[co];
#pragma unroll
    for (int tt = 0; tt < 8; ++tt) {
      const int t = t0 + tq * 8 + tt;
      if (t < p.T_out) p.out[(size_t)t * p.Cout + co] = acc[tt] + b;
    }
  }
}

// [T][C] time-major -> [C][T] (reference layout), fp32; used for the waveform (C=1 is a plain copy) and debug dumps
__global__ void k_dac_to_channel_major(const float* __restrict__ x, float* __restrict__ out, int T, int C) {
  const size_t n = (size_t)T * C;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const size_t t = i / C;
    out[(size_t)c * T + t] = x[i];
  }
}

// =================================================================================================================
// DAC encoder side (modeling_dac.py DacEncoder / DacEncoderBlock / DacVectorQuantize /
// DacResidualVectorQuantizer, driven by autoencoder.py:22-23).  fp32 like the reference's encode (no autocast there).
// The residual units reuse k_dac_conv_simt; these are the two ops the decoder does not have.
// =================================================================================================================

// Strided Conv1d of an encoder block: K = 2 * stride, padding = ceil(stride / 2), T_out = T_in / stride; snake on the input.
// out[t][co] = b[co] + sum_j sum_ci snake(x[t * s - pad + j][ci]) w[j][ci][co].   grid = (ceil(T_out / 8), ceil(Cout / 128)),
// block = 128 (one output channel per thread, 8 output steps per block; the input window of the block is staged in smem).
__global__ void __launch_bounds__(128) k_dac_conv_down_simt(const ConvSimtParams p) {
  extern __shared__ float dn_xs[];                       // [rows][Cin], rows = 7 * s + K
  const int s = p.stride, t0 = blockIdx.x * 8, co = blockIdx.y * 128 + threadIdx.x;
  const int rows = 7 * s + p.K;
  for (int i = threadIdx.x; i < rows * p.Cin; i += 128) {
    const int r = i / p.Cin, ci = i - r * p.Cin;
    const int t = t0 * s - p.pad + r;
    float v = 0.f;
    if (t >= 0 && t < p.T_in) {
      v = p.x[(size_t)t * p.Cin + ci];
      if (p.alpha) v = snake_f(v, p.alpha[ci]);
    }
    dn_xs[i] = v;
  }
  __syncthreads();
  if (co >= p.Cout) return;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int j = 0; j < p.K; ++j) {
    const float* wj = p.w + (size_t)j * p.Cin * p.Cout + co;
    for (int ci = 0; ci < p.Cin; ++ci) {
      const float w = wj[(size_t)ci * p.Cout];
#pragma unroll
      for (int tt = 0; tt < 8; ++tt) acc[tt] = fmaf(dn_xs[(tt * s + j) * p.Cin + ci], w, acc[tt]);
    }
  }
  const float b = p.bias[co];
#pragma unroll
  for (int tt = 0; tt < 8; ++tt)
    if (t0 + tt < p.T_out) p.out[(size_t)(t0 + tt) * p.Cout + co] = acc[tt] + b;
}

// Residual vector quantisation of one latent frame (DacResidualVectorQuantizer.forward in eval mode + DacVectorQuantize.
// decode_latents): for k = 0..n_q-1: e = in_proj_k(residual) [8]; both e and the codebook rows L2-normalised (F.normalize,
// eps 1e-12); dist = -(|e|^2 - 2 e.c) + |c|^2; code = argmax (first maximum); residual -= out_proj_k(codebook_k[code]).
// z: [T][C] time-major latents (C = 1024), codes: int64 [n_q][T].  grid = T, block = 256.
__global__ void __launch_bounds__(256) k_dac_rvq_encode(const float* __restrict__ z, const float* const* __restrict__ in_w /*[n_q] -> [8][C]*/,
                                                        const float* const* __restrict__ in_b, const float* const* __restrict__ codebook /*[1024][8]*/,
                                                        const float* const* __restrict__ out_w /*[C][8]*/, const float* const* __restrict__ out_b,
                                                        long long* __restrict__ codes, int T, int C, int n_q, int n_codes) {
  __shared__ float res[1024];
  __shared__ float e[8], en[8];
  __shared__ float red_v[8];
  __shared__ int red_i[8];
  __shared__ int best;
  const int t = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int c = tid; c < C; c += 256) res[c] = z[(size_t)t * C + c];
  __syncthreads();
  for (int k = 0; k < n_q; ++k) {
    // in_proj: warp w computes e[w] (8 warps = 8 code dims)
    {
      const float* w = in_w[k] + (size_t)warp * C;
      float a = 0.f;
      for (int c = lane; c < C; c += 32) a = fmaf(w[c], res[c], a);
      for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
      if (lane == 0) e[warp] = a + in_b[k][warp];
    }
    __syncthreads();
    if (tid == 0) {
      float n2 = 0.f;
      for (int j = 0; j < 8; ++j) n2 += e[j] * e[j];
      const float inv = 1.0f / fmaxf(sqrtf(n2), 1e-12f);
      for (int j = 0; j < 8; ++j) en[j] = e[j] * inv;
    }
    __syncthreads();
    float l2 = 0.f;
    for (int j = 0; j < 8; ++j) l2 += en[j] * en[j];
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int n = tid; n < n_codes; n += 256) {
      const float* cb = codebook[k] + (size_t)n * 8;
      float c2 = 0.f;
      for (int j = 0; j < 8; ++j) c2 += cb[j] * cb[j];
      const float inv = 1.0f / fmaxf(sqrtf(c2), 1e-12f);
      float dot = 0.f, cn2 = 0.f;
      for (int j = 0; j < 8; ++j) { const float cj = cb[j] * inv; dot = fmaf(en[j], cj, dot); cn2 += cj * cj; }
      const float dist = -(l2 - 2.0f * dot) + cn2;
      if (dist > bv) { bv = dist; bi = n; }               // n ascending per thread: the first maximum is kept
    }
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) { red_v[warp] = bv; red_i[warp] = bi; }
    __syncthreads();
    if (tid == 0) {
      float v = red_v[0];
      int i = red_i[0];
      for (int w = 1; w < 8; ++w)
        if (red_v[w] > v || (red_v[w] == v && red_i[w] < i)) { v = red_v[w]; i = red_i[w]; }
      best = i;
      codes[(size_t)k * T + t] = i;
    }
    __syncthreads();
    {
      const float* cb = codebook[k] + (size_t)best * 8;    // the UN-normalised embedding goes through out_proj
      for (int c = tid; c < C; c += 256) {
        const float* w = out_w[k] + (size_t)c * 8;
        float a = out_b[k][c];
        for (int j = 0; j < 8; ++j) a = fmaf(w[j], cb[j], a);
        res[c] -= a;
      }
    }
    __syncthreads();
  }
}
