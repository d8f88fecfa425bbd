// PrefixConditioner on the device (zonos/conditioning.py:11-52,230-236,257-261,271-273,280-282,297-313).
// The conditioning prefix is a few dozen rows of d_model: everything here is latency-sized.  Arithmetic mirrors the
// reference's bf16 PyTorch ops: bf16 storage, fp32 accumulation inside an op, one rounding to bf16 per op.
#pragma once
#include "zb_common.cuh"

// FourierConditioner.apply_cond: x <- (x - min) / (max - min) (fp32); f = (2 pi x).to(bf16) @ W^T -> bf16; out = [cos f | sin f] bf16.
// grid = n rows, block = 128.  W bf16 [half][in_dim].
__global__ void k_pc_fourier(const float* __restrict__ x, const bf16* __restrict__ w, float min_val, float max_val, bf16* __restrict__ out,
                             int in_dim, int half) {
  const int r = blockIdx.x;
  for (int j = threadIdx.x; j < half; j += blockDim.x) {
    float acc = 0.f;
    for (int i = 0; i < in_dim; ++i) {
      const float xn = (x[(size_t)r * in_dim + i] - min_val) / (max_val - min_val);
      const float xs = bf16_round(bf16_round(xn) * 6.283185307179586f);     // `2 * torch.pi * x.to(bf16)`: a bf16 tensor
      acc = fmaf(xs, __bfloat162float(w[(size_t)j * in_dim + i]), acc);
    }
    const float f = bf16_round(acc);
    out[(size_t)r * 2 * half + j] = __float2bfloat16_rn(cosf(f));
    out[(size_t)r * 2 * half + half + j] = __float2bfloat16_rn(sinf(f));
  }
}

// nn.Embedding gather: out[r] = table[ids[r] - id_offset].  grid = n, block = 128.
__global__ void k_pc_embed(const long long* __restrict__ ids, const bf16* __restrict__ table, bf16* __restrict__ out, int d, long long id_offset,
                           long long n_rows_table) {
  const int r = blockIdx.x;
  long long id = ids[r] - id_offset;
  id = id < 0 ? 0 : (id >= n_rows_table ? n_rows_table - 1 : id);
  for (int e = threadIdx.x * 8; e < d; e += blockDim.x * 8)
    *reinterpret_cast<uint4*>(out + (size_t)r * d + e) = *reinterpret_cast<const uint4*>(table + (size_t)id * d + e);
}

// nn.Linear (+ optional SiLU): out[r][n] = bf16(sum_k in[r][k] W[n][k] + b[n]).  One warp per output column n, up to 8 input
// rows per pass staged in registers-by-loop; grid = ceil(N / 8), block = 256.  K % 8 == 0.
__global__ void __launch_bounds__(256) k_pc_linear(const bf16* __restrict__ in, const bf16* __restrict__ W, const bf16* __restrict__ b,
                                                  bf16* __restrict__ out, int n, int K, int N, int silu) {
  const int lane = threadIdx.x & 31, col = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (col >= N) return;
  const float bias = b ? __bfloat162float(b[col]) : 0.f;
  for (int r0 = 0; r0 < n; r0 += 8) {
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    for (int k = lane * 8; k < K; k += 256) {
      const uint4 wv = *reinterpret_cast<const uint4*>(W + (size_t)col * K + k);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (r0 + i < n) acc[i] = dot8(wv, *reinterpret_cast<const uint4*>(in + (size_t)(r0 + i) * K + k), acc[i]);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float s = warp_sum(acc[i]);
      if (lane == 0 && r0 + i < n) {
        float v = bf16_round(s + bias);
        if (silu) v = bf16_round(v / (1.0f + expf(-v)));
        out[(size_t)(r0 + i) * N + col] = __float2bfloat16_rn(v);
      }
    }
  }
}

// nn.LayerNorm over the last dim (fp32 statistics, bf16 in / out).  grid = n, block = 256.
__global__ void __launch_bounds__(256) k_pc_layernorm(const bf16* __restrict__ in, const bf16* __restrict__ w, const bf16* __restrict__ b, float eps,
                                                     bf16* __restrict__ out, int d) {
  __shared__ float red[16];
  const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float s = 0.f;
  for (int e = tid; e < d; e += 256) s += __bfloat162float(in[(size_t)r * d + e]);
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  float tot = 0.f;
  for (int i = 0; i < 8; ++i) tot += red[i];
  const float mean = tot / (float)d;
  float v = 0.f;
  for (int e = tid; e < d; e += 256) { const float t = __bfloat162float(in[(size_t)r * d + e]) - mean; v += t * t; }
  v = warp_sum(v);
  if (lane == 0) red[8 + warp] = v;
  __syncthreads();
  float vt = 0.f;
  for (int i = 0; i < 8; ++i) vt += red[8 + i];
  const float rstd = rsqrtf(vt / (float)d + eps);
  for (int e = tid; e < d; e += 256)
    out[(size_t)r * d + e] = __float2bfloat16_rn((__bfloat162float(in[(size_t)r * d + e]) - mean) * rstd * __bfloat162float(w[e]) + __bfloat162float(b[e]));
}
