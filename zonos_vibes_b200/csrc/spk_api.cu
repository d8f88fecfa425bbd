// C ABI of libzonos_b200.so — speaker-embedding network ops (zonos/speaker_cloning.py:38-224,356-412:
// ResNet293-SimAM front, attentive statistics pooling, bottleneck + LDA).  Runs once per speaker and is cached by the
// caller (server.py:38-57), so these are plain fp32 SIMT kernels: the reference evaluates this network in fp32 too.
// Op-level entry points on caller-owned device tensors (NCHW, contiguous); zonos_vibes_b200/speaker_cloning.py composes
// them in the reference's order.  BatchNorm is eval-mode and arrives folded: scale = w / sqrt(var + eps), shift = b - mean * scale.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/zonos_b200.h"

namespace {

#define SPK_TH 4          // output rows per block
#define SPK_TW 32         // output columns per block
#define SPK_CO 64         // output channels per block
#define SPK_CI 8          // input channels per shared-memory chunk

// 3x3 convolution, padding 1, stride 1 or 2, no bias, then y = conv * scale[co] + shift[co], optional ReLU.
// grid = (ceil(Wo / 32), ceil(Ho / 4), B * ceil(Cout / 64)), block = 256: thread = 4 output channels x 8 output columns.
__global__ void __launch_bounds__(256) k_spk_conv3x3(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ scale,
                                                     const float* __restrict__ shift, float* __restrict__ out, int Cin, int H, int W, int Cout,
                                                     int Ho, int Wo, int stride, int relu) {
  __shared__ float xs[SPK_CI][SPK_TH * 2 + 1][SPK_TW * 2 + 2];      // sized for stride 2 (stride 1 uses a corner of it)
  __shared__ __align__(16) float ws[SPK_CI][9][SPK_CO];
  const int n_cog = (Cout + SPK_CO - 1) / SPK_CO;
  const int b = blockIdx.z / n_cog, co0 = (blockIdx.z % n_cog) * SPK_CO;
  const int oy0 = blockIdx.y * SPK_TH, ox0 = blockIdx.x * SPK_TW;
  const int tid = threadIdx.x, cg = tid & 15, pg = tid >> 4;         // 16 channel groups x 16 pixel groups
  const int py = pg >> 2, px0 = (pg & 3) * 8;
  const int rows = (SPK_TH - 1) * stride + 3, cols = (SPK_TW - 1) * stride + 3;
  const int iy0 = oy0 * stride - 1, ix0 = ox0 * stride - 1;
  float acc[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  for (int c0 = 0; c0 < Cin; c0 += SPK_CI) {
    __syncthreads();
    for (int i = tid; i < SPK_CI * rows * cols; i += 256) {
      const int ci = i / (rows * cols), r = (i / cols) % rows, c = i % cols;
      const int iy = iy0 + r, ix = ix0 + c;
      float v = 0.f;
      if (c0 + ci < Cin && iy >= 0 && iy < H && ix >= 0 && ix < W) v = x[(((size_t)b * Cin + c0 + ci) * H + iy) * W + ix];
      xs[ci][r][c] = v;
    }
    for (int i = tid; i < SPK_CI * 9 * SPK_CO; i += 256) {
      const int co = i % SPK_CO, k = (i / SPK_CO) % 9, ci = i / (SPK_CO * 9);
      ws[ci][k][co] = (c0 + ci < Cin && co0 + co < Cout) ? w[((size_t)(co0 + co) * Cin + c0 + ci) * 9 + k] : 0.f;
    }
    __syncthreads();
#pragma unroll 1
    for (int ci = 0; ci < SPK_CI; ++ci) {
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const float4 wv = *reinterpret_cast<const float4*>(&ws[ci][ky * 3 + kx][cg * 4]);
          const float* xr = &xs[ci][py * stride + ky][px0 * stride + kx];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float xv = xr[j * stride];
            acc[0][j] = fmaf(wv.x, xv, acc[0][j]); acc[1][j] = fmaf(wv.y, xv, acc[1][j]);
            acc[2][j] = fmaf(wv.z, xv, acc[2][j]); acc[3][j] = fmaf(wv.w, xv, acc[3][j]);
          }
        }
      }
    }
  }
  const int oy = oy0 + py;
  if (oy >= Ho) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int co = co0 + cg * 4 + i;
    if (co >= Cout) continue;
    const float sc = scale[co], sh = shift[co];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int ox = ox0 + px0 + j;
      if (ox < Wo) {
        float v = acc[i][j] * sc + sh;
        if (relu) v = fmaxf(v, 0.f);
        out[(((size_t)b * Cout + co) * Ho + oy) * Wo + ox] = v;
      }
    }
  }
}

// 1x1 convolution with stride (the blocks' downsample branch), folded BatchNorm.  One thread per output element.
__global__ void k_spk_conv1x1(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ scale,
                              const float* __restrict__ shift, float* __restrict__ out, int B, int Cin, int H, int W, int Cout, int Ho, int Wo, int stride) {
  const size_t n = (size_t)B * Cout * Ho * Wo;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % Wo), oy = (int)((i / Wo) % Ho), co = (int)((i / ((size_t)Wo * Ho)) % Cout), b = (int)(i / ((size_t)Wo * Ho * Cout));
    const float* xp = x + ((size_t)b * Cin * H + (size_t)oy * stride) * W + (size_t)ox * stride;
    const float* wp = w + (size_t)co * Cin;
    float a = 0.f;
    for (int ci = 0; ci < Cin; ++ci) a = fmaf(wp[ci], xp[(size_t)ci * H * W], a);
    out[i] = a * scale[co] + shift[co];
  }
}

__device__ __forceinline__ float spk_block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = 0.f;
  for (int i = 0; i < nw; ++i) t += red[i];
  return t;
}
__device__ __forceinline__ float spk_block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = red[0];
  for (int i = 1; i < nw; ++i) t = fmaxf(t, red[i]);
  return t;
}

// SimAMBasicBlock tail (speaker_cloning.py:83-97): out = relu(SimAM(x) + shortcut), SimAM(X) = X * sigmoid(d / (4 (v + lambda)) + 0.5),
// d = (X - mean)^2 over the H x W plane, v = sum(d) / (H W - 1).  grid = B * C planes, block = 256.
__global__ void __launch_bounds__(256) k_spk_simam_add_relu(const float* __restrict__ x, const float* __restrict__ shortcut, float* __restrict__ out,
                                                            int HW, float lambda_p) {
  __shared__ float red[8];
  const size_t base = (size_t)blockIdx.x * HW;
  float s = 0.f;
  for (int i = threadIdx.x; i < HW; i += 256) s += x[base + i];
  const float mean = spk_block_sum(s, red) / (float)HW;
  float q = 0.f;
  for (int i = threadIdx.x; i < HW; i += 256) { const float d = x[base + i] - mean; q = fmaf(d, d, q); }
  const float v = spk_block_sum(q, red) / (float)(HW - 1);
  const float inv = 1.0f / (4.0f * (v + lambda_p));
  for (int i = threadIdx.x; i < HW; i += 256) {
    const float xv = x[base + i], d = (xv - mean) * (xv - mean);
    const float e = d * inv + 0.5f;
    const float g = 1.0f / (1.0f + expf(-e));
    out[base + i] = fmaxf(fmaf(xv, g, shortcut[base + i]), 0.f);
  }
}

// ASP attention, first half (speaker_cloning.py:45-49): h[b][o][t] = bn(relu(b1[o] + sum_i W1[o][i] x[b][i][t])).  One thread per (o, t).
__global__ void k_spk_asp_hidden(const float* __restrict__ x, const float* __restrict__ w1, const float* __restrict__ b1,
                                 const float* __restrict__ bn_scale, const float* __restrict__ bn_shift, float* __restrict__ h, int C, int T, int Hd) {
  const int b = blockIdx.z, o = blockIdx.y, t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const float* xp = x + (size_t)b * C * T + t;
  const float* wp = w1 + (size_t)o * C;
  float a = b1[o];
  for (int i = 0; i < C; ++i) a = fmaf(wp[i], xp[(size_t)i * T], a);
  h[((size_t)b * Hd + o) * T + t] = fmaxf(a, 0.f) * bn_scale[o] + bn_shift[o];
}

// ASP second half (speaker_cloning.py:49-60): logits[t] = b2[c] + sum_j W2[c][j] h[b][j][t]; w = softmax_t; mu = sum x w;
// sg = sqrt(clamp(sum x^2 w - mu^2, 1e-5)); out[b] = [mu (C) | sg (C)].  grid = (C, B), block = 128, dyn smem = T floats.
__global__ void __launch_bounds__(128) k_spk_asp_pool(const float* __restrict__ x, const float* __restrict__ h, const float* __restrict__ w2,
                                                      const float* __restrict__ b2, float* __restrict__ out, int C, int T, int Hd) {
  extern __shared__ float lg[];
  __shared__ float red[4];
  const int c = blockIdx.x, b = blockIdx.y;
  const float* wp = w2 + (size_t)c * Hd;
  float mx = -INFINITY;
  for (int t = threadIdx.x; t < T; t += 128) {
    float a = b2[c];
    for (int j = 0; j < Hd; ++j) a = fmaf(wp[j], h[((size_t)b * Hd + j) * T + t], a);
    lg[t] = a;
    mx = fmaxf(mx, a);
  }
  mx = spk_block_max(mx, red);
  float se = 0.f;
  for (int t = threadIdx.x; t < T; t += 128) { const float e = expf(lg[t] - mx); lg[t] = e; se += e; }
  se = spk_block_sum(se, red);
  const float* xp = x + ((size_t)b * C + c) * T;
  float m1 = 0.f, m2 = 0.f;
  for (int t = threadIdx.x; t < T; t += 128) { const float wv = lg[t] / se, xv = xp[t]; m1 = fmaf(xv, wv, m1); m2 = fmaf(xv * xv, wv, m2); }
  m1 = spk_block_sum(m1, red);
  m2 = spk_block_sum(m2, red);
  if (threadIdx.x == 0) {
    out[(size_t)b * 2 * C + c] = m1;
    out[(size_t)b * 2 * C + C + c] = sqrtf(fmaxf(m2 - m1 * m1, 1e-5f));
  }
}

// nn.Linear: out[b][o] = bias[o] + sum_i W[o][i] x[b][i].  One warp per output; grid = (ceil(O / 8), B), block = 256.
__global__ void __launch_bounds__(256) k_spk_linear(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                                    float* __restrict__ out, int I, int O) {
  const int o = blockIdx.x * 8 + (threadIdx.x >> 5), b = blockIdx.y, lane = threadIdx.x & 31;
  if (o >= O) return;
  const float* wp = w + (size_t)o * I;
  const float* xp = x + (size_t)b * I;
  float a = 0.f;
  for (int i = lane; i < I; i += 32) a = fmaf(wp[i], xp[i], a);
  for (int s = 16; s > 0; s >>= 1) a += __shfl_xor_sync(0xffffffffu, a, s);
  if (lane == 0) out[(size_t)b * O + o] = a + (bias ? bias[o] : 0.f);
}

int spk_done() { return cudaGetLastError() == cudaSuccess ? ZB_OK : ZB_ERR_CUDA; }

}  // namespace

extern "C" {

int zb_spk_conv2d(const float* d_x, const float* d_w, const float* d_scale, const float* d_shift, float* d_out, int B, int Cin, int H, int W,
                  int Cout, int ksize, int stride, int relu, void* stream) {
  if (!d_x || !d_w || !d_scale || !d_shift || !d_out || B <= 0 || Cin <= 0 || Cout <= 0 || H <= 0 || W <= 0) return ZB_ERR_ARG;
  if ((ksize != 3 && ksize != 1) || (stride != 1 && stride != 2)) return ZB_ERR_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  if (ksize == 3) {
    const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
    dim3 grid((Wo + SPK_TW - 1) / SPK_TW, (Ho + SPK_TH - 1) / SPK_TH, B * ((Cout + SPK_CO - 1) / SPK_CO));
    k_spk_conv3x3<<<grid, 256, 0, s>>>(d_x, d_w, d_scale, d_shift, d_out, Cin, H, W, Cout, Ho, Wo, stride, relu);
  } else {
    if (relu) return ZB_ERR_ARG;
    const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
    const size_t n = (size_t)B * Cout * Ho * Wo;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    k_spk_conv1x1<<<(unsigned)blocks, 256, 0, s>>>(d_x, d_w, d_scale, d_shift, d_out, B, Cin, H, W, Cout, Ho, Wo, stride);
  }
  return spk_done();
}

int zb_spk_simam_add_relu(const float* d_x, const float* d_shortcut, float* d_out, int B, int C, int H, int W, float lambda_p, void* stream) {
  if (!d_x || !d_shortcut || !d_out || B <= 0 || C <= 0 || H * W < 2) return ZB_ERR_ARG;
  k_spk_simam_add_relu<<<B * C, 256, 0, (cudaStream_t)stream>>>(d_x, d_shortcut, d_out, H * W, lambda_p);
  return spk_done();
}

int zb_spk_asp(const float* d_x, const float* d_w1, const float* d_b1, const float* d_bn_scale, const float* d_bn_shift, const float* d_w2,
               const float* d_b2, float* d_hidden, float* d_out, int B, int C, int T, int hidden, void* stream) {
  if (!d_x || !d_w1 || !d_b1 || !d_bn_scale || !d_bn_shift || !d_w2 || !d_b2 || !d_hidden || !d_out || B <= 0 || C <= 0 || T <= 0 || hidden <= 0) return ZB_ERR_ARG;
  if ((size_t)T * 4 > 40 * 1024) return ZB_ERR_ARG;                 // the pooled time axis has to fit the softmax scratch (10 240 frames)
  cudaStream_t s = (cudaStream_t)stream;
  k_spk_asp_hidden<<<dim3((T + 127) / 128, hidden, B), 128, 0, s>>>(d_x, d_w1, d_b1, d_bn_scale, d_bn_shift, d_hidden, C, T, hidden);
  k_spk_asp_pool<<<dim3(C, B), 128, (size_t)T * 4, s>>>(d_x, d_hidden, d_w2, d_b2, d_out, C, T, hidden);
  return spk_done();
}

int zb_spk_linear(const float* d_x, const float* d_w, const float* d_bias, float* d_out, int B, int in_features, int out_features, void* stream) {
  if (!d_x || !d_w || !d_out || B <= 0 || in_features <= 0 || out_features <= 0) return ZB_ERR_ARG;
  k_spk_linear<<<dim3((out_features + 7) / 8, B), 256, 0, (cudaStream_t)stream>>>(d_x, d_w, d_bias, d_out, in_features, out_features);
  return spk_done();
}

}  // extern "C"
