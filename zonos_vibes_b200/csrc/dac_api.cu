// C ABI of libzonos_b200.so — DAC 44 kHz codec: decoder (DACAutoencoder.decode, zonos/autoencoder.py:25-27) and, as the
// optional row f3, encoder + residual vector quantiser (DACAutoencoder.encode, autoencoder.py:22-23).
#include <stdarg.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "dac_simt.cuh"
#include "dac_tc.cuh"
#include "dac_tc2.cuh"
#include "dac_ru.cuh"

namespace {

constexpr int kNq = 9, kCodes = 1024, kCodeDim = 8, kHidden = 1024, kDecoder = 1536;
constexpr int kStrides[4] = {8, 8, 4, 2};
constexpr int kDil[3] = {1, 3, 9};
constexpr int kEncHidden = 64;
constexpr int kEncStrides[4] = {2, 4, 8, 8};      // DacConfig.downsampling_ratios

struct RawT {
  float* ptr = nullptr;
  int64_t numel = 0;
  bool provided = false;
};

struct ConvW {            // one Conv1d / ConvTranspose1d
  std::string name;       // HF prefix, e.g. "decoder.block.0.res_unit1.conv1"
  int Cin = 0, Cout = 0, K = 0, dil = 1, pad = 0, stride = 1;   // stride > 1 => transposed
  float* w_kio = nullptr;   // fp32 [K][Cin][Cout]
  __half* w_koi = nullptr;  // fp16 [K][Cout][Cin]
  float* bias = nullptr;    // -> raw tensor
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

}  // namespace

struct zb_dac {
  int device = 0;
  // tuning knobs of the tcgen05 path (zb_dac_set_option; defaults are the measured best, profiles/r01_summary.md)
  int ew_multi = 8;         // epilogue warps of the persistent kernel on multi-tap layers
  int ew_k1 = 8;            // ... on k=1 layers
  int wide_cout = 768;      // layers with Cout >= this run on the one-tile-per-CTA kernel (impl 0)
  int mt2_min_waves = 8;    // 256-row tiles only when that still leaves this many tiles per SM
  int fuse_ru = 1;          // 96 / 192-channel ResidualUnits as ONE kernel (k7 -> snake -> k1 -> + residual, dac_ru.cuh)
  int k1_wide = 1;          // k = 1 layers: all output channels in one tile (A/B switch)
  int max_group = 8;        // utterances decoded together: one launch per layer covers all their tiles (workspace = 5 x 49152 x T x group halves)
  float conv2_bias_h = 0.f; // host copy of decoder.conv2.bias (read back once in zb_dac_finalize)
  ZbErr err;
  std::map<std::string, RawT> raw;
  bool finalized = false;
  float* table = nullptr;           // folded RVQ tables [9][1024][1024] fp32
  ConvW conv1, conv2;
  ConvW convt[4];
  ConvW ru1[4][3], ru2[4][3];       // res_unit{u}.conv1 / .conv2
  // workspace (one utterance at a time)
  int ws_T = 0;
  float* f32buf[3] = {nullptr, nullptr, nullptr};
  __half* f16buf[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  int keep = 0;
  std::map<std::string, std::pair<float*, int64_t>> dbg;   // fp32 [C][T_l] copies (keep=1)
  EncodeTiledFn encode = nullptr;
  int64_t launches = 0;
  // ---- encoder side (optional: zb_dac_encode needs every "encoder.*" / "*.in_proj.*" tensor; decode does not) ----
  std::map<std::string, RawT> raw_enc;
  bool enc_ready = false;
  ConvW e_conv1, e_conv2, e_down[4], e_ru1[4][3], e_ru2[4][3];
  const float** q_ptrs = nullptr;   // device: [5][9] pointers (in_proj w, in_proj b, codebook, out_proj w, out_proj b)
};

static std::string g_dac_create_error;

namespace {

const float* rawp(zb_dac* d, const std::string& n) { return d->raw[n].ptr; }

void declare(zb_dac* d, const std::string& n, int64_t numel) {
  RawT t;
  t.numel = numel;
  d->raw[n] = t;
}

void declare_conv(zb_dac* d, ConvW& c, const std::string& name, int Cin, int Cout, int K, int dil, int stride) {
  c.name = name; c.Cin = Cin; c.Cout = Cout; c.K = K; c.dil = dil; c.stride = stride;
  c.pad = stride > 1 ? (stride + 1) / 2 : ((K - 1) * dil) / 2;
  declare(d, name + ".weight", (int64_t)Cin * Cout * K);
  declare(d, name + ".bias", Cout);
}

void declare_enc(zb_dac* d, const std::string& n, int64_t numel) {
  RawT t;
  t.numel = numel;
  d->raw_enc[n] = t;
}
// encoder conv: stride > 1 here means a strided (down-sampling) Conv1d with K = 2 * stride, pad = ceil(stride / 2)
void declare_conv_enc(zb_dac* d, ConvW& c, const std::string& name, int Cin, int Cout, int K, int dil, int stride) {
  c.name = name; c.Cin = Cin; c.Cout = Cout; c.K = K; c.dil = dil; c.stride = stride;
  c.pad = stride > 1 ? (stride + 1) / 2 : ((K - 1) * dil) / 2;
  declare_enc(d, name + ".weight", (int64_t)Cin * Cout * K);
  declare_enc(d, name + ".bias", Cout);
}

int dbg_store_f32(zb_dac* d, const std::string& name, const float* x_tc, int T, int C, cudaStream_t s) {
  auto& e = d->dbg[name];
  const int64_t n = (int64_t)T * C;
  if (e.second != n) {
    if (e.first) cudaFree(e.first);
    ZB_CUDA(d->err, cudaMalloc((void**)&e.first, n * 4));
    e.second = n;
  }
  k_dac_to_channel_major<<<ZB_NUM_SMS * 4, 256, 0, s>>>(x_tc, e.first, T, C);
  return ZB_OK;
}
int dbg_store_f16(zb_dac* d, const std::string& name, const __half* x_tc, int T, int C, cudaStream_t s) {
  auto& e = d->dbg[name];
  const int64_t n = (int64_t)T * C;
  if (e.second != n) {
    if (e.first) cudaFree(e.first);
    ZB_CUDA(d->err, cudaMalloc((void**)&e.first, n * 4));
    e.second = n;
  }
  k_dac_f16_to_channel_major<<<ZB_NUM_SMS * 4, 256, 0, s>>>(x_tc, e.first, T, C);
  return ZB_OK;
}

int ensure_workspace(zb_dac* d, int T) {
  if (T <= d->ws_T) return ZB_OK;
  ZB_CUDA(d->err, cudaDeviceSynchronize());
  for (auto& p : d->f32buf) { if (p) cudaFree(p); p = nullptr; }
  for (auto& p : d->f16buf) { if (p) cudaFree(p); p = nullptr; }
  d->ws_T = 0;
  const size_t elems = (size_t)49152 * T;    // max over levels of T_l * C_l (256T*192 = 512T*96)
  for (auto& p : d->f16buf) ZB_CUDA(d->err, cudaMalloc((void**)&p, elems * sizeof(__half)));
  d->ws_T = T;
  return ZB_OK;
}
int ensure_workspace_f32(zb_dac* d, int T) {
  if (d->f32buf[0] && T <= d->ws_T) return ZB_OK;
  int rc = ensure_workspace(d, T);
  if (rc) return rc;
  const size_t elems = (size_t)49152 * d->ws_T;
  for (auto& p : d->f32buf) { if (p) cudaFree(p); p = nullptr; }
  for (auto& p : d->f32buf) ZB_CUDA(d->err, cudaMalloc((void**)&p, elems * sizeof(float)));
  return ZB_OK;
}

// ---------------- SIMT path ----------------
int conv_simt(zb_dac* d, const ConvW& c, const float* x, int T_in, const float* alpha, const float* residual, float* out,
              int tanh_out, cudaStream_t s) {
  ConvSimtParams p;
  memset(&p, 0, sizeof(p));
  p.x = x; p.w = c.w_kio; p.bias = c.bias; p.alpha = alpha; p.residual = residual; p.out = out;
  p.T_in = T_in; p.T_out = T_in * c.stride; p.Cin = c.Cin; p.Cout = c.Cout; p.K = c.K; p.dil = c.dil; p.pad = c.pad; p.stride = c.stride;
  p.tanh_out = tanh_out;
  dim3 grid((p.T_out + DSC_TT - 1) / DSC_TT, (c.Cout + DSC_TC - 1) / DSC_TC);
  if (c.stride > 1) k_dac_convt_simt<<<grid, 256, 0, s>>>(p);
  else k_dac_conv_simt<<<grid, 256, 0, s>>>(p);
  d->launches++;
  ZB_CUDA(d->err, cudaGetLastError());
  return ZB_OK;
}

int decode_one_simt(zb_dac* d, const long long* codes, float* wav, int T, cudaStream_t s) {
  int rc;
  float *a = d->f32buf[0], *b = d->f32buf[1], *y = d->f32buf[2];
  k_dac_latent<<<T, 256, 0, s>>>(codes, d->table, a, nullptr, T, kHidden, kNq, kCodes);
  d->launches++;
  if (d->keep && (rc = dbg_store_f32(d, "latent", a, T, kHidden, s))) return rc;
  if ((rc = conv_simt(d, d->conv1, a, T, nullptr, nullptr, b, 0, s))) return rc;
  if (d->keep && (rc = dbg_store_f32(d, "conv1", b, T, kDecoder, s))) return rc;
  std::swap(a, b);                                      // a = current activation (raw)
  int Tl = T;
  for (int j = 0; j < 4; ++j) {
    const std::string bp = "decoder.block." + std::to_string(j);
    if ((rc = conv_simt(d, d->convt[j], a, Tl, rawp(d, bp + ".snake1.alpha"), nullptr, b, 0, s))) return rc;
    Tl *= kStrides[j];
    std::swap(a, b);
    for (int u = 0; u < 3; ++u) {
      const std::string rp = bp + ".res_unit" + std::to_string(u + 1);
      if ((rc = conv_simt(d, d->ru1[j][u], a, Tl, rawp(d, rp + ".snake1.alpha"), nullptr, y, 0, s))) return rc;
      if ((rc = conv_simt(d, d->ru2[j][u], y, Tl, rawp(d, rp + ".snake2.alpha"), a, b, 0, s))) return rc;
      std::swap(a, b);
    }
    if (d->keep && (rc = dbg_store_f32(d, "block" + std::to_string(j), a, Tl, d->convt[j].Cout, s))) return rc;
  }
  // final conv writes [T*512][1] == the waveform row of this utterance
  if ((rc = conv_simt(d, d->conv2, a, Tl, rawp(d, "decoder.snake1.alpha"), nullptr, wav, 1, s))) return rc;
  return ZB_OK;
}

// ---------------- tcgen05 path ----------------
int make_map(zb_dac* d, CUtensorMap* m, const void* base, uint64_t inner, uint64_t rows, uint32_t box_rows) {
  cuuint64_t dims[2] = {inner, rows};
  cuuint64_t strides[1] = {inner * 2};
  cuuint32_t box[2] = {TC_BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = d->encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return d->err.fail(ZB_ERR_CUDA, "cuTensorMapEncodeTiled failed: %d (inner=%llu rows=%llu box=%u)", (int)r,
                                            (unsigned long long)inner, (unsigned long long)rows, box_rows);
  return ZB_OK;
}

// activations of a group of utterances: [nb][T][C] fp16, as a 3-D map {C, T, nb}: rows outside [0, T) of an utterance are
// zero-filled by TMA (the conv's zero padding), so a tile never reads its neighbour utterance
int make_map_x(zb_dac* d, CUtensorMap* m, const void* base, uint64_t C, uint64_t T, uint64_t nb, uint32_t box_rows) {
  cuuint64_t dims[3] = {C, T, nb};
  cuuint64_t strides[2] = {C * 2, T * C * 2};
  cuuint32_t box[3] = {TC_BK, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = d->encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return d->err.fail(ZB_ERR_CUDA, "cuTensorMapEncodeTiled (activations) failed: %d (C=%llu T=%llu nb=%llu box=%u)", (int)r,
                                            (unsigned long long)C, (unsigned long long)T, (unsigned long long)nb, box_rows);
  return ZB_OK;
}

int pick_n_tile(int Cout) {
  if (Cout % 256 == 0) return 256;
  if (Cout % 192 == 0) return 192;
  if (Cout % 128 == 0) return 128;
  if (Cout % 96 == 0) return 96;
  if (Cout % 64 == 0) return 64;
  return 32;
}

int conv_tc(zb_dac* d, const ConvW& c, const __half* x, int nb, int T_in, const __half* residual, __half* out_raw, __half* out_act,
            const float* alpha_next, cudaStream_t s) {
  TcConvParams p;
  memset(&p, 0, sizeof(p));
  p.ntaps = c.stride > 1 ? 2 : c.K;
  p.dil = c.dil; p.pad = c.pad; p.stride = c.stride;
  p.kchunks = (c.Cin + TC_BK - 1) / TC_BK;
  p.T_out = T_in * c.stride;
  p.m_rows = c.stride > 1 ? T_in + 1 : T_in;
  p.Cout = c.Cout;
  p.n_tile = pick_n_tile(c.Cout);
  p.tmem_cols = p.n_tile <= 32 ? 32 : p.n_tile <= 64 ? 64 : p.n_tile <= 128 ? 128 : 256;
  const int stage_bytes = TC_A_BYTES + p.n_tile * 128;
  int stages = (108 * 1024 - TC_TAB_BYTES) / stage_bytes;   // two CTAs per SM
  if (stages < 2) stages = 2;
  if (stages > 8) stages = 8;
  p.stages = stages;
  p.bias = c.bias; p.residual = residual; p.out_raw = out_raw; p.out_act = out_act; p.alpha_next = alpha_next;
  CUtensorMap mx, mw;
  int rc;
  if ((rc = make_map_x(d, &mx, x, c.Cin, T_in, nb, TC_BM))) return rc;
  if ((rc = make_map(d, &mw, c.w_koi, c.Cin, (uint64_t)c.K * c.Cout, p.n_tile))) return rc;
  const size_t smem = 1024 + (size_t)stages * stage_bytes + 256 + TC_TAB_BYTES;
  p.m_tiles = (p.m_rows + TC_BM - 1) / TC_BM;
  dim3 grid(p.m_tiles * nb, c.Cout / p.n_tile, c.stride > 1 ? c.stride : 1);
  k_dac_conv_tc<<<grid, TC_THREADS, smem, s>>>(mx, mw, p);
  d->launches++;
  ZB_CUDA(d->err, cudaGetLastError());
  return ZB_OK;
}


// second-generation persistent kernel (dac_tc2.cuh): window reuse across taps, 256-row tiles, double-buffered TMEM
int conv_tc2(zb_dac* d, const ConvW& c, const __half* x, int nb, int T_in, const __half* residual, __half* out_raw, __half* out_act,
             const float* alpha_next, cudaStream_t s) {
  if (c.Cout > TC2_MAX_COUT) return d->err.fail(ZB_ERR_ARG, "conv_tc2: Cout %d > %d", c.Cout, TC2_MAX_COUT);
  TcConv2Params p;
  memset(&p, 0, sizeof(p));
  p.ntaps = c.stride > 1 ? 2 : c.K;
  p.dil = c.dil; p.pad = c.pad; p.stride = c.stride;
  p.kchunks = (c.Cin + TC_BK - 1) / TC_BK;
  p.T_out = T_in * c.stride;
  p.m_rows = c.stride > 1 ? T_in + 1 : T_in;
  p.Cout = c.Cout;
  p.n_tile = c.Cout % 128 == 0 ? 128 : c.Cout % 96 == 0 ? 96 : c.Cout % 64 == 0 ? 64 : 32;
  // k = 1 layers are pure streams (Y + residual in, raw + activated out; the weights are nothing): every output-channel
  // tile re-reads the whole activation window, so they take ALL channels in one tile (up to 256 accumulator columns,
  // 128-row tiles) - the ncu capture of the 192-channel k = 1 layer showed 846 MB read where 506 MB are needed
  // (profiles/r02_summary.md).  384 channels: two tiles of 192 instead of three of 128.
  const bool k1 = c.K == 1 && c.stride == 1 && d->k1_wide;
  if (k1) p.n_tile = c.Cout <= 256 ? c.Cout : (c.Cout % 192 == 0 ? 192 : p.n_tile);
  p.n_tiles = c.Cout / p.n_tile;
  p.phases = c.stride > 1 ? c.stride : 1;
  p.acc_cols = k1 ? 256 : 128; p.acc_stages = 2;
  // 256-row tiles halve the weight traffic; keep 128-row tiles when there would be too few tiles to balance 148 SMs
  p.nb = nb;
  const long tiles2 = (long)((p.m_rows + 2 * TC_BM - 1) / (2 * TC_BM)) * p.n_tiles * p.phases * nb;
  p.mt = (!k1 && tiles2 >= (long)d->mt2_min_waves * ZB_NUM_SMS) ? 2 : 1;
  p.m_tiles = (p.m_rows + p.mt * TC_BM - 1) / (p.mt * TC_BM);
  p.lead = c.stride > 1 ? p.ntaps - 1 : c.pad;
  const int win_rows = p.mt * TC_BM + (c.stride > 1 ? p.ntaps - 1 : (c.K - 1) * c.dil);
  p.n_aload = win_rows > 256 ? 2 : 1;
  p.box_rows = (((win_rows + p.n_aload - 1) / p.n_aload) + 7) & ~7;
  p.a_bytes = p.n_aload * p.box_rows * 128;
  p.a_slots = k1 ? TC2_MAX_A_SLOTS : (p.kchunks >= 3 ? 3 : 2);      // k = 1: bytes in flight are what bounds the stream
  const int b_bytes = p.n_tile * 128;
  const int budget = 227 * 1024 - 1024 - 512 - TC2_TAB_BYTES;
  if (p.a_slots * p.a_bytes + 4 * b_bytes > budget) p.a_slots = 2;
  int bs = (budget - p.a_slots * p.a_bytes) / b_bytes;
  if (bs > TC2_MAX_B_STAGES) bs = TC2_MAX_B_STAGES;
  if (bs < 2) return d->err.fail(ZB_ERR_ARG, "conv_tc2: shared memory budget");
  p.b_stages = bs;
  p.bias = c.bias; p.residual = residual; p.out_raw = out_raw; p.out_act = out_act; p.alpha_next = alpha_next;
  CUtensorMap mx, mw;
  int rc;
  if ((rc = make_map_x(d, &mx, x, c.Cin, T_in, nb, p.box_rows))) return rc;
  if ((rc = make_map(d, &mw, c.w_koi, c.Cin, (uint64_t)c.K * c.Cout, p.n_tile))) return rc;
  const size_t smem = 1024 + (size_t)p.a_slots * p.a_bytes + (size_t)p.b_stages * b_bytes + 512 + TC2_TAB_BYTES;
  const long tiles = (long)p.m_tiles * p.n_tiles * p.phases * nb;
  const int grid = (int)(tiles < ZB_NUM_SMS ? tiles : ZB_NUM_SMS);
  if ((p.ntaps > 1 ? d->ew_multi : d->ew_k1) == 16) k_dac_conv_tc2<16><<<grid, 64 + 32 * 16, smem, s>>>(mx, mw, p);
  else k_dac_conv_tc2<8><<<grid, 64 + 32 * 8, smem, s>>>(mx, mw, p);
  d->launches++;
  ZB_CUDA(d->err, cudaGetLastError());
  return ZB_OK;
}

// Fused ResidualUnit (dac_ru.cuh): x_act = snake1(x) [nb][T][C] in, x raw residual in, out raw (optional) + activated out.
bool ru_fusable(const zb_dac* d, const ConvW& c1, const ConvW& c2) {
  return d->fuse_ru && (c1.Cout == 96 || c1.Cout == 192) && c1.Cin == c1.Cout && c2.Cin == c1.Cout && c2.Cout == c1.Cout && c1.K == 7 && c2.K == 1 &&
         c1.stride == 1 && c2.stride == 1;
}
int conv_ru(zb_dac* d, const ConvW& c1, const ConvW& c2, const __half* x_act, const __half* x_raw, int nb, int T, __half* out_raw, __half* out_act,
            const float* alpha2, const float* alpha_next, cudaStream_t s) {
  RuParams p;
  memset(&p, 0, sizeof(p));
  p.C = c1.Cout; p.mt = p.C == 96 ? 2 : 1; p.kchunks = (p.C + TC_BK - 1) / TC_BK;
  p.dil = c1.dil; p.pad = c1.pad; p.T = T; p.nb = nb;
  const int tile_rows = p.mt * TC_BM;
  p.m_tiles = (T + tile_rows - 1) / tile_rows;
  const int win_rows = tile_rows + (c1.K - 1) * c1.dil;
  p.n_aload = win_rows > 256 ? 2 : 1;
  p.box_rows = (((win_rows + p.n_aload - 1) / p.n_aload) + 7) & ~7;
  p.a_bytes = p.n_aload * p.box_rows * 128;
  p.a_slots = p.C == 96 ? 2 : 3;
  const int b_bytes = p.C * 128, y_bytes = p.kchunks * tile_rows * 128;
  const int budget = 227 * 1024 - 1024 - 512 - RU_TAB_BYTES - y_bytes - p.a_slots * p.a_bytes;
  int bs = budget / b_bytes;
  if (bs > RU_MAX_B_STAGES) bs = RU_MAX_B_STAGES;
  if (bs < 3) return d->err.fail(ZB_ERR_ARG, "conv_ru: shared memory budget");
  p.b_stages = bs;
  p.bias1 = c1.bias; p.alpha2 = alpha2; p.bias2 = c2.bias; p.alpha_next = alpha_next;
  p.residual = x_raw; p.out_raw = out_raw; p.out_act = out_act;
  CUtensorMap mx, mw1, mw2;
  int rc;
  if ((rc = make_map_x(d, &mx, x_act, p.C, T, nb, p.box_rows))) return rc;
  if ((rc = make_map(d, &mw1, c1.w_koi, p.C, (uint64_t)c1.K * p.C, p.C))) return rc;
  if ((rc = make_map(d, &mw2, c2.w_koi, p.C, (uint64_t)p.C, RU_N2))) return rc;
  const size_t smem = 1024 + (size_t)p.a_slots * p.a_bytes + (size_t)p.b_stages * b_bytes + (size_t)y_bytes + 512 + RU_TAB_BYTES;
  const long tiles = (long)p.m_tiles * nb;
  const int grid = (int)(tiles < ZB_NUM_SMS ? tiles : ZB_NUM_SMS);
  k_dac_ru_fused<<<grid, RU_THREADS, smem, s>>>(mx, mw1, mw2, p);
  d->launches++;
  ZB_CUDA(d->err, cudaGetLastError());
  return ZB_OK;
}

// One group of `nb` utterances: every layer is ONE launch whose tiles cover all of them (tile index = (utterance, time
// tile, channel tile, phase)), so the tail of a layer is paid once per group instead of once per utterance, and the small
// layers (conv1 at T = 2580: 120 tiles for 148 SMs) fill the machine.  `bufs` = the five fp16 activation buffers of
// 49152 * T * nb halves each (the context workspace, or the static array of zb_dac_decode_static);
// keep_dbg = copy intermediates of the group's LAST utterance for zb_dac_get_intermediate (allocates: context workspace only).
int decode_group_tc(zb_dac* d, const long long* codes, float* wav, int nb, int T, cudaStream_t s, int impl, __half* const* bufs, bool keep_dbg) {
  auto conv = [&](const ConvW& c, const __half* x, int T_in, const __half* residual, __half* out_raw, __half* out_act, const float* alpha_next) {
    // impl 0: per layer - the wide layers (Cout >= 768) run best on the N=256 one-tile-per-CTA kernel (two CTAs per SM),
    // everything else on the persistent window-reuse kernel; impl 2 / 3 force one kernel everywhere (A/B runs)
    const bool v1 = impl == 2 || (impl == 0 && c.Cout >= d->wide_cout);
    return v1 ? conv_tc(d, c, x, nb, T_in, residual, out_raw, out_act, alpha_next, s) : conv_tc2(d, c, x, nb, T_in, residual, out_raw, out_act, alpha_next, s);
  };
  int rc;
  __half *A = bufs[0], *R = bufs[1], *Y = bufs[2], *A2 = bufs[3], *R2 = bufs[4];
  const bool keep = keep_dbg;
  const size_t last = (size_t)(nb - 1);                          // intermediates are kept for the last utterance of the group
  // latent (raw, fp16) -> conv1 -> act = snake(., block0.snake1)
  k_dac_latent<<<nb * T, 256, 0, s>>>(codes, d->table, nullptr, R, T, kHidden, kNq, kCodes);
  d->launches++;
  if (keep && (rc = dbg_store_f16(d, "latent", R + last * T * kHidden, T, kHidden, s))) return rc;
  if ((rc = conv(d->conv1, R, T, nullptr, keep ? R2 : nullptr, A, rawp(d, "decoder.block.0.snake1.alpha")))) return rc;
  if (keep && (rc = dbg_store_f16(d, "conv1", R2 + last * T * kDecoder, T, kDecoder, s))) return rc;
  int Tl = T;
  for (int j = 0; j < 4; ++j) {
    const std::string bp = "decoder.block." + std::to_string(j);
    // A (snaked) -> ConvT -> R (raw, residual of unit 1), A2 = snake(., ru1.snake1)
    if ((rc = conv(d->convt[j], A, Tl, nullptr, R, A2, rawp(d, bp + ".res_unit1.snake1.alpha")))) return rc;
    Tl *= kStrides[j];
    std::swap(A, A2);
    for (int u = 0; u < 3; ++u) {
      const std::string rp = bp + ".res_unit" + std::to_string(u + 1);
      // conv2 (k1) + residual R -> R2 (raw, only if a later unit needs it) and A2 = snake(., next alpha)
      std::string next_alpha;
      if (u < 2) next_alpha = bp + ".res_unit" + std::to_string(u + 2) + ".snake1.alpha";
      else if (j < 3) next_alpha = "decoder.block." + std::to_string(j + 1) + ".snake1.alpha";
      else next_alpha = "decoder.snake1.alpha";
      const bool need_raw = (u < 2) || keep;
      const bool fused = impl == 0 && ru_fusable(d, d->ru1[j][u], d->ru2[j][u]);
      // conv1 (k7, dilated): A -> Y = snake(., snake2)
      if (!fused && (rc = conv(d->ru1[j][u], A, Tl, nullptr, nullptr, Y, rawp(d, rp + ".snake2.alpha")))) return rc;
      if (impl == 0 && ru_fusable(d, d->ru1[j][u], d->ru2[j][u])) {
        // fused unit: A (activated) + R (raw residual) -> R2 (raw, if needed) and A2 = snake(., next alpha); Y stays on chip
        if ((rc = conv_ru(d, d->ru1[j][u], d->ru2[j][u], A, R, nb, Tl, need_raw ? R2 : nullptr, A2, rawp(d, rp + ".snake2.alpha"), rawp(d, next_alpha), s))) return rc;
        std::swap(A, A2);
        std::swap(R, R2);
        continue;
      }
      if ((rc = conv(d->ru2[j][u], Y, Tl, R, need_raw ? R2 : nullptr, A2, rawp(d, next_alpha)))) return rc;
      std::swap(A, A2);
      std::swap(R, R2);
    }
    if (keep && (rc = dbg_store_f16(d, "block" + std::to_string(j), R + last * Tl * d->convt[j].Cout, Tl, d->convt[j].Cout, s))) return rc;
  }
  // final Conv1d(96 -> 1, k7) + tanh on the snaked activation (bias cached on the host at finalize: no sync here)
  const ConvW& c = d->conv2;
  {
    const size_t fsm = (size_t)(FC_TILE + c.K - 1) * ((c.Cin / 2) | 1) * 4 + (size_t)c.K * c.Cin * 4;
    k_dac_final_conv<<<dim3((Tl + FC_TILE - 1) / FC_TILE, nb), FC_TILE, fsm, s>>>(A, c.w_kio, d->conv2_bias_h, wav, Tl, c.Cin, c.K, c.pad);
  }
  d->launches++;
  ZB_CUDA(d->err, cudaGetLastError());
  return ZB_OK;
}

}  // namespace

// =============================================================================================
extern "C" {

const char* zb_dac_last_error(const zb_dac* d) { return d ? d->err.msg.c_str() : g_dac_create_error.c_str(); }

int zb_dac_create(int device, zb_dac** out) {
  if (!out) return ZB_ERR_ARG;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= device) {
    g_dac_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (libzonos_b200 has no CPU fallback)";
    return ZB_ERR_CUDA;
  }
  cudaSetDevice(device);
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  if (prop.major != 10) { g_dac_create_error = "libzonos_b200 is built for sm_100a only"; return ZB_ERR_CUDA; }
  zb_dac* d = new zb_dac();
  d->device = device;
  for (int k = 0; k < kNq; ++k) {
    const std::string q = "quantizer.quantizers." + std::to_string(k);
    declare(d, q + ".codebook.weight", (int64_t)kCodes * kCodeDim);
    declare(d, q + ".out_proj.weight", (int64_t)kHidden * kCodeDim);
    declare(d, q + ".out_proj.bias", kHidden);
  }
  declare_conv(d, d->conv1, "decoder.conv1", kHidden, kDecoder, 7, 1, 1);
  int C = kDecoder;
  for (int j = 0; j < 4; ++j) {
    const std::string bp = "decoder.block." + std::to_string(j);
    declare(d, bp + ".snake1.alpha", C);
    declare_conv(d, d->convt[j], bp + ".conv_t1", C, C / 2, 2 * kStrides[j], 1, kStrides[j]);
    C /= 2;
    for (int u = 0; u < 3; ++u) {
      const std::string rp = bp + ".res_unit" + std::to_string(u + 1);
      declare(d, rp + ".snake1.alpha", C);
      declare_conv(d, d->ru1[j][u], rp + ".conv1", C, C, 7, kDil[u], 1);
      declare(d, rp + ".snake2.alpha", C);
      declare_conv(d, d->ru2[j][u], rp + ".conv2", C, C, 1, 1, 1);
    }
  }
  declare(d, "decoder.snake1.alpha", C);
  declare_conv(d, d->conv2, "decoder.conv2", C, 1, 7, 1, 1);
  {
    // encoder (DacEncoder): conv1 1 -> 64 k7; 4 x [3 ResidualUnits(C), Snake, Conv1d C -> 2C k 2s stride s]; Snake; conv2 1024 -> 1024 k3
    for (int k = 0; k < kNq; ++k) {
      const std::string q = "quantizer.quantizers." + std::to_string(k);
      declare_enc(d, q + ".in_proj.weight", (int64_t)kCodeDim * kHidden);
      declare_enc(d, q + ".in_proj.bias", kCodeDim);
    }
    declare_conv_enc(d, d->e_conv1, "encoder.conv1", 1, kEncHidden, 7, 1, 1);
    int Ce = kEncHidden;
    for (int j = 0; j < 4; ++j) {
      const std::string bp = "encoder.block." + std::to_string(j);
      for (int u = 0; u < 3; ++u) {
        const std::string rp = bp + ".res_unit" + std::to_string(u + 1);
        declare_enc(d, rp + ".snake1.alpha", Ce);
        declare_conv_enc(d, d->e_ru1[j][u], rp + ".conv1", Ce, Ce, 7, kDil[u], 1);
        declare_enc(d, rp + ".snake2.alpha", Ce);
        declare_conv_enc(d, d->e_ru2[j][u], rp + ".conv2", Ce, Ce, 1, 1, 1);
      }
      declare_enc(d, bp + ".snake1.alpha", Ce);
      declare_conv_enc(d, d->e_down[j], bp + ".conv1", Ce, 2 * Ce, 2 * kEncStrides[j], 1, kEncStrides[j]);
      Ce *= 2;
    }
    declare_enc(d, "encoder.snake1.alpha", Ce);
    declare_conv_enc(d, d->e_conv2, "encoder.conv2", Ce, kHidden, 3, 1, 1);
  }
  // driver entry point for tensor-map encoding (no -lcuda link dependency)
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
    g_dac_create_error = "cuTensorMapEncodeTiled not available from the driver";
    delete d;
    return ZB_ERR_CUDA;
  }
  d->encode = reinterpret_cast<EncodeTiledFn>(fn);
  *out = d;
  return ZB_OK;
}

int zb_dac_destroy(zb_dac* d) {
  if (!d) return ZB_OK;
  cudaSetDevice(d->device);
  cudaDeviceSynchronize();
  for (auto& kv : d->raw) if (kv.second.ptr) cudaFree(kv.second.ptr);
  for (auto& kv : d->dbg) if (kv.second.first) cudaFree(kv.second.first);
  if (d->table) cudaFree(d->table);
  std::vector<ConvW*> all = {&d->conv1, &d->conv2};
  for (int j = 0; j < 4; ++j) { all.push_back(&d->convt[j]); for (int u = 0; u < 3; ++u) { all.push_back(&d->ru1[j][u]); all.push_back(&d->ru2[j][u]); } }
  all.push_back(&d->e_conv1); all.push_back(&d->e_conv2);
  for (int j = 0; j < 4; ++j) { all.push_back(&d->e_down[j]); for (int u = 0; u < 3; ++u) { all.push_back(&d->e_ru1[j][u]); all.push_back(&d->e_ru2[j][u]); } }
  for (ConvW* c : all) { if (c->w_kio) cudaFree(c->w_kio); if (c->w_koi) cudaFree(c->w_koi); }
  for (auto& kv : d->raw_enc) if (kv.second.ptr) cudaFree(kv.second.ptr);
  if (d->q_ptrs) cudaFree(d->q_ptrs);
  for (auto& p : d->f32buf) if (p) cudaFree(p);
  for (auto& p : d->f16buf) if (p) cudaFree(p);
  delete d;
  return ZB_OK;
}

int zb_dac_set_tensor(zb_dac* d, const char* name, const float* d_ptr, int64_t numel, void* stream) {
  if (!d || !name || !d_ptr) return ZB_ERR_ARG;
  auto it = d->raw.find(name);
  if (it == d->raw.end()) {
    it = d->raw_enc.find(name);
    if (it == d->raw_enc.end()) return d->err.fail(ZB_ERR_ARG, "unknown DAC tensor '%s'", name);
  }
  RawT& t = it->second;
  if (numel != t.numel) return d->err.fail(ZB_ERR_ARG, "DAC tensor '%s': numel %lld, expected %lld", name, (long long)numel, (long long)t.numel);
  ZB_CUDA(d->err, cudaSetDevice(d->device));
  if (!t.ptr) ZB_CUDA(d->err, cudaMalloc((void**)&t.ptr, (size_t)numel * 4));
  ZB_CUDA(d->err, cudaMemcpyAsync(t.ptr, d_ptr, (size_t)numel * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  t.provided = true;
  d->finalized = false;
  return ZB_OK;
}

int zb_dac_finalize(zb_dac* d, void* stream) {
  if (!d) return ZB_ERR_ARG;
  for (auto& kv : d->raw)
    if (!kv.second.provided) return d->err.fail(ZB_ERR_STATE, "DAC tensor '%s' was not provided", kv.first.c_str());
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(d->err, cudaSetDevice(d->device));
  if (!d->table) ZB_CUDA(d->err, cudaMalloc((void**)&d->table, (size_t)kNq * kCodes * kHidden * 4));
  for (int k = 0; k < kNq; ++k) {
    const std::string q = "quantizer.quantizers." + std::to_string(k);
    k_dac_fold<<<kCodes, 256, 0, s>>>(rawp(d, q + ".codebook.weight"), rawp(d, q + ".out_proj.weight"), rawp(d, q + ".out_proj.bias"),
                                      d->table + (size_t)k * kCodes * kHidden, kHidden, kCodes, kCodeDim);
  }
  std::vector<ConvW*> all = {&d->conv1, &d->conv2};
  for (int j = 0; j < 4; ++j) { all.push_back(&d->convt[j]); for (int u = 0; u < 3; ++u) { all.push_back(&d->ru1[j][u]); all.push_back(&d->ru2[j][u]); } }
  for (ConvW* c : all) {
    const size_t n = (size_t)c->Cin * c->Cout * c->K;
    if (!c->w_kio) ZB_CUDA(d->err, cudaMalloc((void**)&c->w_kio, n * 4));
    if (!c->w_koi) ZB_CUDA(d->err, cudaMalloc((void**)&c->w_koi, n * 2));
    c->bias = d->raw[c->name + ".bias"].ptr;
    k_dac_repack<<<ZB_NUM_SMS * 4, 256, 0, s>>>(rawp(d, c->name + ".weight"), c->w_kio, c->w_koi, c->Cout, c->Cin, c->K, c->stride > 1 ? 1 : 0);
  }
  // encoder side: only when every one of its tensors was provided (a decode-only user never sets them)
  d->enc_ready = false;
  {
    bool all_enc = true;
    for (auto& kv : d->raw_enc) all_enc = all_enc && kv.second.provided;
    if (all_enc) {
      std::vector<ConvW*> enc = {&d->e_conv1, &d->e_conv2};
      for (int j = 0; j < 4; ++j) { enc.push_back(&d->e_down[j]); for (int u = 0; u < 3; ++u) { enc.push_back(&d->e_ru1[j][u]); enc.push_back(&d->e_ru2[j][u]); } }
      for (ConvW* c : enc) {
        const size_t n = (size_t)c->Cin * c->Cout * c->K;
        if (!c->w_kio) ZB_CUDA(d->err, cudaMalloc((void**)&c->w_kio, n * 4));
        if (!c->w_koi) ZB_CUDA(d->err, cudaMalloc((void**)&c->w_koi, n * 2));
        c->bias = d->raw_enc[c->name + ".bias"].ptr;
        k_dac_repack<<<ZB_NUM_SMS * 4, 256, 0, s>>>(d->raw_enc[c->name + ".weight"].ptr, c->w_kio, c->w_koi, c->Cout, c->Cin, c->K, 0);
      }
      const float* h[5 * kNq];
      for (int k = 0; k < kNq; ++k) {
        const std::string q = "quantizer.quantizers." + std::to_string(k);
        h[0 * kNq + k] = d->raw_enc[q + ".in_proj.weight"].ptr;
        h[1 * kNq + k] = d->raw_enc[q + ".in_proj.bias"].ptr;
        h[2 * kNq + k] = rawp(d, q + ".codebook.weight");
        h[3 * kNq + k] = rawp(d, q + ".out_proj.weight");
        h[4 * kNq + k] = rawp(d, q + ".out_proj.bias");
      }
      if (!d->q_ptrs) ZB_CUDA(d->err, cudaMalloc((void**)&d->q_ptrs, sizeof(h)));
      ZB_CUDA(d->err, cudaMemcpyAsync(d->q_ptrs, h, sizeof(h), cudaMemcpyHostToDevice, s));
      ZB_CUDA(d->err, cudaStreamSynchronize(s));          // `h` is a stack array
      ZB_CUDA(d->err, cudaFuncSetAttribute(k_dac_conv_down_simt, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
      d->enc_ready = true;
    }
  }
  ZB_CUDA(d->err, cudaGetLastError());
  ZB_CUDA(d->err, cudaMemcpyAsync(&d->conv2_bias_h, d->conv2.bias, 4, cudaMemcpyDeviceToHost, s));
  ZB_CUDA(d->err, cudaStreamSynchronize(s));
  // opt-in to large dynamic smem for the tensor-core kernel
  ZB_CUDA(d->err, cudaFuncSetAttribute(k_dac_conv_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
  ZB_CUDA(d->err, cudaFuncSetAttribute(k_dac_conv_tc2<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  ZB_CUDA(d->err, cudaFuncSetAttribute(k_dac_conv_tc2<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  ZB_CUDA(d->err, cudaFuncSetAttribute(k_dac_final_conv, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  ZB_CUDA(d->err, cudaFuncSetAttribute(k_dac_ru_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  d->finalized = true;
  return ZB_OK;
}

int zb_dac_decode(zb_dac* d, const int64_t* d_codes, float* d_wav, int B, int T, int impl, void* stream) {
  if (!d || !d_codes || !d_wav || B < 0 || T < 0) return ZB_ERR_ARG;
  if (!d->finalized) return d->err.fail(ZB_ERR_STATE, "zb_dac_finalize first");
  if (B == 0 || T == 0) return ZB_OK;
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(d->err, cudaSetDevice(d->device));
  const int group = impl == 1 ? 1 : (B < d->max_group ? B : (d->max_group < 1 ? 1 : d->max_group));
  if ((long long)T * group > 0x7fffffffLL / 64) return d->err.fail(ZB_ERR_ARG, "T * group too large");
  int rc = impl == 1 ? ensure_workspace_f32(d, T) : ensure_workspace(d, T * group);
  if (rc) return rc;
  for (int b = 0; b < B; b += group) {
    const int nb = B - b < group ? B - b : group;
    const long long* codes = reinterpret_cast<const long long*>(d_codes) + (size_t)b * kNq * T;
    float* wav = d_wav + (size_t)b * 512 * T;
    rc = impl == 1 ? decode_one_simt(d, codes, wav, T, s) : decode_group_tc(d, codes, wav, nb, T, s, impl, d->f16buf, d->keep != 0);
    if (rc) return rc;
  }
  return ZB_OK;
}

// ---- encoder + residual vector quantiser (row f3): wav fp32 [B][1][S], S % 512 == 0 -> codes int64 [B][9][S / 512] --------
namespace {
int conv_down_simt(zb_dac* d, const ConvW& c, const float* x, int T_in, const float* alpha, float* out, cudaStream_t s) {
  ConvSimtParams p;
  memset(&p, 0, sizeof(p));
  p.x = x; p.w = c.w_kio; p.bias = c.bias; p.alpha = alpha; p.out = out;
  p.T_in = T_in; p.T_out = T_in / c.stride; p.Cin = c.Cin; p.Cout = c.Cout; p.K = c.K; p.dil = 1; p.pad = c.pad; p.stride = c.stride;
  const size_t smem = (size_t)(7 * c.stride + c.K) * c.Cin * sizeof(float);
  dim3 grid((p.T_out + 7) / 8, (c.Cout + 127) / 128);
  k_dac_conv_down_simt<<<grid, 128, smem, s>>>(p);
  d->launches++;
  ZB_CUDA(d->err, cudaGetLastError());
  return ZB_OK;
}
const float* rawe(zb_dac* d, const std::string& n) { return d->raw_enc[n].ptr; }
}  // namespace

int zb_dac_encode(zb_dac* d, const float* d_wav, int64_t* d_codes, int B, int64_t S, void* stream) {
  if (!d || B < 0 || S < 0 || (B > 0 && S > 0 && (!d_wav || !d_codes))) return ZB_ERR_ARG;
  if (!d->finalized) return d->err.fail(ZB_ERR_STATE, "zb_dac_finalize first");
  if (!d->enc_ready) return d->err.fail(ZB_ERR_STATE, "zb_dac_encode: the encoder tensors (encoder.*, quantizer.quantizers.*.in_proj.*) were not provided");
  if (S % 512) return d->err.fail(ZB_ERR_ARG, "zb_dac_encode: %lld samples is not a multiple of the hop length 512 (autoencoder.py:17-20 pads)", (long long)S);
  if (B == 0 || S == 0) return ZB_OK;
  if (S / 512 > 0x7fffffffLL / 49152) return d->err.fail(ZB_ERR_ARG, "zb_dac_encode: utterance too long");
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(d->err, cudaSetDevice(d->device));
  const int T = (int)(S / 512);
  int rc = ensure_workspace_f32(d, T);
  if (rc) return rc;
  for (int b = 0; b < B; ++b) {
    float *a = d->f32buf[0], *o = d->f32buf[1], *y = d->f32buf[2];
    int Tl = (int)S;
    if ((rc = conv_simt(d, d->e_conv1, d_wav + (size_t)b * S, Tl, nullptr, nullptr, a, 0, s))) return rc;     // [S][1] -> [S][64]
    for (int j = 0; j < 4; ++j) {
      const std::string bp = "encoder.block." + std::to_string(j);
      for (int u = 0; u < 3; ++u) {
        const std::string rp = bp + ".res_unit" + std::to_string(u + 1);
        if ((rc = conv_simt(d, d->e_ru1[j][u], a, Tl, rawe(d, rp + ".snake1.alpha"), nullptr, y, 0, s))) return rc;
        if ((rc = conv_simt(d, d->e_ru2[j][u], y, Tl, rawe(d, rp + ".snake2.alpha"), a, o, 0, s))) return rc;
        std::swap(a, o);
      }
      if ((rc = conv_down_simt(d, d->e_down[j], a, Tl, rawe(d, bp + ".snake1.alpha"), o, s))) return rc;
      Tl /= kEncStrides[j];
      std::swap(a, o);
    }
    if ((rc = conv_simt(d, d->e_conv2, a, Tl, rawe(d, "encoder.snake1.alpha"), nullptr, o, 0, s))) return rc;   // [T][1024] latents
    k_dac_rvq_encode<<<T, 256, 0, s>>>(o, d->q_ptrs, d->q_ptrs + kNq, d->q_ptrs + 2 * kNq, d->q_ptrs + 3 * kNq, d->q_ptrs + 4 * kNq,
                                       reinterpret_cast<long long*>(d_codes) + (size_t)b * kNq * T, T, kHidden, kNq, kCodes);
    d->launches++;
    ZB_CUDA(d->err, cudaGetLastError());
  }
  return ZB_OK;
}

// ---- allocation-free variant of zb_dac_decode (no counterpart in the reference; see include/zonos_b200.h) ---------
// The five fp16 activation buffers that zb_dac_decode keeps in a cudaMalloc'd, grow-on-demand context workspace live in
// ONE file-scope device array here, sized for ZB_DAC_STATIC_MAX_FRAMES frames per utterance.  Every GPU that loads the
// library holds these 23.6 MB until the process exits.  In host code `g_dac_static_ws` is only a stand-in for the
// symbol: the device address is resolved once per device with cudaGetSymbolAddress (that device current) and passed to
// the kernels like any other pointer.  Nothing has to start at zero: every layer writes all T_l * C_l elements of its
// output before the next layer reads them through a tensor map bounded to exactly those rows.
static __device__ __align__(256) __half g_dac_static_ws[(size_t)5 * 49152 * ZB_DAC_STATIC_MAX_FRAMES];
static __half* g_dac_static_ptr[64] = {nullptr};

int zb_dac_decode_static(zb_dac* d, const int64_t* d_codes, float* d_wav, int B, int T, void* stream) {
  if (!d || B < 0 || T < 0) return ZB_ERR_ARG;
  if (!d->finalized) return d->err.fail(ZB_ERR_STATE, "zb_dac_finalize first");
  if (T > ZB_DAC_STATIC_MAX_FRAMES)
    return d->err.fail(ZB_ERR_ARG, "zb_dac_decode_static: T = %d frames exceeds the static workspace (at most %d frames per utterance); "
                                   "use zb_dac_decode, which allocates", T, ZB_DAC_STATIC_MAX_FRAMES);
  if (d->device < 0 || d->device >= 64) return d->err.fail(ZB_ERR_ARG, "zb_dac_decode_static: device index %d out of range", d->device);
  cudaStream_t s = (cudaStream_t)stream;
  ZB_CUDA(d->err, cudaSetDevice(d->device));
  if (!g_dac_static_ptr[d->device]) {                 // set-up, once per device: symbol -> device address (no allocation)
    void* ptr = nullptr;
    ZB_CUDA(d->err, cudaGetSymbolAddress(&ptr, g_dac_static_ws));
    if (!ptr) return d->err.fail(ZB_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static DAC workspace");
    g_dac_static_ptr[d->device] = static_cast<__half*>(ptr);
  }
  if (B == 0 || T == 0) return ZB_OK;                 // (B = 0 is the explicit set-up call)
  if (!d_codes || !d_wav) return ZB_ERR_ARG;
  __half* bufs[5];
  for (int i = 0; i < 5; ++i) bufs[i] = g_dac_static_ptr[d->device] + (size_t)i * 49152 * ZB_DAC_STATIC_MAX_FRAMES;
  for (int b = 0; b < B; ++b) {
    const long long* codes = reinterpret_cast<const long long*>(d_codes) + (size_t)b * kNq * T;
    int rc = decode_group_tc(d, codes, d_wav + (size_t)b * 512 * T, 1, T, s, 0, bufs, false);
    if (rc) return rc;
  }
  return ZB_OK;
}

int zb_dac_get_intermediate(zb_dac* d, const char* name, float* d_out, int64_t numel, void* stream) {
  if (!d || !name || !d_out) return ZB_ERR_ARG;
  auto it = d->dbg.find(name);
  if (it == d->dbg.end() || !it->second.first) return d->err.fail(ZB_ERR_ARG, "no intermediate '%s' (set keep=1 and decode first)", name);
  if (numel != it->second.second) return d->err.fail(ZB_ERR_ARG, "intermediate '%s' has %lld elements", name, (long long)it->second.second);
  ZB_CUDA(d->err, cudaMemcpyAsync(d_out, it->second.first, (size_t)numel * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return ZB_OK;
}

int zb_dac_set_option(zb_dac* d, const char* key, int value) {
  if (!d || !key) return ZB_ERR_ARG;
  const std::string k(key);
  if (k == "keep_intermediates") { d->keep = value; return ZB_OK; }
  if (k == "ew_multi" || k == "ew_k1") {
    if (value != 8 && value != 16) return d->err.fail(ZB_ERR_ARG, "%s must be 8 or 16", key);
    (k == "ew_multi" ? d->ew_multi : d->ew_k1) = value;
    return ZB_OK;
  }
  if (k == "wide_cout") { d->wide_cout = value; return ZB_OK; }
  if (k == "mt2_min_waves") { d->mt2_min_waves = value; return ZB_OK; }
  if (k == "fuse_ru") { d->fuse_ru = value; return ZB_OK; }
  if (k == "k1_wide") { d->k1_wide = value; return ZB_OK; }
  if (k == "max_group") { d->max_group = value < 1 ? 1 : value; return ZB_OK; }
  return d->err.fail(ZB_ERR_ARG, "unknown option '%s'", key);
}

int64_t zb_dac_launch_count(const zb_dac* d) { return d ? d->launches : 0; }

}  // extern "C"
