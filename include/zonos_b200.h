/*
 * zonos_b200.h — C ABI of libzonos_b200.so, the B200 (sm_100a) hot path behind the reference's
 * `Zonos.generate(conditioning)` / `model.autoencoder.decode(codes)` API
 * (BreakTheBeta/Zonos_Vibes: zonos/model.py:218-315, zonos/autoencoder.py:25-27).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes only.  Every pointer named `d_*` is a DEVICE pointer owned by the
 *     caller (e.g. `tensor.data_ptr()`); the library owns only its context arena (weights, KV pages, scratch).
 *   - Every call returns 0 on success or a negative zb_status; `zb_last_error()` gives the text.  Nothing throws.
 *   - `stream` is a `cudaStream_t` passed as `void*`; all work is asynchronous on that stream unless stated.
 *   - A context is bound to one device and is not re-entrant (mirrors the reference, model.py:39-44).
 *   - There is NO CPU fallback: without a CUDA device every compute entry point fails with ZB_ERR_CUDA.
 *
 * Each entry point cites the reference interface (file:line) it replaces.
 */
#ifndef ZONOS_B200_H
#define ZONOS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ZB_ABI_VERSION 1

typedef enum {
  ZB_OK = 0,
  ZB_ERR_ARG = -1,      /* bad argument / unsupported shape                */
  ZB_ERR_CUDA = -2,     /* CUDA runtime error (text in zb_last_error)      */
  ZB_ERR_STATE = -3,    /* call sequence error (e.g. weights not finalised)*/
  ZB_ERR_NOMEM = -4
} zb_status;

typedef enum { ZB_BF16 = 0, ZB_F32 = 1, ZB_I64 = 2, ZB_I32 = 3 } zb_dtype;

/* Architecture contract = the fields of zonos/config.py:28-62 the hot path reads.  */
typedef struct {
  int32_t d_model;          /* BackboneConfig.d_model                         */
  int32_t n_layer;          /* BackboneConfig.n_layer                         */
  int32_t num_heads;        /* attn_cfg["num_heads"]                          */
  int32_t num_heads_kv;     /* attn_cfg["num_heads_kv"]                       */
  int32_t ffn;              /* attn_mlp_d_intermediate                        */
  int32_t n_codebooks;      /* 9  (autoencoder.py:14)                         */
  int32_t vocab;            /* 1026 logits per head after the pad hook (model.py:46-51) */
  int32_t eos_token_id;     /* 1024 */
  int32_t masked_token_id;  /* 1025 */
  float   norm_eps;         /* BackboneConfig.norm_epsilon                    */
  int32_t rope_table_len;   /* 16384 (backbone/_torch.py:67)                  */
  int32_t reserved[5];
} zb_config;

/* kwargs of sample_from_logits (zonos/sampling.py:117-129). */
typedef struct {
  float   temperature;
  float   top_p;
  int32_t top_k;
  float   min_p;
  float   linear, conf, quad;
  float   repetition_penalty;
  int32_t repetition_penalty_window;
  int32_t reserved[3];
} zb_sampling;

/* Hybrid (Mamba2 + attention) backbone: BackboneConfig.ssm_cfg / attn_layer_idx (zonos/config.py:28-39) and the Mamba2
 * constructor defaults of mamba-ssm 2.2.4 that zonos/backbone/_mamba_ssm.py:15-33 relies on.  Layers listed in
 * attn_layer_idx are attention + GatedMLP(ffn) blocks, every other layer is a Mamba2 block without MLP
 * (d_intermediate = 0).  LayerNorm blocks only (rms_norm = 0). */
typedef struct {
  int32_t d_state;          /* 128 */
  int32_t d_conv;           /* 4   */
  int32_t expand;           /* 2   */
  int32_t headdim;          /* 64  */
  int32_t ngroups;          /* 1   */
  int32_t n_attn_layers;
  int32_t attn_layer_idx[64];
  int32_t reserved[4];
} zb_hybrid;

typedef struct zb_ctx zb_ctx;

/* ---- lifetime ------------------------------------------------------------------------------ */
int  zb_abi_version(void);
int  zb_ctx_create(const zb_config* cfg, int device, zb_ctx** out);
/* MambaSSMZonosBackbone (zonos/backbone/_mamba_ssm.py:8-57).  Extra tensors per Mamba2 layer i (names of mamba-ssm's
 * Mamba2): backbone.layers.{i}.mixer.{in_proj.weight [2*d_inner+2*G*N+H, d], conv1d.weight [conv_dim,1,d_conv],
 * conv1d.bias, norm.weight [d_inner], out_proj.weight [d, d_inner]} bf16 and {dt_bias, A_log, D} [H] fp32. */
int  zb_ctx_create_hybrid(const zb_config* cfg, const zb_hybrid* hyb, int device, zb_ctx** out);
int  zb_ctx_destroy(zb_ctx* ctx);
const char* zb_last_error(const zb_ctx* ctx);   /* ctx may be NULL: last create() error */

/* ---- weights: replaces Zonos.from_local's load_state_dict (model.py:78-88) ------------------
 * `name` is the reference state_dict key: backbone.layers.{i}.norm.weight, ...,
 * embeddings.{k}.weight, heads.{k}.weight (1025 or 1026 rows; padded like utils.py:12-27),
 * and "rope.freqs_cis" = the fp32 table of backbone/_torch.py:9-15 [rope_table_len, Dh/2, 2].
 * The data is copied into the context arena (device-to-device, on `stream`). */
int  zb_set_tensor(zb_ctx* ctx, const char* name, const void* d_ptr, int64_t numel, int dtype, void* stream);
int  zb_weights_finalize(zb_ctx* ctx, void* stream);          /* checks every tensor was provided */

/* ---- KV cache: replaces setup_cache/allocate_inference_cache (model.py:198-202, _torch.py:64-71)
 * rows = 2*B with classifier-free guidance.  Pages of 128 tokens, bf16, head-major inside a page. */
int  zb_kv_alloc(zb_ctx* ctx, int rows, int max_seqlen);
int  zb_kv_reset(zb_ctx* ctx, void* stream);                  /* seqlen_offset = lengths_per_sample = 0 */

/* ---- plugin level: TorchZonosBackbone.forward (backbone/_torch.py:73-80) ---------------------
 * d_hidden_in/out: bf16 [rows, S, d_model].  Appends S tokens to the KV cache at `seqlen_offset` and uses
 * RoPE positions d_lengths[r] + s (int32 [rows]; NULL = seqlen_offset for every row).  Output is post-norm_f. */
int  zb_backbone_forward(zb_ctx* ctx, const void* d_hidden_in, void* d_hidden_out, int rows, int S,
                         int seqlen_offset, const int32_t* d_lengths, void* stream);

/* ---- delay pattern (codebook_pattern.py:5-12), int64 ----------------------------------------- */
int  zb_delay_apply(const int64_t* d_codes, int64_t* d_out, int B, int n_q, int T, int64_t mask_token, void* stream);
int  zb_delay_revert(const int64_t* d_delayed, int64_t* d_out, int B, int n_q, int L, void* stream);

/* ---- sampling (sampling.py:117-182) on given logits -------------------------------------------
 * d_logits fp32 [B, n_q, V]; d_history int64 [B, n_q, hist_len] or NULL (= generated_tokens);
 * d_tokens int64 [B, n_q] out; d_probs_out fp32 [B, n_q, V] or NULL: the filtered, renormalised
 * distribution the draw is made from (for mask parity tests).  Philox4x32-10 keyed by (seed, step). */
int  zb_sample(const float* d_logits, const int64_t* d_history, int hist_len, int64_t* d_tokens,
               float* d_probs_out, int B, int n_q, int V, const zb_sampling* sp,
               uint64_t seed, uint64_t step, void* stream);

/* ---- logits of one forward pass: Zonos._compute_logits (model.py:103-116) ---------------------
 * Teacher-forced step used by parity tests and by generate: embeds column `col` of d_delayed
 * (int64 [B, n_q, L]) for rows b and B+b (model.py:141-142), runs the backbone at the current KV
 * position, applies heads + CFG + pad mask (+ logit_bias when apply_eos_bias, model.py:266-267,280),
 * advances the position, writes fp32 [B, n_q, vocab] to d_logits_out. */
int  zb_decode_logits(zb_ctx* ctx, const int64_t* d_delayed, int B, int L, int col, float cfg_scale,
                      int apply_eos_bias, float* d_logits_out, void* stream);

/* ---- generation: Zonos.generate (model.py:218-315) --------------------------------------------
 * zb_gen_begin: d_prefix bf16 [2B, Lc, d_model] (rows [0,B) cond, [B,2B) uncond, model.py:204-212);
 *   d_delayed int64 [B, n_q, L] already delay-patterned with -1 = unknown (model.py:246-251);
 *   prefix_cols = P+1 known leading columns.  Runs the prefill (model.py:253-264), samples without
 *   logit_bias/repetition penalty (model.py:256) and writes column P+1.
 * zb_gen_steps: runs up to n_steps iterations of the loop model.py:276-307 entirely on the device
 *   (EOS wind-down and frame write-back included; no host sync inside).  compat_scatter=1 reproduces
 *   masked_scatter_'s sequential routing, 0 = positional.
 * zb_gen_poll: returns the loop state after the OLDEST zb_gen_steps call that has not been polled yet (it waits for that
 *   call only: up to 4 calls may be in flight, so the host can enqueue call k+1 before it waits for call k); with no call
 *   in flight it synchronises `stream` and reads the current state. */
int  zb_gen_begin(zb_ctx* ctx, const void* d_prefix, int B, int Lc, int64_t* d_delayed, int L, int prefix_cols,
                  float cfg_scale, const zb_sampling* sp, uint64_t seed, int compat_scatter, void* stream);
int  zb_gen_steps(zb_ctx* ctx, int n_steps, void* stream);
int  zb_gen_poll(zb_ctx* ctx, int* done, int* col, int32_t* h_out_len /* [B] or NULL */, void* stream);

/* One application of the device-side frame write-back + EOS wind-down (model.py:283-301) on caller-owned state:
 * d_tokens int64 [B, n_q] = the sampled next_token, d_remaining / d_stopping / d_out_len int32 [B]; writes column
 * col+1 of d_delayed.  Exposed for exact-integer parity tests of the loop state machine. */
int  zb_update_frame(zb_ctx* ctx, int64_t* d_delayed, int B, int L, int col, const int64_t* d_tokens, int32_t* d_remaining,
                     int32_t* d_stopping, int32_t* d_out_len, int eos_logic, int compat_scatter, void* stream);

/* ---- DAC 44 kHz decoder: DACAutoencoder.decode (autoencoder.py:25-27) --------------------------
 * Weights use the HF DacModel keys, fp32 on the device; repacked at load. */
typedef struct zb_dac zb_dac;
int  zb_dac_create(int device, zb_dac** out);
int  zb_dac_destroy(zb_dac* dac);
const char* zb_dac_last_error(const zb_dac* dac);
int  zb_dac_set_tensor(zb_dac* dac, const char* name, const float* d_ptr, int64_t numel, void* stream);
int  zb_dac_finalize(zb_dac* dac, void* stream);
/* d_codes int64 [B, 9, T] in [0,1023]; d_wav fp32 [B, 1, 512*T].  impl: 0 = tcgen05 fp16 (default),
 * 1 = SIMT fp32 cross-check, 2 / 3 = force the one-tile-per-CTA / the persistent tcgen05 kernel for every layer. */
int  zb_dac_decode(zb_dac* dac, const int64_t* d_codes, float* d_wav, int B, int T, int impl, void* stream);
/* Allocation-free variant of zb_dac_decode (tcgen05 path).  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid it).  The activation scratch that zb_dac_decode keeps in a grow-on-demand cudaMalloc'd
 * workspace lives in a file-scope `static __device__` array sized for ZB_DAC_STATIC_MAX_FRAMES frames per utterance
 * (5 x 49152 x 48 fp16 = 23.6 MB, held by every GPU that loads the library until the process exits).  T above the
 * maximum is rejected with ZB_ERR_ARG and a message naming the limit; B is unlimited (utterances are decoded one after
 * the other).  The call makes no cudaMalloc / cudaFree / allocating library call; the first call per device resolves the
 * array's device address (cudaGetSymbolAddress) - call once with B = 0 outside any capture to do that up front.
 * All calls share the one array, which keeps the last call's contents: nothing in it has to start at zero (each
 * layer fully writes what the next reads), but two calls that overlap on the GPU (different streams, or different
 * zb_dac contexts on the same device) RACE on it - serialise them.  Same inputs / outputs as zb_dac_decode. */
#define ZB_DAC_STATIC_MAX_FRAMES 48
int  zb_dac_decode_static(zb_dac* dac, const int64_t* d_codes, float* d_wav, int B, int T, void* stream);
/* SURVEY §8 row f3 - DACAutoencoder.encode (zonos/autoencoder.py:22-23 -> HF DacModel.encode: DacEncoder + DacResidualVectorQuantizer,
 * modeling_dac.py): d_wav fp32 [B][1][S] at 44.1 kHz, S a multiple of 512 (the reference's preprocess pads, autoencoder.py:17-20)
 * -> d_codes int64 [B][9][S / 512].  fp32 arithmetic like the reference (no autocast on encode).  Needs the encoder tensors
 * ("encoder.*" and "quantizer.quantizers.K.in_proj.*", set with zb_dac_set_tensor before zb_dac_finalize); a context that was
 * given the decoder tensors only returns ZB_ERR_STATE here and decodes as before. */
int  zb_dac_encode(zb_dac* dac, const float* d_wav, int64_t* d_codes, int B, int64_t S, void* stream);
/* debug/parity: copy an intermediate of the last decode ("latent", "conv1", "block0".."block3") as fp32
 * [B, C, T_l] (reference layout) into d_out; returns ZB_ERR_ARG if the name is unknown. */
int  zb_dac_get_intermediate(zb_dac* dac, const char* name, float* d_out, int64_t numel, void* stream);
int  zb_dac_set_option(zb_dac* dac, const char* key, int value);      /* "keep_intermediates" */
int64_t zb_dac_launch_count(const zb_dac* dac);

/* ---- serving shell (server.py:142-147) -----------------------------------------------------------------
 * fp32 waveform -> 16-bit PCM exactly as the reference's /tts handler does it on the host with numpy:
 * clip to [-1, 1], multiply by 32767, convert to int16 with truncation toward zero.  NaN -> 0.  n samples, device pointers. */
int  zb_wav_to_pcm16(const float* d_wav, int16_t* d_pcm, int64_t n, void* stream);

/* ---- speaker-embedding network (zonos/speaker_cloning.py) ------------------------------------------------
 * Op-level entry points on caller-owned fp32 device tensors (NCHW, contiguous); zonos_vibes_b200/speaker_cloning.py composes
 * them in the reference's order (ResNet293_based.forward, :217-224).  BatchNorm is eval-mode and arrives folded:
 * scale = weight / sqrt(running_var + eps), shift = bias - running_mean * scale.
 *   zb_spk_conv2d          nn.Conv2d(bias=False) 3x3 pad 1 (stride 1 | 2) or 1x1 pad 0 (the downsample branch) + folded
 *                          BatchNorm2d + optional ReLU (:69-81, :169-170): x [B, Cin, H, W], w [Cout, Cin, k, k] -> [B, Cout, Ho, Wo]
 *   zb_spk_simam_add_relu  SimAMBasicBlock tail (:86-97): out = relu(x * sigmoid(d / (4 (v + lambda)) + 0.5) + shortcut) per (b, c) plane
 *   zb_spk_asp             attentive statistics pooling (:38-61): x [B, C, T]; attention Conv1d(C, hidden, 1) -> ReLU -> BatchNorm1d
 *                          (folded) -> Conv1d(hidden, C, 1) -> softmax over T; out [B, 2 C] = [mean | std]; d_hidden = scratch [B, hidden, T]
 *   zb_spk_linear          nn.Linear (the bottleneck :212 and the LDA projection :400-402): x [B, in] -> [B, out], bias may be NULL */
int  zb_spk_conv2d(const float* d_x, const float* d_w, const float* d_scale, const float* d_shift, float* d_out, int B, int Cin, int H, int W,
                   int Cout, int ksize, int stride, int relu, void* stream);
int  zb_spk_simam_add_relu(const float* d_x, const float* d_shortcut, float* d_out, int B, int C, int H, int W, float lambda_p, void* stream);
int  zb_spk_asp(const float* d_x, const float* d_w1, const float* d_b1, const float* d_bn_scale, const float* d_bn_shift, const float* d_w2,
                const float* d_b2, float* d_hidden, float* d_out, int B, int C, int T, int hidden, void* stream);
int  zb_spk_linear(const float* d_x, const float* d_w, const float* d_bias, float* d_out, int B, int in_features, int out_features, void* stream);

/* ---- PrefixConditioner (zonos/conditioning.py) -------------------------------------------------------
 * The four ops the reference's PrefixConditioner.forward (conditioning.py:297-313) is made of, on caller-owned bf16 device
 * tensors; zonos_vibes_b200/conditioning.py composes them in the reference's order.  Text -> phoneme ids stays on the host.
 *   zb_pc_fourier    FourierConditioner.apply_cond (conditioning.py:257-261): d_x fp32 [n, in_dim], weight bf16 [half, in_dim]
 *                    -> bf16 [n, 2*half] = [cos | sin] of 2*pi*((x - min)/(max - min)) @ W^T
 *   zb_pc_embed      nn.Embedding of EspeakPhonemeConditioner / IntegerConditioner (:230-236, :271-273): ids int64 [n]
 *                    (id_offset = min_val is subtracted), table bf16 [table_rows, d] -> bf16 [n, d]
 *   zb_pc_linear     nn.Linear (+ SiLU when silu != 0) of Conditioner.project (:24-35): in bf16 [n, K], w bf16 [N, K],
 *                    b bf16 [N] or NULL -> bf16 [n, N]
 *   zb_pc_layernorm  PrefixConditioner.norm (:301,313): bf16 [n, d] -> bf16 [n, d] */
int  zb_pc_fourier(const float* d_x, const void* d_weight, float min_val, float max_val, void* d_out, int n, int in_dim, int half, void* stream);
int  zb_pc_embed(const int64_t* d_ids, const void* d_table, void* d_out, int n, int d, int64_t id_offset, int64_t table_rows, void* stream);
int  zb_pc_linear(const void* d_in, const void* d_w, const void* d_b, void* d_out, int n, int K, int N, int silu, void* stream);
int  zb_pc_layernorm(const void* d_in, const void* d_w, const void* d_b, float eps, void* d_out, int n, int d, void* stream);

/* ---- introspection used by bench.py (launch counting, roofline byte counts) ------------------- */
int64_t zb_weight_bytes(const zb_ctx* ctx);            /* algorithmic weight bytes streamed per decode step */
int64_t zb_launch_count(const zb_ctx* ctx);            /* kernels launched by this context so far (graph nodes count per replay) */
/* Tuning / cross-check switches (defaults in parentheses; every combination is parity-tested in tests/test_gpu_parity.py):
 *   which path:   "mega" (1) persistent kernel for <= 8 rows; "bt_min_rows" (9) batched tcgen05 path from this many rows;
 *                 "graph" (1) one CUDA graph per step on the per-op / batched paths; "gemv" (1) mma.sync GEMVs on the per-op path;
 *                 "bt_prefill" (1) prefill as blocks of rows on the batched path; "chunked_prefill" (1) hybrid chunked SSD scan
 *   persistent:   "ll" (3) which activations travel as epoch-tagged words (0 = grid barriers ... 3 = all); "chain" (0);
 *                 "att_chunk", "att_single", "ring", "psleep", "evict", "l2_prefetch", "pdl" (1)
 *   batched:      "bt_cta2" (1) bit 0: split-K GEMMs on CTA pairs (k_bt_gemm2), bit 1: fc1 too; "bt_cluster" (1) 2 / 4 = activation
 *                 tile multicast across a cluster (k_bt_gemm); "attn_tma" (1) 0 = ld.global attention, 1 = TMA-staged K/V from 64 rows,
 *                 2 = always, 3 = always with early slot release; "attn_mma" (1) 0 = SIMT attention; "nsplit" (0 = auto) KV splits
 *   debug:        "trace", "trace_quiet" (see zb_last_trace)
 * Unknown keys return ZB_ERR_ARG. */
int  zb_set_option(zb_ctx* ctx, const char* key, int value);
/* isolated device timing of one kernel family ("in_proj", "attn", "out_proj", "fc1", "fc2"), `iters` launches cycling
 * over the layers, CUDA events on `stream`; total ms in *ms_out (synchronises).  Scratch outputs only. */
int  zb_bench_kernel(zb_ctx* ctx, const char* name, int iters, float* ms_out, void* stream);
/* device-timed (CUDA events on `stream`) duration of the zb_gen_steps call that the last zb_gen_poll returned, ms; <0 if none */
float zb_last_steps_ms(zb_ctx* ctx);
/* bench.py's per-phase table: with options "trace" = 1 and "trace_quiet" = 1 a persistent-kernel launch stamps
 * %globaltimer (ns) at every phase boundary of one step on CTA 0; this copies those stamps (order documented in
 * profiles/r02_summary.md) and returns how many there are.  Traced launches use a separate, slower instantiation. */
int zb_last_trace(zb_ctx* ctx, uint64_t* out, int max_slots);

#ifdef __cplusplus
}
#endif
#endif /* ZONOS_B200_H */
