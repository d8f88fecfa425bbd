"""CPU oracle for the Zonos generation hot path (transformer backbone, heads, sampling, delay pattern).

TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import this module.  The product path
(`zonos_vibes_b200/`) never imports it and has no CPU fallback.

What it is: a functional restatement, in plain torch CPU ops, of the reference's algorithm
(BreakTheBeta/Zonos_Vibes; file:line cited per function, relative to the reference root).  The
reference is itself pure PyTorch, so the restatement uses the same ATen primitives at the same
precision (bf16 storage, fp32 accumulate inside ATen) and is therefore expected to be *bit-identical*
to the reference's CPU path; `tests/test_oracle_pinned.py` checks exactly that against golden
vectors frozen from the real reference by `oracle/make_golden.py` (and live against
`/root/reference` when it is present).  Parity status: PINNED by those self-generated vectors — the
reference ships no golden vectors of its own.

It is written against an explicit `state_dict` (reference key names) instead of
`nn.Module`s so that identical weights can be fed to the CUDA path and to the oracle.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch
import torch.nn.functional as F

N_CODEBOOKS = 9          # autoencoder.num_codebooks, zonos/autoencoder.py:14
EMB_ROWS = 1026          # zonos/model.py:36
HEAD_ROWS = 1025         # zonos/model.py:37 (padded to 1026 by the load hook, model.py:46-51)


# --------------------------------------------------------------------------------------
# config / synthetic weights
# --------------------------------------------------------------------------------------
@dataclass
class OracleConfig:
    """Mirror of the fields of zonos/config.py:28-62 that the transformer hot path reads."""
    d_model: int = 2048
    n_layer: int = 26
    num_heads: int = 16
    num_heads_kv: int = 4
    attn_mlp_d_intermediate: int = 8192
    norm_epsilon: float = 1e-5
    eos_token_id: int = 1024
    masked_token_id: int = 1025
    pad_vocab_to_multiple_of: int = 8

    @property
    def head_dim(self) -> int:
        return self.d_model // self.num_heads

    def to_reference_dict(self) -> dict:
        """The JSON dict `ZonosConfig.from_dict` (zonos/config.py:56-62) takes."""
        return dict(
            backbone=dict(d_model=self.d_model, d_intermediate=0,
                          attn_mlp_d_intermediate=self.attn_mlp_d_intermediate, n_layer=self.n_layer,
                          ssm_cfg={}, attn_layer_idx=[],
                          attn_cfg=dict(num_heads=self.num_heads, num_heads_kv=self.num_heads_kv),
                          rms_norm=False, residual_in_fp32=False, norm_epsilon=self.norm_epsilon),
            prefix_conditioner=dict(conditioners=[], projection="none"),
            eos_token_id=self.eos_token_id, masked_token_id=self.masked_token_id,
            pad_vocab_to_multiple_of=self.pad_vocab_to_multiple_of,
        )


FULL_CONFIG = OracleConfig()                                   # Zonos-v0.1-transformer
TINY_CONFIG = OracleConfig(d_model=256, n_layer=2, num_heads=2, num_heads_kv=1,
                           attn_mlp_d_intermediate=512)        # Dh=128 like the real model
SMALL_CONFIG = OracleConfig(d_model=512, n_layer=3, num_heads=4, num_heads_kv=2,
                            attn_mlp_d_intermediate=1024)


def find_multiple(n: int, k: int) -> int:
    """zonos/utils.py:6-9."""
    return n if (k == 0 or n % k == 0) else n + k - (n % k)


def synth_state_dict(cfg: OracleConfig, seed: int = 0, dtype=torch.bfloat16) -> dict:
    """Deterministic random-init weights with the reference's state_dict keys/shapes.

    Linear weights ~ U(-1/sqrt(in), 1/sqrt(in)) (nn.Linear default scale), embeddings ~ N(0,1)
    (nn.Embedding default), LayerNorm weight 1+0.1 N(0,1), bias 0.1 N(0,1) so that affine terms are
    exercised.  Generated on CPU with a private generator: identical on every box with the same torch.
    """
    g = torch.Generator(device="cpu").manual_seed(seed)
    d, f = cfg.d_model, cfg.attn_mlp_d_intermediate
    qkv = (cfg.num_heads + 2 * cfg.num_heads_kv) * cfg.head_dim

    def lin(out_f, in_f):
        b = 1.0 / math.sqrt(in_f)
        return ((torch.rand(out_f, in_f, generator=g) * 2 - 1) * b).to(dtype)

    def ln(prefix, sd):
        sd[prefix + ".weight"] = (1.0 + 0.1 * torch.randn(d, generator=g)).to(dtype)
        sd[prefix + ".bias"] = (0.1 * torch.randn(d, generator=g)).to(dtype)

    sd: dict = {}
    for i in range(cfg.n_layer):
        p = f"backbone.layers.{i}"
        ln(p + ".norm", sd)
        sd[p + ".mixer.in_proj.weight"] = lin(qkv, d)
        sd[p + ".mixer.out_proj.weight"] = lin(d, cfg.num_heads * cfg.head_dim)
        ln(p + ".norm2", sd)
        sd[p + ".mlp.fc1.weight"] = lin(2 * f, d)
        sd[p + ".mlp.fc2.weight"] = lin(d, f)
    ln("backbone.norm_f", sd)
    for k in range(N_CODEBOOKS):
        sd[f"embeddings.{k}.weight"] = torch.randn(EMB_ROWS, d, generator=g).to(dtype)
    for k in range(N_CODEBOOKS):
        sd[f"heads.{k}.weight"] = lin(HEAD_ROWS, d)
    return sd


def synth_prefix_conditioning(cfg: OracleConfig, batch: int = 1, lc: int = 24, seed: int = 1) -> torch.Tensor:
    """Synthetic `prefix_conditioning` [2B, Lc, d] bf16 (rows [0,B) cond, [B,2B) uncond; model.py:204-212).
    PrefixConditioner ends in a LayerNorm (conditioning.py:297-313) so N(0,1) entries are the right scale."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    return torch.randn(2 * batch, lc, cfg.d_model, generator=g).to(torch.bfloat16)


# --------------------------------------------------------------------------------------
# delay pattern  (zonos/codebook_pattern.py)
# --------------------------------------------------------------------------------------
def apply_delay_pattern(codes: torch.Tensor, mask_token: int) -> torch.Tensor:
    """codebook_pattern.py:5-7 — right-pad by n_q with `mask_token`, then circularly shift row k by k+1.
    out[b,k,t] = padded[b,k,(t-(k+1)) mod (T+n_q)]."""
    b, nq, t = codes.shape
    length = t + nq
    padded = torch.full((b, nq, length), mask_token, dtype=codes.dtype)
    padded[..., :t] = codes
    idx = (torch.arange(length)[None, :] - (torch.arange(nq)[:, None] + 1)) % length      # [nq, L]
    return torch.gather(padded, 2, idx[None].expand(b, -1, -1))


def revert_delay_pattern(codes: torch.Tensor) -> torch.Tensor:
    """codebook_pattern.py:10-12 — out[b,k,i] = codes[b,k,i+k+1], i in [0, L-n_q)."""
    b, nq, length = codes.shape
    t = length - nq
    idx = torch.arange(t)[None, :] + torch.arange(nq)[:, None] + 1
    return torch.gather(codes, 2, idx[None].expand(b, -1, -1))


# --------------------------------------------------------------------------------------
# sampling  (zonos/sampling.py)
# --------------------------------------------------------------------------------------
def repetition_penalty(logits, generated_tokens, penalty: float, window: int):
    """sampling.py:99-114 — factor = penalty ** (#occurrences in the last `window` columns),
    ids clamped to vocab-1; logits<=0 are multiplied, >0 divided."""
    recent = generated_tokens[..., -window:].clamp_max(logits.shape[-1] - 1).to(torch.int64)
    factors = torch.ones_like(logits).scatter_reduce(2, recent, torch.full_like(logits, penalty), reduce="prod")
    return torch.where(logits <= 0, logits * factors, logits / factors)


def unified_probs(probs, linear: float, conf: float, quad: float):
    """sampling.py:29-43 — NovelAI unified sampler."""
    logp = torch.log(probs.clamp_min(1e-20))
    entropy = -(probs * logp).sum(-1, keepdim=True)
    return (logp * (linear + entropy * conf) - logp ** 2 * quad).softmax(-1)


def top_k_probs(probs, k: int):
    """sampling.py:45-61 — keep p >= k-th largest (ties kept), renormalise."""
    kth = torch.topk(probs, min(k, probs.size(-1))).values[..., -1:]
    probs = torch.where(probs < kth, 0.0, probs)
    return probs / probs.sum(-1, keepdim=True)


def top_p_probs(probs, p: float):
    """sampling.py:64-79 — descending sort, zero entries whose EXCLUSIVE cumsum > p, renormalise."""
    srt, idx = torch.sort(probs, dim=-1, descending=True)
    drop = torch.cumsum(srt, dim=-1) - srt > p
    srt = srt * (~drop).float()
    probs = probs.scatter(-1, idx, srt)
    return probs / probs.sum(-1, keepdim=True)


def min_p_probs(probs, min_p: float):
    """sampling.py:82-96 — zero p < min_p * max(p), renormalise."""
    top = probs.max(-1, keepdim=True).values
    probs = probs.masked_fill(probs < min_p * top, 0.0)
    return probs / probs.sum(-1, keepdim=True)


def filtered_probs(logits, temperature=1.0, top_p=0.0, top_k=0, min_p=0.0, linear=0.0, conf=0.0, quad=0.0,
                   generated_tokens=None, repetition_penalty_=3.0, repetition_penalty_window=2):
    """The deterministic part of sampling.py:117-182 (everything before the multinomial draw).
    Returns (logits_after_penalty, probs or None when temperature == 0)."""
    if repetition_penalty_ != 1.0 and generated_tokens is not None:
        logits = repetition_penalty(logits, generated_tokens, repetition_penalty_, repetition_penalty_window)
    if temperature <= 0:
        return logits, None
    probs = torch.softmax(logits / temperature, dim=-1)
    if linear > 0.0:
        probs = unified_probs(probs, linear, conf, quad)
    if top_p > 0:
        probs = top_p_probs(probs, top_p)
    if top_k > 0:
        probs = top_k_probs(probs, top_k)
    if min_p > 0:
        probs = min_p_probs(probs, min_p)
    return logits, probs


def sample_from_logits(logits, temperature=1.0, top_p=0.0, top_k=0, min_p=0.0, linear=0.0, conf=0.0, quad=0.0,
                       generated_tokens=None, repetition_penalty=3.0, repetition_penalty_window=2,
                       generator=None):
    """sampling.py:117-182; the draw is the exponential race of sampling.py:19-21: argmax(p / Exp(1))."""
    logits, probs = filtered_probs(logits, temperature, top_p, top_k, min_p, linear, conf, quad,
                                   generated_tokens, repetition_penalty, repetition_penalty_window)
    if probs is None:
        return torch.argmax(logits, dim=-1, keepdim=True)
    q = torch.empty_like(probs).exponential_(1, generator=generator)
    return torch.argmax(probs / q, dim=-1, keepdim=True).to(torch.int64)


# --------------------------------------------------------------------------------------
# transformer backbone  (zonos/backbone/_torch.py)
# --------------------------------------------------------------------------------------
def precompute_freqs_cis(seq_len: int, n_elem: int, base: float = 10000.0) -> torch.Tensor:
    """_torch.py:9-15 — [seq_len, n_elem/2, 2] fp32 (cos, sin) of pos * base^(-2i/n_elem)."""
    inv = 1.0 / (base ** (torch.arange(0, n_elem, 2)[: n_elem // 2].float() / n_elem))
    ang = torch.outer(torch.arange(seq_len), inv)
    cis = torch.polar(torch.ones_like(ang), ang)
    return torch.stack([cis.real, cis.imag], dim=-1)


def apply_rotary_emb(x: torch.Tensor, freqs: torch.Tensor) -> torch.Tensor:
    """_torch.py:18-30 — adjacent-pair rotation in fp32, cast back.  x [B,S,H,Dh]; freqs [B,S,Dh/2,2]."""
    xs = x.float().reshape(*x.shape[:-1], -1, 2)
    fr = freqs.view(-1, xs.size(1), 1, xs.size(3), 2)
    x0, x1, c, s = xs[..., 0], xs[..., 1], fr[..., 0], fr[..., 1]
    out = torch.stack([x0 * c - x1 * s, x1 * c + x0 * s], dim=-1)
    return out.flatten(3).type_as(x)


@dataclass
class OracleInferenceParams:
    """zonos/config.py:7-25."""
    max_seqlen: int
    max_batch_size: int
    seqlen_offset: int = 0
    key_value_memory_dict: dict = field(default_factory=dict)
    lengths_per_sample: torch.Tensor | None = None


class OracleZonos:
    """Functional restatement of `Zonos` (zonos/model.py) + `TorchZonosBackbone` (zonos/backbone/_torch.py)."""

    def __init__(self, cfg: OracleConfig, state_dict: dict):
        self.cfg = cfg
        self.sd = dict(state_dict)
        # model.py:46-51 + utils.py:12-27: heads get `rows % multiple` zero rows appended (1025 -> 1026);
        # embeddings are (mis)checked on dim 1 and left alone.
        m = cfg.pad_vocab_to_multiple_of
        if m:
            for k in range(N_CODEBOOKS):
                w = self.sd[f"heads.{k}.weight"]
                if w.shape[0] % m != 0:
                    self.sd[f"heads.{k}.weight"] = F.pad(w, (0, 0, 0, w.shape[0] % m))
        self.freqs_cis = None

    # ---- model.py:198-202 / _torch.py:64-71,96-97
    def setup_cache(self, batch_size: int, max_seqlen: int, dtype=torch.bfloat16) -> OracleInferenceParams:
        c = self.cfg
        max_seqlen = find_multiple(max_seqlen, 8)
        self.freqs_cis = precompute_freqs_cis(16384, c.head_dim)
        kv = {i: torch.zeros(batch_size, max_seqlen, 2, c.num_heads_kv, c.head_dim, dtype=dtype)
              for i in range(c.n_layer)}
        return OracleInferenceParams(max_seqlen, batch_size, 0, kv,
                                     torch.zeros(batch_size, dtype=torch.int32))

    # ---- _torch.py:105-141
    def _attention(self, i: int, x, params: OracleInferenceParams, freqs):
        c, sd = self.cfg, self.sd
        b, s, _ = x.shape
        qs, ks = c.num_heads * c.head_dim, c.num_heads_kv * c.head_dim
        q, k, v = F.linear(x, sd[f"backbone.layers.{i}.mixer.in_proj.weight"]).split([qs, ks, ks], dim=-1)
        q = apply_rotary_emb(q.view(b, s, c.num_heads, c.head_dim), freqs)
        k = apply_rotary_emb(k.view(b, s, c.num_heads_kv, c.head_dim), freqs)
        v = v.view(b, s, c.num_heads_kv, c.head_dim)
        cache = params.key_value_memory_dict[i]                       # _torch.py:33-49
        start, end = params.seqlen_offset, params.seqlen_offset + s
        cache[:b, start:end, 0] = k
        cache[:b, start:end, 1] = v
        kk, vv = cache[:b, :end].unbind(dim=-3)
        y = F.scaled_dot_product_attention(q.transpose(1, 2), kk.transpose(1, 2), vv.transpose(1, 2),
                                           is_causal=s > 1, enable_gqa=True)
        y = y.transpose(1, 2).contiguous().view(b, s, qs)
        return F.linear(y, sd[f"backbone.layers.{i}.mixer.out_proj.weight"])

    # ---- _torch.py:144-152
    def _mlp(self, i: int, x):
        sd = self.sd
        y, gate = F.linear(x, sd[f"backbone.layers.{i}.mlp.fc1.weight"]).chunk(2, dim=-1)
        return F.linear(y * F.silu(gate), sd[f"backbone.layers.{i}.mlp.fc2.weight"])

    def _ln(self, prefix: str, x):
        return F.layer_norm(x, (self.cfg.d_model,), self.sd[prefix + ".weight"], self.sd[prefix + ".bias"],
                            self.cfg.norm_epsilon)

    # ---- _torch.py:73-80, 99-102
    def backbone(self, hidden, params: OracleInferenceParams, collect: dict | None = None):
        pos = torch.arange(hidden.shape[1])[None, :] + params.lengths_per_sample[:, None].long()
        freqs = self.freqs_cis[pos].expand(hidden.shape[0], -1, -1, -1)
        x = hidden
        for i in range(self.cfg.n_layer):
            p = f"backbone.layers.{i}"
            x = x + self._attention(i, self._ln(p + ".norm", x), params, freqs)
            x = x + self._mlp(i, self._ln(p + ".norm2", x))
            if collect is not None:
                collect[i] = x.clone()
        return self._ln("backbone.norm_f", x)

    # ---- model.py:97-98 (python sum: ((0+E0)+E1)+... in bf16)
    def embed_codes(self, codes):
        out = 0
        for k in range(N_CODEBOOKS):
            out = out + F.embedding(codes[:, k], self.sd[f"embeddings.{k}.weight"])
        return out

    # ---- model.py:100-116
    def compute_logits(self, hidden, params, cfg_scale: float):
        last = self.backbone(hidden, params)[:, -1, :].unsqueeze(1)
        logits = torch.stack([F.linear(last, self.sd[f"heads.{k}.weight"]) for k in range(N_CODEBOOKS)], dim=1)
        logits = logits.squeeze(2).float()
        if cfg_scale != 1.0:
            cond, uncond = logits.chunk(2)
            logits = uncond + (cond - uncond) * cfg_scale
        logits[..., 1025:].fill_(-torch.inf)
        return logits

    # ---- model.py:141-143 (non-graph path)
    def decode_one_token(self, input_ids, params, cfg_scale: float):
        h = self.embed_codes(input_ids).repeat(2, 1, 1)
        return self.compute_logits(h, params, cfg_scale)

    # ---- model.py:181-196  (B=1 only: the reference's expand() raises for B>1)
    def prefill(self, prefix_hidden, input_ids, params, cfg_scale: float):
        ids = input_ids.expand(prefix_hidden.shape[0], -1, -1)
        h = torch.cat([prefix_hidden, self.embed_codes(ids)], dim=1)
        return self.compute_logits(h, params, cfg_scale)

    # ---- model.py:218-315
    @torch.inference_mode()
    def generate(self, prefix_conditioning, audio_prefix_codes=None, max_new_tokens: int = 86 * 30,
                 cfg_scale: float = 2.0, sampling_params: dict | None = None, generator=None,
                 callback=None, trace: dict | None = None):
        """Batch-1 generation.  `trace`, when given, receives per-step logits (after logit_bias) and the
        written frames — used to freeze golden vectors and for teacher-forced parity."""
        assert cfg_scale != 1
        assert prefix_conditioning.shape[0] == 2, "oracle = batch-1 runs (SURVEY §8c)"
        c = self.cfg
        sp = dict(min_p=0.1) if sampling_params is None else dict(sampling_params)
        plen = 0 if audio_prefix_codes is None else audio_prefix_codes.shape[2]
        t_audio = plen + max_new_tokens
        seq_len = prefix_conditioning.shape[1] + t_audio + 9
        params = self.setup_cache(2, seq_len)
        codes = torch.full((1, 9, t_audio), -1, dtype=torch.int64)
        if audio_prefix_codes is not None:
            codes[..., :plen] = audio_prefix_codes
        delayed = apply_delay_pattern(codes, c.masked_token_id)

        logits = self.prefill(prefix_conditioning, delayed[..., : plen + 1], params, cfg_scale)
        if trace is not None:
            trace.setdefault("logits", []).append(logits.clone())
        nxt = sample_from_logits(logits, generator=generator, **sp)
        offset = plen + 1
        frame = delayed[..., offset: offset + 1]
        frame.masked_scatter_(frame == -1, nxt)
        prefix_length = prefix_conditioning.shape[1] + plen + 1
        params.seqlen_offset += prefix_length
        params.lengths_per_sample[:] += prefix_length

        bias = torch.zeros_like(logits)
        bias[:, 1:, c.eos_token_id] = -torch.inf
        stopping = False
        max_steps = delayed.shape[2] - offset
        remaining = max_steps
        step = 0
        while remaining > 0:
            offset += 1
            logits = self.decode_one_token(delayed[..., offset - 1: offset], params, cfg_scale) + bias
            if trace is not None:
                trace["logits"].append(logits.clone())
            nxt = sample_from_logits(logits, generated_tokens=delayed[..., :offset], generator=generator, **sp)
            if int(nxt[0, 0, 0]) == c.eos_token_id:
                remaining = min(remaining, 9)
                stopping = True
            if stopping:
                idx = min(9 - remaining, 8)
                nxt[0, :idx] = c.masked_token_id
                nxt[0, idx] = c.eos_token_id
            frame = delayed[..., offset: offset + 1]
            frame.masked_scatter_(frame == -1, nxt)
            params.seqlen_offset += 1
            params.lengths_per_sample[:] += 1
            remaining -= 1
            step += 1
            if callback is not None and not callback(frame, step, max_steps):
                break
        if trace is not None:
            trace["delayed"] = delayed.clone()
            trace["offset"] = offset
        out = revert_delay_pattern(delayed)
        out.masked_fill_(out >= 1024, 0)
        return out[..., : offset - 9]

    @torch.inference_mode()
    def teacher_forced_logits(self, prefix_conditioning, delayed: torch.Tensor, n_steps: int, cfg_scale: float = 2.0):
        """Per-step logits (prefill + n_steps decode steps, logit_bias applied on decode steps) when the
        model is fed the given, fully known `delayed` codes [1,9,L] — the parity fixture of SURVEY §8c."""
        c = self.cfg
        params = self.setup_cache(2, prefix_conditioning.shape[1] + delayed.shape[2] + 9)
        out = [self.prefill(prefix_conditioning, delayed[..., :1], params, cfg_scale)]
        plen = prefix_conditioning.shape[1] + 1
        params.seqlen_offset += plen
        params.lengths_per_sample[:] += plen
        bias = torch.zeros_like(out[0])
        bias[:, 1:, c.eos_token_id] = -torch.inf
        for s in range(n_steps):
            out.append(self.decode_one_token(delayed[..., s + 1: s + 2], params, cfg_scale) + bias)
            params.seqlen_offset += 1
            params.lengths_per_sample[:] += 1
        return torch.cat(out, dim=0)          # [1+n_steps, 9, 1026]
