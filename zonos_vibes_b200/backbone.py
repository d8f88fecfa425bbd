"""B200 backbone behind the reference's backbone plugin API (zonos/backbone/__init__.py:1-12, _torch.py:52-80).

`BACKBONES["b200"] = B200ZonosBackbone` can be handed to the reference's own `Zonos(config, backbone_cls)`; the class
keeps the reference's parameter names (`layers.{i}.norm/mixer/norm2/mlp`, `norm_f`) so the reference's
`load_state_dict` fills it, and implements `allocate_inference_cache` / `forward` on top of the native context.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from . import _lib
from .config import BackboneConfig, InferenceParams

ROPE_TABLE_LEN = 16384       # zonos/backbone/_torch.py:67
N_CODEBOOKS = 9              # zonos/autoencoder.py:14
VOCAB = 1026                 # head width after the pad hook (zonos/model.py:46-51, utils.py:12-27)


def precompute_freqs_cis(seq_len: int, n_elem: int, base: float = 10000.0) -> torch.Tensor:
    """RoPE table [seq_len, n_elem/2, 2] fp32 = (cos, sin)(pos * base^(-2i/n_elem)); host-side setup data, built
    with the reference's formula (zonos/backbone/_torch.py:9-15) so positions round identically."""
    inv_freq = 1.0 / (base ** (torch.arange(0, n_elem, 2)[: n_elem // 2].float() / n_elem))
    angle = torch.outer(torch.arange(seq_len), inv_freq)
    unit = torch.polar(torch.ones_like(angle), angle)          # same primitive as the reference => same roundings
    return torch.stack([unit.real, unit.imag], dim=-1).float().contiguous()


class NativeContext:
    """Owns one zb_ctx (weights arena + KV pages + scratch) on one device."""

    def __init__(self, cfg: BackboneConfig, device: torch.device, eos_token_id: int = 1024, masked_token_id: int = 1025):
        self.lib = _lib.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.ZbError("zonos_vibes_b200 needs a CUDA (sm_100) device; there is no CPU fallback")
        zc = _lib.ZbConfig()
        zc.d_model, zc.n_layer = cfg.d_model, cfg.n_layer
        zc.num_heads, zc.num_heads_kv = cfg.attn_cfg["num_heads"], cfg.attn_cfg["num_heads_kv"]
        zc.ffn, zc.n_codebooks, zc.vocab = cfg.attn_mlp_d_intermediate, N_CODEBOOKS, VOCAB
        zc.eos_token_id, zc.masked_token_id = eos_token_id, masked_token_id
        zc.norm_eps, zc.rope_table_len = cfg.norm_epsilon, ROPE_TABLE_LEN
        self.zc = zc
        self.handle = C.c_void_p()
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        if cfg.ssm_cfg:
            # hybrid: MambaSSMZonosBackbone (zonos/backbone/_mamba_ssm.py:8-57) with mamba-ssm's Mamba2 defaults
            m = mamba2_dims(cfg)
            if cfg.rms_norm:
                raise _lib.ZbError("the hybrid path implements LayerNorm blocks (rms_norm=False)")
            if cfg.residual_in_fp32:
                raise _lib.ZbError("the hybrid path keeps a bf16 residual stream (residual_in_fp32=False); "
                                   "mamba-ssm's fp32 residual is not implemented")
            if len(cfg.attn_layer_idx) > 64:
                raise _lib.ZbError("at most 64 attention layers")
            zh = _lib.ZbHybrid()
            zh.d_state, zh.d_conv, zh.expand, zh.headdim, zh.ngroups = m["d_state"], m["d_conv"], m["expand"], m["headdim"], m["ngroups"]
            zh.n_attn_layers = len(cfg.attn_layer_idx)
            for j, li in enumerate(cfg.attn_layer_idx):
                zh.attn_layer_idx[j] = int(li)
            _lib.check(self.lib.zb_ctx_create_hybrid(C.byref(zc), C.byref(zh), idx, C.byref(self.handle)))
        else:
            _lib.check(self.lib.zb_ctx_create(C.byref(zc), idx, C.byref(self.handle)))
        self.rows = 0
        self.finalized = False
        head_dim = cfg.d_model // zc.num_heads
        table = precompute_freqs_cis(ROPE_TABLE_LEN, head_dim)
        # Hybrid attention layers are mamba-ssm's MHA, which takes its rotary settings from attn_cfg
        # (_mamba_ssm.py:15-33): rotary_emb_dim (0 = no rotary) and rotary_emb_interleaved (default False = half-split
        # pairs (i, i + Dh/2)).  The torch backbone ignores both keys (_torch.py:18-30: interleaved, full head dim).
        self.rope_half_split = False
        if cfg.ssm_cfg:
            rd = int(cfg.attn_cfg.get("rotary_emb_dim", head_dim))
            if rd == 0:
                table = torch.stack([torch.ones_like(table[..., 0]), torch.zeros_like(table[..., 1])], dim=-1).contiguous()
            elif rd != head_dim:
                raise _lib.ZbError(f"hybrid: rotary_emb_dim {rd} must be 0 or the head dim ({head_dim})")
            self.rope_half_split = rd != 0 and not bool(cfg.attn_cfg.get("rotary_emb_interleaved", False))
        self.head_dim = head_dim
        self.set_tensor("rope.freqs_cis", table.to(self.device))

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self.lib.zb_ctx_destroy(self.handle)
                self.handle = None
        except Exception:  # noqa: BLE001
            pass

    def check(self, rc: int) -> None:
        _lib.check(rc, self.handle)

    @property
    def stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    F32_SUFFIXES = ("rope.freqs_cis", ".mixer.dt_bias", ".mixer.A_log", ".mixer.D")

    def _half_split_to_interleaved(self, w: torch.Tensor) -> torch.Tensor:
        """Half-split RoPE pairs rows (i, i + Dh/2) of every q / k head; the kernels rotate adjacent rows (2i, 2i+1).
        Permuting the q and k rows of in_proj identically (new 2i <- old i, new 2i+1 <- old i + Dh/2) makes the two the
        same rotation, and q.k is invariant under a permutation applied to both.  V rows are untouched."""
        dh = self.head_dim
        n_qk = (self.zc.num_heads + self.zc.num_heads_kv) * dh
        perm = torch.stack([torch.arange(dh // 2), torch.arange(dh // 2) + dh // 2], dim=1).flatten()
        idx = (torch.arange(n_qk // dh)[:, None] * dh + perm[None, :]).flatten()
        idx = torch.cat([idx, torch.arange(n_qk, w.shape[0])]).to(w.device)
        return w[idx]

    def set_tensor(self, name: str, t: torch.Tensor) -> None:
        want = torch.float32 if name.endswith(self.F32_SUFFIXES) else torch.bfloat16
        if self.rope_half_split and name.endswith(".mixer.in_proj.weight") and t.shape[0] == (self.zc.num_heads + 2 * self.zc.num_heads_kv) * self.head_dim:
            t = self._half_split_to_interleaved(t.detach())
        t = t.detach().to(self.device, want).contiguous()
        with torch.cuda.device(self.device):
            self.check(self.lib.zb_set_tensor(self.handle, name.encode(), t.data_ptr(), t.numel(),
                                              _lib.ZB_F32 if want == torch.float32 else _lib.ZB_BF16, self.stream))
            torch.cuda.current_stream(self.device).synchronize()     # `t` may be a temporary

    def finalize(self) -> None:
        with torch.cuda.device(self.device):
            self.check(self.lib.zb_weights_finalize(self.handle, self.stream))
        self.finalized = True

    def kv_alloc(self, rows: int, max_seqlen: int) -> None:
        with torch.cuda.device(self.device):
            self.check(self.lib.zb_kv_alloc(self.handle, rows, max_seqlen))
            self.check(self.lib.zb_kv_reset(self.handle, self.stream))
        self.rows = rows

    def set_option(self, key: str, value: int) -> None:
        self.check(self.lib.zb_set_option(self.handle, key.encode(), int(value)))


def mamba2_dims(cfg: BackboneConfig) -> dict:
    """Mamba2 constructor arguments: ssm_cfg overrides on top of mamba-ssm 2.2.4's defaults (d_state=128, d_conv=4,
    expand=2, headdim=64, ngroups=1)."""
    sc = dict(cfg.ssm_cfg)
    layer = sc.pop("layer", "Mamba2")
    if layer != "Mamba2":
        raise _lib.ZbError(f"ssm_cfg layer {layer!r} is not supported (Mamba2 only)")
    m = dict(d_state=128, d_conv=4, expand=2, headdim=64, ngroups=1)
    m.update({k: v for k, v in sc.items() if k in m})
    m["d_inner"] = m["expand"] * cfg.d_model
    m["nheads"] = m["d_inner"] // m["headdim"]
    m["conv_dim"] = m["d_inner"] + 2 * m["ngroups"] * m["d_state"]
    m["d_in_proj"] = 2 * m["d_inner"] + 2 * m["ngroups"] * m["d_state"] + m["nheads"]
    return m


class _Norm(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(d))
        self.bias = nn.Parameter(torch.zeros(d))


class _Proj(nn.Module):
    def __init__(self, out_f, in_f):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(out_f, in_f).uniform_(-in_f ** -0.5, in_f ** -0.5))


class _Mixer(nn.Module):
    def __init__(self, cfg: BackboneConfig):
        super().__init__()
        h, hkv = cfg.attn_cfg["num_heads"], cfg.attn_cfg["num_heads_kv"]
        dh = cfg.d_model // h
        self.in_proj = _Proj((h + 2 * hkv) * dh, cfg.d_model)
        self.out_proj = _Proj(cfg.d_model, h * dh)


class _Mlp(nn.Module):
    def __init__(self, cfg: BackboneConfig):
        super().__init__()
        self.fc1 = _Proj(2 * cfg.attn_mlp_d_intermediate, cfg.d_model)
        self.fc2 = _Proj(cfg.d_model, cfg.attn_mlp_d_intermediate)


class _Conv1d(nn.Module):
    def __init__(self, channels, k):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(channels, 1, k).uniform_(-k ** -0.5, k ** -0.5))
        self.bias = nn.Parameter(torch.zeros(channels))


class _GatedNorm(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(d))


class _Mamba2Mixer(nn.Module):
    """Parameter holder with mamba_ssm.modules.mamba2.Mamba2's names and shapes."""

    def __init__(self, cfg: BackboneConfig):
        super().__init__()
        m = mamba2_dims(cfg)
        self.in_proj = _Proj(m["d_in_proj"], cfg.d_model)
        self.conv1d = _Conv1d(m["conv_dim"], m["d_conv"])
        self.dt_bias = nn.Parameter(torch.zeros(m["nheads"]))
        self.A_log = nn.Parameter(torch.zeros(m["nheads"]))
        self.D = nn.Parameter(torch.ones(m["nheads"]))
        self.norm = _GatedNorm(m["d_inner"])
        self.out_proj = _Proj(cfg.d_model, m["d_inner"])


class _Block(nn.Module):
    def __init__(self, cfg: BackboneConfig, mamba: bool = False):
        super().__init__()
        self.norm = _Norm(cfg.d_model)
        if mamba:                                  # _mamba_ssm.py:19-21: d_intermediate = 0 => no MLP on Mamba2 layers
            self.mixer = _Mamba2Mixer(cfg)
            return
        self.mixer = _Mixer(cfg)
        self.norm2 = _Norm(cfg.d_model)
        self.mlp = _Mlp(cfg)


class B200ZonosBackbone(nn.Module):
    """Drop-in for `TorchZonosBackbone` (zonos/backbone/_torch.py:52-80): same constructor, same state_dict keys,
    same `allocate_inference_cache` / `forward` contract.  Parameters are plain holders: the arithmetic runs in
    libzonos_b200 from a repacked device arena that is (re)filled from them on first use."""

    supported_architectures = ["transformer", "hybrid"]

    def __init__(self, config: BackboneConfig):
        super().__init__()
        self.config = config
        if config.ssm_cfg and config.d_intermediate != 0:
            raise _lib.ZbError("hybrid: Mamba2 layers with an MLP (d_intermediate != 0) are not supported")
        hybrid = bool(config.ssm_cfg)
        self.layers = nn.ModuleList(_Block(config, mamba=hybrid and i not in config.attn_layer_idx) for i in range(config.n_layer))
        self.norm_f = _Norm(config.d_model)
        self._ctx: NativeContext | None = None
        self._extra_weights: dict[str, torch.Tensor] = {}
        self.eos_token_id, self.masked_token_id = 1024, 1025

    # -- native context ---------------------------------------------------------------------------
    def native(self) -> NativeContext:
        if self._ctx is None:
            dev = self.norm_f.weight.device
            ctx = NativeContext(self.config, dev, self.eos_token_id, self.masked_token_id)
            for name, p in self.state_dict().items():
                ctx.set_tensor("backbone." + name, p)
            d = self.config.d_model
            zeros = torch.zeros(VOCAB, d, dtype=torch.bfloat16, device=ctx.device)
            for k in range(N_CODEBOOKS):      # heads/embeddings belong to `Zonos`; zero unless it provides them
                ctx.set_tensor(f"heads.{k}.weight", self._extra_weights.get(f"heads.{k}.weight", zeros))
                ctx.set_tensor(f"embeddings.{k}.weight", self._extra_weights.get(f"embeddings.{k}.weight", zeros))
            ctx.finalize()
            self._ctx = ctx
        return self._ctx

    def invalidate(self) -> None:
        """Call after changing parameters (e.g. load_state_dict) so the arena is refilled on next use."""
        self._ctx = None

    def _load_from_state_dict(self, *args, **kwargs):
        self._ctx = None
        return super()._load_from_state_dict(*args, **kwargs)

    # -- plugin API --------------------------------------------------------------------------------
    def allocate_inference_cache(self, batch_size: int, max_seqlen: int, dtype: torch.dtype = torch.bfloat16):
        """_torch.py:64-71: returns {layer_idx: handle}.  The pages live in the native context (bf16, 128-token pages)."""
        if dtype != torch.bfloat16:
            raise _lib.ZbError("the B200 KV cache is bf16")
        ctx = self.native()
        ctx.kv_alloc(batch_size, max_seqlen)
        return {i: ("zb_kv", i) for i in range(self.config.n_layer)}

    def forward(self, hidden_states: torch.Tensor, inference_params: InferenceParams) -> torch.Tensor:
        """_torch.py:73-80: [rows, S, d] bf16 -> [rows, S, d] (post-norm_f); appends S tokens to the KV cache at
        `inference_params.seqlen_offset`, RoPE position = lengths_per_sample[r] + s."""
        ctx = self.native()
        x = hidden_states.to(torch.bfloat16).contiguous()
        rows, s, d = x.shape
        out = torch.empty_like(x)
        lengths = inference_params.lengths_per_sample
        lp = None
        if lengths is not None:
            lengths = lengths.to(ctx.device, torch.int32).contiguous()
            lp = lengths.data_ptr()
        with torch.cuda.device(ctx.device):
            ctx.check(ctx.lib.zb_backbone_forward(ctx.handle, x.data_ptr(), out.data_ptr(), rows, s,
                                                  int(inference_params.seqlen_offset), lp, ctx.stream))
        return out


BACKBONES = {"b200": B200ZonosBackbone}
