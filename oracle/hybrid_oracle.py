"""CPU oracle of the reference's HYBRID backbone (Mamba2 + attention layers) — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this file; the product
(zonos_vibes_b200/) never does.

What it restates
  * the stack wiring of `zonos/backbone/_mamba_ssm.py:8-57` (MambaSSMZonosBackbone): `create_block` per layer with
    `fused_add_norm=True`, Mamba2 mixer on every layer except `attn_layer_idx`, where the mixer is MHA and the block
    gets a GatedMLP of `attn_mlp_d_intermediate`; `d_intermediate` (0 for the hybrid checkpoint) applies to the Mamba
    layers => no MLP there; final `layer_norm_fn(hidden, norm_f, residual)`.
  * the arithmetic of `mamba-ssm == 2.2.4` (pinned by the reference's uv.lock:681-682; NOT vendored, NOT installed,
    no network).  Its published algorithm is restated from the pure-torch port with the same parameter layout that IS
    on this box: `transformers/models/mamba2/modeling_mamba2.py:398-585` (`Mamba2Mixer.torch_forward`):
      in_proj -> [z, xBC, dt]                                              (:409-413)
      depthwise causal conv (k=d_conv) + SiLU on xBC, rolling conv state    (:419-436)
      dt = softplus(dt + dt_bias); dA = exp(dt * A), A = -exp(A_log)        (:447-460)
      state = state * dA + dt * B (x) x ; y = state . C + D * x             (:461-502)
      gated RMSNorm: rmsnorm(y * silu(z)) * w                               (:103-118,579)
      out_proj                                                              (:584)
    and `mamba_ssm.modules.block.Block.forward` (fused add + norm, residual kept separately):
      residual = hidden + residual ; hidden = norm(residual) ; hidden = mixer(hidden) ; [same with norm2 / mlp]
  * attention layers: `flash_attn.modules.mha.MHA` semantics = the torch backbone's Attention (`_torch.py:105-141`) with
    rotary dimension `rotary_emb_dim` (interleaved pairs configurable) — reused from zonos_oracle.

Pinning.  `tests/test_oracle_pinned.py::test_mamba2_mixer_matches_hf_port` checks `mamba2_mixer_*` here against
`transformers`' Mamba2Mixer.torch_forward (chunked SSD over a whole sequence) on identical random weights.  The STACK
(which layers are attention, dims of the hybrid checkpoint) cannot be pinned: the checkpoint's config.json is not in the
repo and mamba-ssm is absent — header says so: **stack parity unpinned**.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch
import torch.nn.functional as F


@dataclass
class HybridConfig:
    """BackboneConfig fields (zonos/config.py:28-39) + Mamba2 defaults (ssm_cfg={"layer": "Mamba2"})."""
    d_model: int = 2048
    n_layer: int = 46
    attn_layer_idx: list = field(default_factory=lambda: [9, 18, 27, 36, 45])
    num_heads: int = 16
    num_heads_kv: int = 4
    attn_mlp_d_intermediate: int = 8192
    d_intermediate: int = 0
    d_state: int = 128
    d_conv: int = 4
    expand: int = 2
    headdim: int = 64
    ngroups: int = 1
    norm_epsilon: float = 1e-5
    rms_norm: bool = False
    # attn_cfg keys forwarded to mamba-ssm's MHA by create_block (_mamba_ssm.py:15-33); MHA's own default is
    # rotary_emb_interleaved=False (GPT-NeoX half-split pairs (i, i + rotary_dim/2), flash_attn/layers/rotary.py:14-35)
    rotary_emb_dim: int = 128
    rotary_emb_interleaved: bool = False

    def to_reference_dict(self) -> dict:
        """The dict ZonosConfig.from_dict (zonos/config.py:56-62) expects for this architecture."""
        return dict(backbone=dict(d_model=self.d_model, d_intermediate=self.d_intermediate, attn_mlp_d_intermediate=self.attn_mlp_d_intermediate,
                                  n_layer=self.n_layer, ssm_cfg=dict(layer="Mamba2", d_state=self.d_state, d_conv=self.d_conv, expand=self.expand,
                                                                     headdim=self.headdim, ngroups=self.ngroups),
                                  attn_layer_idx=list(self.attn_layer_idx), attn_cfg=dict(num_heads=self.num_heads, num_heads_kv=self.num_heads_kv, rotary_emb_dim=self.rotary_emb_dim,
                                                rotary_emb_interleaved=self.rotary_emb_interleaved),
                                  rms_norm=self.rms_norm, residual_in_fp32=False, norm_epsilon=self.norm_epsilon),
                    prefix_conditioner=dict(conditioners=[], projection="none"), eos_token_id=1024, masked_token_id=1025)

    @property
    def d_inner(self) -> int:
        return self.expand * self.d_model

    @property
    def nheads(self) -> int:
        return self.d_inner // self.headdim

    @property
    def conv_dim(self) -> int:
        return self.d_inner + 2 * self.ngroups * self.d_state

    @property
    def d_in_proj(self) -> int:
        return 2 * self.d_inner + 2 * self.ngroups * self.d_state + self.nheads


HYBRID_TINY = HybridConfig(d_model=256, n_layer=4, attn_layer_idx=[2], num_heads=2, num_heads_kv=1, attn_mlp_d_intermediate=512,
                           d_state=128, headdim=64)


def synth_mamba2_params(cfg: HybridConfig, seed: int, dtype=torch.bfloat16) -> dict:
    """Random-init parameters of one Mamba2 mixer with mamba_ssm.modules.mamba2.Mamba2's names and shapes."""
    g = torch.Generator().manual_seed(seed)
    d, di = cfg.d_model, cfg.d_inner
    p = {}
    p["in_proj.weight"] = ((torch.rand(cfg.d_in_proj, d, generator=g) * 2 - 1) / math.sqrt(d)).to(dtype)
    p["conv1d.weight"] = ((torch.rand(cfg.conv_dim, 1, cfg.d_conv, generator=g) * 2 - 1) / math.sqrt(cfg.d_conv)).to(dtype)
    p["conv1d.bias"] = (0.1 * torch.randn(cfg.conv_dim, generator=g)).to(dtype)
    dt = torch.exp(torch.rand(cfg.nheads, generator=g) * (math.log(0.1) - math.log(0.001)) + math.log(0.001)).clamp(min=1e-4)
    # stored in the model dtype like every other parameter of the reference module (`.to(device, bf16)`), used in fp32
    p["dt_bias"] = (dt + torch.log(-torch.expm1(-dt))).to(dtype).float()
    p["A_log"] = torch.log(1 + 15 * torch.rand(cfg.nheads, generator=g)).to(dtype).float()
    p["D"] = (torch.ones(cfg.nheads) + 0.1 * torch.randn(cfg.nheads, generator=g)).to(dtype).float()
    p["norm.weight"] = (1 + 0.1 * torch.randn(di, generator=g)).to(dtype)
    p["out_proj.weight"] = ((torch.rand(d, di, generator=g) * 2 - 1) / math.sqrt(di)).to(dtype)
    return p


def mamba2_alloc_state(cfg: HybridConfig, batch: int) -> dict:
    return {"conv": torch.zeros(batch, cfg.conv_dim, cfg.d_conv), "ssm": torch.zeros(batch, cfg.nheads, cfg.headdim, cfg.d_state)}


def mamba2_mixer_step(cfg: HybridConfig, p: dict, state: dict, u: torch.Tensor) -> torch.Tensor:
    """One token through the Mamba2 mixer (mamba_ssm Mamba2.step; HF torch_forward cached branch :419-436,447-502).
    u: [B, d_model] (bf16).  state: fp32 conv [B, conv_dim, d_conv], ssm [B, nheads, headdim, d_state].  Returns [B, d_model]."""
    dtype = u.dtype
    B = u.shape[0]
    di, ds, ng, nh, hd = cfg.d_inner, cfg.d_state, cfg.ngroups, cfg.nheads, cfg.headdim
    zxbcdt = F.linear(u, p["in_proj.weight"])                                          # bf16 GEMV, fp32 accumulate
    z, xBC, dt = torch.split(zxbcdt, [di, cfg.conv_dim, nh], dim=-1)
    # depthwise conv over the last d_conv inputs (rolling state), + bias, SiLU
    conv = state["conv"]
    conv.copy_(torch.roll(conv, shifts=-1, dims=-1))
    conv[:, :, -1] = xBC.float()
    xBC = torch.sum(conv * p["conv1d.weight"].float().squeeze(1), dim=-1) + p["conv1d.bias"].float()
    xBC = F.silu(xBC).to(dtype)
    x, Bm, Cm = torch.split(xBC.float(), [di, ng * ds, ng * ds], dim=-1)
    A = -torch.exp(p["A_log"].float())                                                  # [nh]
    dtf = F.softplus(dt.float() + p["dt_bias"].float())                                 # [B, nh]
    dA = torch.exp(dtf * A)                                                             # [B, nh]
    x = x.view(B, nh, hd)
    Bg = Bm.view(B, ng, ds).repeat_interleave(nh // ng, dim=1)                          # [B, nh, ds]
    Cg = Cm.view(B, ng, ds).repeat_interleave(nh // ng, dim=1)
    ssm = state["ssm"]
    ssm.copy_(ssm * dA[:, :, None, None] + (dtf[:, :, None] * x)[..., None] * Bg[:, :, None, :])
    y = torch.einsum("bhpn,bhn->bhp", ssm, Cg) + p["D"].float()[None, :, None] * x
    y = y.reshape(B, di).to(dtype)
    # gated RMSNorm (norm_before_gate=False): rmsnorm(y * silu(z)) * w, group = d_inner / ngroups
    yz = y.float() * F.silu(z.float())
    yz = yz.view(B, ng, di // ng)
    yz = yz * torch.rsqrt(yz.pow(2).mean(-1, keepdim=True) + 1e-5)
    y = (yz.view(B, di) * p["norm.weight"].float()).to(dtype)
    return F.linear(y, p["out_proj.weight"])


def mamba2_mixer_sequence(cfg: HybridConfig, p: dict, state: dict, u: torch.Tensor) -> torch.Tensor:
    """u [B, S, d_model] -> [B, S, d_model] by stepping (the chunked SSD scan of the prefill computes the same recurrence)."""
    return torch.stack([mamba2_mixer_step(cfg, p, state, u[:, t]) for t in range(u.shape[1])], dim=1)


# ---------------------------------------------------------------------------------------------------------------
# the hybrid stack (_mamba_ssm.py:43-57 + mamba_ssm Block)
# ---------------------------------------------------------------------------------------------------------------
def _norm(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor | None, eps: float, rms: bool) -> torch.Tensor:
    xf = x.float()
    if rms:
        y = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps) * w.float()
    else:
        y = F.layer_norm(xf, (xf.shape[-1],), w.float(), b.float() if b is not None else None, eps)
    return y.to(x.dtype)


def synth_hybrid_state_dict(cfg: HybridConfig, seed: int = 0, dtype=torch.bfloat16) -> dict:
    """Keys follow mamba_ssm's Block (norm / mixer / norm2 / mlp), like the reference checkpoint."""
    g = torch.Generator().manual_seed(seed)
    d = cfg.d_model
    sd = {}

    def ln(prefix):
        sd[prefix + ".weight"] = (1 + 0.1 * torch.randn(d, generator=g)).to(dtype)
        sd[prefix + ".bias"] = (0.1 * torch.randn(d, generator=g)).to(dtype)

    def lin(o, i):
        return ((torch.rand(o, i, generator=g) * 2 - 1) / math.sqrt(i)).to(dtype)

    for i in range(cfg.n_layer):
        pfx = f"backbone.layers.{i}"
        ln(pfx + ".norm")
        if i in cfg.attn_layer_idx:
            qkv = (cfg.num_heads + 2 * cfg.num_heads_kv) * 128
            sd[pfx + ".mixer.in_proj.weight"] = lin(qkv, d)
            sd[pfx + ".mixer.out_proj.weight"] = lin(d, cfg.num_heads * 128)
            ln(pfx + ".norm2")
            sd[pfx + ".mlp.fc1.weight"] = lin(2 * cfg.attn_mlp_d_intermediate, d)
            sd[pfx + ".mlp.fc2.weight"] = lin(d, cfg.attn_mlp_d_intermediate)
        else:
            for k, v in synth_mamba2_params(cfg, 1000 * (seed + 1) + i, dtype).items():
                sd[f"{pfx}.mixer.{k}"] = v
    ln("backbone.norm_f")
    return sd


class HybridOracleBackbone:
    """forward(hidden [B, S, d], ...) token by token; state dict per layer: Mamba2 {conv, ssm} or attention KV lists."""

    def __init__(self, cfg: HybridConfig, sd: dict):
        self.cfg, self.sd = cfg, sd
        self.state = None
        self.pos = 0

    def allocate(self, batch: int, max_seqlen: int):
        c = self.cfg
        self.state = {}
        for i in range(c.n_layer):
            if i in c.attn_layer_idx:
                self.state[i] = {"k": torch.zeros(batch, max_seqlen, c.num_heads_kv, 128, dtype=torch.bfloat16),
                                 "v": torch.zeros(batch, max_seqlen, c.num_heads_kv, 128, dtype=torch.bfloat16)}
            else:
                self.state[i] = mamba2_alloc_state(c, batch)
        self.pos = 0

    def _attn(self, i: int, h: torch.Tensor) -> torch.Tensor:
        """MHA semantics for one token: RoPE over the full head dim (interleaved or half-split pairs per attn_cfg), GQA."""
        c, sd = self.cfg, self.sd
        B = h.shape[0]
        H, Hkv = c.num_heads, c.num_heads_kv
        qkv = F.linear(h, sd[f"backbone.layers.{i}.mixer.in_proj.weight"])
        q, k, v = qkv.split([H * 128, Hkv * 128, Hkv * 128], dim=-1)
        q, k, v = q.view(B, H, 128), k.view(B, Hkv, 128), v.view(B, Hkv, 128)
        freqs = 1.0 / (10000 ** (torch.arange(0, 128, 2)[:64].float() / 128))
        ang = self.pos * freqs
        cos, sin = torch.cos(ang), torch.sin(ang)

        def rope(x):
            if c.rotary_emb_dim == 0:
                return x
            assert c.rotary_emb_dim == 128, "rotary over the full head dim (or none)"
            if c.rotary_emb_interleaved:              # pairs (2i, 2i+1), like _torch.py:18-30
                xf = x.float().reshape(*x.shape[:-1], 64, 2)
                out = torch.stack([xf[..., 0] * cos - xf[..., 1] * sin, xf[..., 1] * cos + xf[..., 0] * sin], -1)
                return out.flatten(-2).to(x.dtype)
            x1, x2 = x.float()[..., :64], x.float()[..., 64:]          # pairs (i, i + 64): rotate_half form
            return torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin], -1).to(x.dtype)

        q, k = rope(q), rope(k)
        st = self.state[i]
        st["k"][:, self.pos] = k
        st["v"][:, self.pos] = v
        K = st["k"][:, : self.pos + 1].transpose(1, 2)              # [B, Hkv, ctx, 128]
        V = st["v"][:, : self.pos + 1].transpose(1, 2)
        y = F.scaled_dot_product_attention(q.unsqueeze(2), K, V, enable_gqa=True)      # [B, H, 1, 128]
        y = y.squeeze(2).reshape(B, H * 128)
        return F.linear(y, sd[f"backbone.layers.{i}.mixer.out_proj.weight"])

    def step(self, h: torch.Tensor) -> torch.Tensor:
        """One token [B, d] -> post-norm_f [B, d]."""
        c, sd = self.cfg, self.sd
        residual = None
        hidden = h
        for i in range(c.n_layer):
            pfx = f"backbone.layers.{i}"
            residual = hidden if residual is None else (hidden.float() + residual.float()).to(hidden.dtype)
            hidden = _norm(residual, sd[pfx + ".norm.weight"], sd.get(pfx + ".norm.bias"), c.norm_epsilon, c.rms_norm)
            if i in c.attn_layer_idx:
                hidden = self._attn(i, hidden)
                residual = (hidden.float() + residual.float()).to(hidden.dtype)
                hidden = _norm(residual, sd[pfx + ".norm2.weight"], sd.get(pfx + ".norm2.bias"), c.norm_epsilon, c.rms_norm)
                y, gate = F.linear(hidden, sd[pfx + ".mlp.fc1.weight"]).chunk(2, dim=-1)
                hidden = F.linear(y * F.silu(gate), sd[pfx + ".mlp.fc2.weight"])
            else:
                p = {k[len(pfx + ".mixer."):]: v for k, v in sd.items() if k.startswith(pfx + ".mixer.")}
                hidden = mamba2_mixer_step(c, p, self.state[i], hidden)
        residual = (hidden.float() + residual.float()).to(hidden.dtype)
        self.pos += 1
        return _norm(residual, sd["backbone.norm_f.weight"], sd.get("backbone.norm_f.bias"), c.norm_epsilon, c.rms_norm)

    def forward(self, hidden: torch.Tensor) -> torch.Tensor:
        return torch.stack([self.step(hidden[:, t]) for t in range(hidden.shape[1])], dim=1)
