"""Data contract of the hot path.

The four records below carry exactly the fields (names, order, defaults) of the reference's configuration records
(zonos/config.py:7-62), because checkpoints' `config.json` files and the reference's callers construct them positionally and
by keyword.  They are declared from field tables so that the contract is stated once, with what each field means for THIS
implementation, and validated where the native library would otherwise fail late.
"""
from __future__ import annotations

import dataclasses as _dc
from typing import Any, Literal

import torch

_REQUIRED = _dc.MISSING


def _record(name: str, doc: str, table: list[tuple[str, Any, Any, str]], methods: dict | None = None):
    """Builds a dataclass from rows of (field, type, default | factory | _REQUIRED, meaning)."""
    fields = []
    for fname, ftype, default, _meaning in table:
        if default is _REQUIRED:
            fields.append((fname, ftype))
        elif callable(default):
            fields.append((fname, ftype, _dc.field(default_factory=default)))
        else:
            fields.append((fname, ftype, _dc.field(default=default)))
    cls = _dc.make_dataclass(name, fields, namespace=dict(methods or {}))
    cls.__doc__ = doc + "\n\n" + "\n".join(f"    {f}: {m}" for f, _t, _d, m in table)
    cls.__module__ = __name__
    return cls


# ---- per-call decoding state handed to a backbone (reference: zonos/config.py:7-25) --------------------------------------
def _reset(self, max_seqlen, max_batch_size):
    self.max_seqlen, self.max_batch_size, self.seqlen_offset = max_seqlen, max_batch_size, 0
    if self.lengths_per_sample is not None:
        self.lengths_per_sample.zero_()


InferenceParams = _record(
    "InferenceParams", "What `backbone.forward(hidden, inference_params)` needs to know about the cache (plugin API, INTEGRATION.md 2b).",
    [
        ("max_seqlen", int, _REQUIRED, "tokens the KV cache was allocated for"),
        ("max_batch_size", int, _REQUIRED, "rows of the cache (2 x utterances with CFG)"),
        ("seqlen_offset", int, 0, "tokens already in the cache (host integer, advanced by the caller)"),
        ("batch_size_offset", int, 0, "kept for signature compatibility; always 0 here"),
        ("key_value_memory_dict", dict, dict, "layer index -> opaque handle: the pages live in the native context"),
        ("lengths_per_sample", "torch.Tensor | None", None, "int32 device tensor [rows]: per-row length, added to the RoPE position"),
    ],
    {"reset": _reset},
)

# ---- backbone architecture (reference: zonos/config.py:28-39) ----------------------------------------------------------------
BackboneConfig = _record(
    "BackboneConfig", "Architecture of the autoregressive backbone (transformer, or hybrid Mamba2 + attention when `ssm_cfg` is set).",
    [
        ("d_model", int, 1024, "residual width; a multiple of 256, at most 2048 in the native kernels"),
        ("d_intermediate", int, 0, "MLP width of Mamba2 blocks (0: none, the only value the hybrid checkpoints use)"),
        ("attn_mlp_d_intermediate", int, 0, "SwiGLU width of attention blocks (fc1 has twice as many rows)"),
        ("n_layer", int, 16, "number of blocks"),
        ("ssm_cfg", dict, dict, "non-empty ({'layer': 'Mamba2', ...}) selects the hybrid stack"),
        ("attn_layer_idx", list, list, "hybrid: which blocks are attention blocks"),
        ("attn_cfg", dict, dict, "num_heads, num_heads_kv, and for the hybrid rotary_emb_dim / rotary_emb_interleaved"),
        ("rms_norm", bool, False, "must stay False (LayerNorm blocks only)"),
        ("residual_in_fp32", bool, False, "must stay False for the hybrid (bf16 residual stream)"),
        ("norm_epsilon", float, 1e-5, "LayerNorm epsilon"),
    ],
)

# ---- conditioning prefix (reference: zonos/config.py:42-45) --------------------------------------------------------------------
PrefixConditionerConfig = _record(
    "PrefixConditionerConfig", "Which conditioners build the prefix (zonos_vibes_b200/conditioning.py) and how it is projected.",
    [
        ("conditioners", "list[dict]", _REQUIRED, "one dict per conditioner: type, name, and its constructor arguments"),
        ("projection", 'Literal["none", "linear", "mlp"]', _REQUIRED, "projection applied to the concatenated prefix"),
    ],
)


# ---- the model (reference: zonos/config.py:48-62) --------------------------------------------------------------------------------
def _zonos_from_dict(cls, d: dict):
    """From a checkpoint's config.json: nested dicts become the records above; unknown top-level keys are an error here
    rather than a TypeError three frames down."""
    rest = {k: v for k, v in d.items() if k not in ("backbone", "prefix_conditioner")}
    known = {f.name for f in _dc.fields(cls)}
    extra = sorted(set(rest) - known)
    if extra:
        raise TypeError(f"ZonosConfig.from_dict: unknown keys {extra}")
    return cls(BackboneConfig(**d["backbone"]), PrefixConditionerConfig(**d["prefix_conditioner"]), **rest)


def _zonos_to_dict(self) -> dict:
    """Inverse of from_dict (json-serialisable)."""
    return _dc.asdict(self)


ZonosConfig = _record(
    "ZonosConfig", "Top-level model configuration.",
    [
        ("backbone", BackboneConfig, _REQUIRED, "see BackboneConfig"),
        ("prefix_conditioner", PrefixConditionerConfig, _REQUIRED, "see PrefixConditionerConfig"),
        ("eos_token_id", int, 1024, "end-of-speech code; 1025 disables EOS (fixed-length benchmark runs)"),
        ("masked_token_id", int, 1025, "the delay pattern's fill value"),
        ("pad_vocab_to_multiple_of", int, 8, "embeddings / heads are padded to 1026 rows"),
    ],
    {"from_dict": classmethod(_zonos_from_dict), "to_dict": _zonos_to_dict},
)

__all__ = ["InferenceParams", "BackboneConfig", "PrefixConditionerConfig", "ZonosConfig"]
