"""`PrefixConditioner` on the device — mirror of the reference's zonos/conditioning.py:11-52,230-313.

Same classes, constructor arguments, `state_dict` keys and `forward` contract as the reference, so the
`prefix_conditioner.*` entries of a reference checkpoint load into it; the arithmetic (embedding gathers, Fourier
features, per-conditioner and prefix projections, final LayerNorm) runs in libzonos_b200 (`zb_pc_*`, csrc/prefix_kernels.cuh).

What stays on the host, as in the reference's "espeak containment zone" (conditioning.py:55-250): text normalisation and
phonemisation.  `EspeakPhonemeConditioner` therefore takes either phoneme-id tensors (`("ids", LongTensor[B, S])`, what
the reference's `tokenize_phonemes` returns) or `(texts, languages)` together with a `phonemize_to_ids` callable supplied
by the caller (e.g. `lambda t, l: zonos.conditioning.tokenize_phonemes(zonos.conditioning.phonemize(t, l))[0]`).
"""
from __future__ import annotations

from typing import Any, Callable, Literal

import torch
import torch.nn as nn

from . import _lib
from .config import PrefixConditionerConfig

PHONEME_VOCAB = 189          # len(SPECIAL_TOKEN_IDS) + len(symbols) in zonos/conditioning.py (4 + 185)


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _need_cuda(t: torch.Tensor) -> None:
    if not t.is_cuda:
        raise _lib.ZbError("zonos_vibes_b200 conditioners run on a CUDA (sm_100) device; there is no CPU fallback")


def _linear(x: torch.Tensor, lin: nn.Linear, silu: bool = False) -> torch.Tensor:
    """nn.Linear (+ SiLU) on [n, K] bf16 rows through zb_pc_linear."""
    _need_cuda(x)
    x = x.to(torch.bfloat16).contiguous()
    n, k = x.shape
    w = lin.weight.detach().to(torch.bfloat16).contiguous()
    b = lin.bias.detach().to(torch.bfloat16).contiguous() if lin.bias is not None else None
    out = torch.empty((n, w.shape[0]), dtype=torch.bfloat16, device=x.device)
    lib = _lib.load()
    with torch.cuda.device(x.device):
        _lib.check(lib.zb_pc_linear(x.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None, out.data_ptr(), n, k, w.shape[0],
                                    1 if silu else 0, _stream(x)))
    return out


def _project(x: torch.Tensor, project: nn.Module) -> torch.Tensor:
    """Conditioner.project (conditioning.py:24-35): Identity, Linear or Linear-SiLU-Linear, on [..., K]."""
    if isinstance(project, nn.Identity):
        return x
    lead = x.shape[:-1]
    rows = x.reshape(-1, x.shape[-1])
    if isinstance(project, nn.Linear):
        rows = _linear(rows, project)
    else:
        rows = _linear(_linear(rows, project[0], silu=True), project[2])
    return rows.view(*lead, rows.shape[-1])


class Conditioner(nn.Module):
    """conditioning.py:11-52."""

    def __init__(self, output_dim: int, name: str, cond_dim: int | None = None, projection: Literal["none", "linear", "mlp"] = "none",
                 uncond_type: Literal["learned", "none"] = "none", **kwargs):
        super().__init__()
        self.name = name
        self.output_dim = output_dim
        self.cond_dim = cond_dim = cond_dim or output_dim
        if projection == "linear":
            self.project = nn.Linear(cond_dim, output_dim)
        elif projection == "mlp":
            self.project = nn.Sequential(nn.Linear(cond_dim, output_dim), nn.SiLU(), nn.Linear(output_dim, output_dim))
        else:
            self.project = nn.Identity()
        self.uncond_vector = None
        if uncond_type == "learned":
            self.uncond_vector = nn.Parameter(torch.zeros(output_dim))

    def apply_cond(self, *inputs: Any) -> torch.Tensor:
        raise NotImplementedError()

    def forward(self, inputs: tuple[Any, ...] | None) -> torch.Tensor:
        if inputs is None:
            assert self.uncond_vector is not None
            return self.uncond_vector.data.view(1, 1, -1)
        return _project(self.apply_cond(*inputs), self.project)


class EspeakPhonemeConditioner(Conditioner):
    """conditioning.py:218-236.  Phonemisation stays on the host (module docstring)."""

    def __init__(self, output_dim: int, phonemize_to_ids: Callable[[list, list], torch.Tensor] | None = None, **kwargs):
        super().__init__(output_dim, **kwargs)
        self.phoneme_embedder = nn.Embedding(PHONEME_VOCAB, output_dim)
        self.phonemize_to_ids = phonemize_to_ids

    def apply_cond(self, texts_or_tag, languages_or_ids) -> torch.Tensor:
        if isinstance(texts_or_tag, str) and texts_or_tag == "ids":
            ids = languages_or_ids
        else:
            if self.phonemize_to_ids is None:
                raise _lib.ZbError("EspeakPhonemeConditioner: pass ('ids', phoneme_ids) or construct with phonemize_to_ids= "
                                   "(text normalisation / espeak stay on the host, outside the accelerated path)")
            ids = self.phonemize_to_ids(texts_or_tag, languages_or_ids)
        table = self.phoneme_embedder.weight
        _need_cuda(table)
        ids = ids.to(table.device, torch.int64).contiguous()
        b, s = ids.shape
        tb = table.detach().to(torch.bfloat16).contiguous()
        out = torch.empty((b, s, tb.shape[1]), dtype=torch.bfloat16, device=table.device)
        lib = _lib.load()
        with torch.cuda.device(table.device):
            _lib.check(lib.zb_pc_embed(ids.data_ptr(), tb.data_ptr(), out.data_ptr(), b * s, tb.shape[1], 0, tb.shape[0], _stream(tb)))
        return out


class FourierConditioner(Conditioner):
    """conditioning.py:242-261."""

    def __init__(self, output_dim: int, input_dim: int = 1, std: float = 1.0, min_val: float = 0.0, max_val: float = 1.0, **kwargs):
        assert output_dim % 2 == 0
        super().__init__(output_dim, **kwargs)
        self.register_buffer("weight", torch.randn([output_dim // 2, input_dim]) * std)
        self.input_dim, self.min_val, self.max_val = input_dim, min_val, max_val

    def apply_cond(self, x: torch.Tensor) -> torch.Tensor:
        assert x.shape[-1] == self.input_dim
        _need_cuda(self.weight)
        w = self.weight.detach().to(torch.bfloat16).contiguous()
        xf = x.to(w.device, torch.float32).contiguous()
        lead = xf.shape[:-1]
        n = xf.numel() // self.input_dim
        half = w.shape[0]
        out = torch.empty((*lead, 2 * half), dtype=torch.bfloat16, device=w.device)
        lib = _lib.load()
        with torch.cuda.device(w.device):
            _lib.check(lib.zb_pc_fourier(xf.data_ptr(), w.data_ptr(), float(self.min_val), float(self.max_val), out.data_ptr(), n, self.input_dim, half,
                                         _stream(w)))
        return out


class IntegerConditioner(Conditioner):
    """conditioning.py:264-273."""

    def __init__(self, output_dim: int, min_val: int = 0, max_val: int = 512, **kwargs):
        super().__init__(output_dim, **kwargs)
        self.min_val, self.max_val = min_val, max_val
        self.int_embedder = nn.Embedding(max_val - min_val + 1, output_dim)

    def apply_cond(self, x: torch.Tensor) -> torch.Tensor:
        assert x.shape[-1] == 1
        table = self.int_embedder.weight
        _need_cuda(table)
        ids = x.squeeze(-1).to(table.device, torch.int64).contiguous()
        tb = table.detach().to(torch.bfloat16).contiguous()
        out = torch.empty((*ids.shape, tb.shape[1]), dtype=torch.bfloat16, device=table.device)
        lib = _lib.load()
        with torch.cuda.device(table.device):
            _lib.check(lib.zb_pc_embed(ids.data_ptr(), tb.data_ptr(), out.data_ptr(), ids.numel(), tb.shape[1], int(self.min_val), tb.shape[0], _stream(tb)))
        return out


class PassthroughConditioner(Conditioner):
    """conditioning.py:276-282."""

    def apply_cond(self, x: torch.Tensor) -> torch.Tensor:
        assert x.shape[-1] == self.cond_dim
        return x


_cond_cls_map = {
    "PassthroughConditioner": PassthroughConditioner,
    "EspeakPhonemeConditioner": EspeakPhonemeConditioner,
    "FourierConditioner": FourierConditioner,
    "IntegerConditioner": IntegerConditioner,
}


def build_conditioners(conditioners: list[dict], output_dim: int) -> list[Conditioner]:
    """conditioning.py:293-294."""
    return [_cond_cls_map[config["type"]](output_dim, **config) for config in conditioners]


class PrefixConditioner(Conditioner):
    """conditioning.py:297-313: concatenate the conditioners' outputs along the sequence, project, LayerNorm."""

    def __init__(self, config: PrefixConditionerConfig, output_dim: int):
        super().__init__(output_dim, "prefix", projection=config.projection)
        self.conditioners = nn.ModuleList(build_conditioners(config.conditioners, output_dim))
        self.norm = nn.LayerNorm(output_dim)
        self.required_keys = {c.name for c in self.conditioners if c.uncond_vector is None}

    def forward(self, cond_dict: dict) -> torch.Tensor:
        if not set(cond_dict).issuperset(self.required_keys):
            raise ValueError(f"Missing required keys: {self.required_keys - set(cond_dict)}")
        conds = []
        for conditioner in self.conditioners:
            conds.append(conditioner(cond_dict.get(conditioner.name)))
        max_bsz = max(map(len, conds))
        assert all(c.shape[0] in (max_bsz, 1) for c in conds)
        dev = self.norm.weight.device
        conds = [c.to(dev, torch.bfloat16).expand(max_bsz, -1, -1) for c in conds]
        x = _project(torch.cat(conds, dim=-2).contiguous(), self.project)
        b, s, d = x.shape
        x = x.contiguous()
        out = torch.empty_like(x)
        w = self.norm.weight.detach().to(torch.bfloat16).contiguous()
        bias = self.norm.bias.detach().to(torch.bfloat16).contiguous()
        lib = _lib.load()
        with torch.cuda.device(dev):
            _lib.check(lib.zb_pc_layernorm(x.data_ptr(), w.data_ptr(), bias.data_ptr(), float(self.norm.eps), out.data_ptr(), b * s, d, _stream(x)))
        return out


# espeak-ng language codes in the order that defines `language_id` (conditioning.py:316-328); data, not code.
supported_language_codes = (
    "af am an ar as az ba bg bn bpy bs ca cmn cs cy da de el en-029 en-gb en-gb-scotland en-gb-x-gbclan "
    "en-gb-x-gbcwmd en-gb-x-rp en-us eo es es-419 et eu fa fa-latn fi fr-be fr-ch fr-fr ga gd gn grc gu hak hi hr ht "
    "hu hy hyw ia id is it ja jbo ka kk kl kn ko kok ku ky la lfn lt lv mi mk ml mr ms mt my nb nci ne nl om or pa "
    "pap pl pt pt-br py quc ro ru ru-lv sd shn si sk sl sq sr sv sw ta te tn tr tt ur uz vi vi-vn-x-central "
    "vi-vn-x-south yue"
).split()


def make_cond_dict(text: str = "It would be nice to have time for testing, indeed.", language: str = "en-us",
                   speaker: torch.Tensor | None = None,
                   emotion: list[float] = [0.3077, 0.0256, 0.0256, 0.0256, 0.0256, 0.0256, 0.2564, 0.3077],
                   fmax: float = 22050.0, pitch_std: float = 20.0, speaking_rate: float = 15.0, vqscore_8: list[float] = [0.78] * 8,
                   ctc_loss: float = 0.0, dnsmos_ovrl: float = 4.0, speaker_noised: bool = False,
                   unconditional_keys=frozenset({"vqscore_8", "dnsmos_ovrl"}), device: torch.device | str = "cuda",
                   phoneme_ids: torch.Tensor | None = None) -> dict:
    """conditioning.py:331-405: same keyword arguments, defaults and value shapes ([1, 1, D] tensors on `device`, emotion
    normalised to sum 1).  `phoneme_ids` (not in the reference) replaces the `(texts, languages)` espeak entry with
    `("ids", ids)` for callers that phonemise elsewhere."""
    assert language.lower() in supported_language_codes, "Please pick a supported language"
    cond_dict: dict = {
        "espeak": ("ids", phoneme_ids) if phoneme_ids is not None else ([text], [language]),
        "speaker": speaker, "emotion": emotion, "fmax": fmax, "pitch_std": pitch_std, "speaking_rate": speaking_rate,
        "language_id": supported_language_codes.index(language.lower()), "vqscore_8": vqscore_8, "ctc_loss": ctc_loss,
        "dnsmos_ovrl": dnsmos_ovrl, "speaker_noised": int(speaker_noised),
    }
    for k in unconditional_keys:
        cond_dict.pop(k, None)
    for k, v in cond_dict.items():
        if k == "espeak":
            continue
        if isinstance(v, (float, int, list)):
            v = torch.tensor(v)
        if isinstance(v, torch.Tensor):
            cond_dict[k] = v.view(1, 1, -1).to(device)
        if k == "emotion":
            cond_dict[k] = cond_dict[k] / cond_dict[k].sum(dim=-1)
    return cond_dict
