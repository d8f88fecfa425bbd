"""Fused sampler — same call signature as the reference's zonos/sampling.py:117-182 `sample_from_logits`."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


def sample_from_logits(
    logits: torch.Tensor,
    temperature: float = 1.0,
    top_p: float = 0.0,
    top_k: int = 0,
    min_p: float = 0.0,
    linear: float = 0.0,
    conf: float = 0.0,
    quad: float = 0.0,
    generated_tokens: torch.Tensor | None = None,
    repetition_penalty: float = 3.0,
    repetition_penalty_window: int = 2,
    *,
    seed: int | None = None,
    step: int = 0,
    return_probs: bool = False,
):
    """logits fp32 [B, n_q, V] on CUDA -> tokens int64 [B, n_q, 1].

    Repetition penalty, temperature softmax, unified sampler, top-p, top-k, min-p and the exponential-race draw run
    in ONE kernel launch (one CTA per (b, codebook)).  The draw uses a Philox4x32-10 stream keyed by (seed, step)
    instead of torch's global generator; `seed=None` derives one from torch's CPU generator so
    that `torch.manual_seed` still makes runs reproducible.  `return_probs=True` also returns the filtered,
    renormalised distribution (or, for temperature == 0, the penalised logits) for mask-parity checks.
    """
    if not logits.is_cuda:
        raise _lib.ZbError("zonos_vibes_b200 runs on CUDA tensors only (no CPU fallback)")
    lib = _lib.load()
    logits = logits.contiguous().float()
    b, nq, v = logits.shape
    hist_ptr, hist_len = None, 0
    if generated_tokens is not None:
        generated_tokens = generated_tokens.contiguous().to(torch.int64)
        hist_ptr, hist_len = generated_tokens.data_ptr(), generated_tokens.shape[-1]
    if seed is None:
        seed = int(torch.randint(0, 2**62, (1,)).item())
    tokens = torch.empty((b, nq), dtype=torch.int64, device=logits.device)
    probs = torch.empty_like(logits) if return_probs else None
    sp = _lib.make_sampling(temperature, top_p, top_k, min_p, linear, conf, quad, repetition_penalty,
                            repetition_penalty_window)
    with torch.cuda.device(logits.device):
        _lib.check(lib.zb_sample(logits.data_ptr(), hist_ptr, hist_len, tokens.data_ptr(),
                                 probs.data_ptr() if probs is not None else None, b, nq, v, C.byref(sp),
                                 int(seed), int(step), _lib.stream_ptr(logits.device)))
    tokens = tokens.unsqueeze(-1)
    return (tokens, probs) if return_probs else tokens
