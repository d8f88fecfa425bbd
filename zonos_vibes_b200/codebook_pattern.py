"""Delay pattern on the device — same signatures as the reference's zonos/codebook_pattern.py:5-12."""
from __future__ import annotations

import torch

from . import _lib


def _need_cuda(t: torch.Tensor) -> None:
    if not t.is_cuda:
        raise _lib.ZbError("zonos_vibes_b200 runs on CUDA tensors only (no CPU fallback)")


def apply_delay_pattern(codes: torch.Tensor, mask_token: int) -> torch.Tensor:
    """codes int64 [B, n_q, T] -> [B, n_q, T + n_q] (codebook_pattern.py:5-7)."""
    _need_cuda(codes)
    codes = codes.contiguous().to(torch.int64)
    b, nq, t = codes.shape
    out = torch.empty((b, nq, t + nq), dtype=torch.int64, device=codes.device)
    with torch.cuda.device(codes.device):
        _lib.check(_lib.load().zb_delay_apply(codes.data_ptr(), out.data_ptr(), b, nq, t, int(mask_token),
                                              _lib.stream_ptr(codes.device)))
    return out


def revert_delay_pattern(codes: torch.Tensor) -> torch.Tensor:
    """codes int64 [B, n_q, L] -> [B, n_q, L - n_q] (codebook_pattern.py:10-12)."""
    _need_cuda(codes)
    codes = codes.contiguous().to(torch.int64)
    b, nq, length = codes.shape
    out = torch.empty((b, nq, length - nq), dtype=torch.int64, device=codes.device)
    with torch.cuda.device(codes.device):
        _lib.check(_lib.load().zb_delay_revert(codes.data_ptr(), out.data_ptr(), b, nq, length,
                                               _lib.stream_ptr(codes.device)))
    return out
