"""GPU: the allocation-free C-ABI entry point `zb_dac_decode_static` (include/zonos_b200.h) against the existing
`zb_dac_decode`, through ctypes, on two calls in a row.  Collected last on purpose (file name):
a failure here must not hide the parity tests.  Passes alone in a fresh process and after the others.
"""
import ctypes as C

import pytest
import torch

import dac_oracle as do

pytestmark = pytest.mark.gpu

MAX_FRAMES = 48          # ZB_DAC_STATIC_MAX_FRAMES in include/zonos_b200.h
TOL = 1e-2               # the DAC tolerance of tests/test_gpu_parity.py (max-abs on identical codes)


def test_dac_decode_static_matches_dac_decode(cuda):
    import zonos_vibes_b200 as z
    from zonos_vibes_b200 import _lib

    dac = z.DACAutoencoder(do.synth_dac_state_dict(2), cuda)
    lib, h = dac.lib, dac.handle
    stream = torch.cuda.current_stream(cuda).cuda_stream
    assert lib.zb_dac_decode_static(h, None, None, 0, 0, stream) == 0          # set-up call: resolves the symbol address
    gen = torch.Generator().manual_seed(77)
    prev = None
    for call, (b, t) in enumerate([(2, 40), (1, MAX_FRAMES)]):                 # two calls in a row share the static scratch
        codes = torch.randint(0, 1024, (b, 9, t), generator=gen).to(cuda)
        want = torch.empty((b, 1, 512 * t), dtype=torch.float32, device=cuda)
        got = torch.full_like(want, float("nan"))
        _lib.check(lib.zb_dac_decode(h, codes.data_ptr(), want.data_ptr(), b, t, 0, stream), h, dac=True)
        torch.cuda.synchronize(cuda)
        free0, _ = torch.cuda.mem_get_info(cuda)
        _lib.check(lib.zb_dac_decode_static(h, codes.data_ptr(), got.data_ptr(), b, t, stream), h, dac=True)
        torch.cuda.synchronize(cuda)
        free1, _ = torch.cuda.mem_get_info(cuda)
        assert free1 >= free0, f"call {call}: device memory shrank by {free0 - free1} B inside the no-alloc entry point"
        assert bool(torch.isfinite(got).all())
        err = (got - want).abs().max().item()
        assert err <= TOL, f"call {call}: max-abs {err} vs zb_dac_decode"
        ref = do.dac_decode(do.synth_dac_state_dict(2), codes[:1].cpu())      # and against the oracle itself
        assert (got[:1].cpu() - ref).abs().max().item() <= TOL
        if prev is not None:
            assert got.shape != prev.shape                                     # the second call really reused the array
        prev = got
    # sizes beyond the static maximum are rejected with a clear error, and nothing is launched
    t = MAX_FRAMES + 1
    codes = torch.zeros((1, 9, t), dtype=torch.int64, device=cuda)
    out = torch.zeros((1, 1, 512 * t), dtype=torch.float32, device=cuda)
    rc = lib.zb_dac_decode_static(h, codes.data_ptr(), out.data_ptr(), 1, t, stream)
    assert rc == -1
    msg = lib.zb_dac_last_error(h).decode()
    assert "exceeds the static workspace" in msg and str(MAX_FRAMES) in msg
    torch.cuda.synchronize(cuda)
    assert float(out.abs().max()) == 0.0
