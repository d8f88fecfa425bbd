"""ctypes binding of libzonos_b200.so (the C ABI declared in include/zonos_b200.h).

There is no CPU fallback: if the shared library is missing and cannot be built, or no sm_100 device is
present when a context is created, the import / call fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libzonos_b200.so")

ZB_BF16, ZB_F32, ZB_I64, ZB_I32 = 0, 1, 2, 3


class ZbConfig(C.Structure):
    _fields_ = [
        ("d_model", C.c_int32), ("n_layer", C.c_int32), ("num_heads", C.c_int32), ("num_heads_kv", C.c_int32),
        ("ffn", C.c_int32), ("n_codebooks", C.c_int32), ("vocab", C.c_int32), ("eos_token_id", C.c_int32),
        ("masked_token_id", C.c_int32), ("norm_eps", C.c_float), ("rope_table_len", C.c_int32),
        ("reserved", C.c_int32 * 5),
    ]


class ZbHybrid(C.Structure):
    _fields_ = [
        ("d_state", C.c_int32), ("d_conv", C.c_int32), ("expand", C.c_int32), ("headdim", C.c_int32), ("ngroups", C.c_int32),
        ("n_attn_layers", C.c_int32), ("attn_layer_idx", C.c_int32 * 64), ("reserved", C.c_int32 * 4),
    ]


class ZbSampling(C.Structure):
    _fields_ = [
        ("temperature", C.c_float), ("top_p", C.c_float), ("top_k", C.c_int32), ("min_p", C.c_float),
        ("linear", C.c_float), ("conf", C.c_float), ("quad", C.c_float), ("repetition_penalty", C.c_float),
        ("repetition_penalty_window", C.c_int32), ("reserved", C.c_int32 * 3),
    ]


# name -> (restype, argtypes); exactly the entry points of include/zonos_b200.h
_P, _I, _I64, _U64, _F = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_float
SIGNATURES = {
    "zb_abi_version": (_I, []),
    "zb_ctx_create": (_I, [C.POINTER(ZbConfig), _I, C.POINTER(_P)]),
    "zb_ctx_create_hybrid": (_I, [C.POINTER(ZbConfig), C.POINTER(ZbHybrid), _I, C.POINTER(_P)]),
    "zb_ctx_destroy": (_I, [_P]),
    "zb_last_error": (C.c_char_p, [_P]),
    "zb_set_tensor": (_I, [_P, C.c_char_p, _P, _I64, _I, _P]),
    "zb_weights_finalize": (_I, [_P, _P]),
    "zb_kv_alloc": (_I, [_P, _I, _I]),
    "zb_kv_reset": (_I, [_P, _P]),
    "zb_backbone_forward": (_I, [_P, _P, _P, _I, _I, _I, _P, _P]),
    "zb_delay_apply": (_I, [_P, _P, _I, _I, _I, _I64, _P]),
    "zb_delay_revert": (_I, [_P, _P, _I, _I, _I, _P]),
    "zb_sample": (_I, [_P, _P, _I, _P, _P, _I, _I, _I, C.POINTER(ZbSampling), _U64, _U64, _P]),
    "zb_decode_logits": (_I, [_P, _P, _I, _I, _I, _F, _I, _P, _P]),
    "zb_gen_begin": (_I, [_P, _P, _I, _I, _P, _I, _I, _F, C.POINTER(ZbSampling), _U64, _I, _P]),
    "zb_gen_steps": (_I, [_P, _I, _P]),
    "zb_gen_poll": (_I, [_P, C.POINTER(_I), C.POINTER(_I), _P, _P]),
    "zb_update_frame": (_I, [_P, _P, _I, _I, _I, _P, _P, _P, _P, _I, _I, _P]),
    "zb_dac_create": (_I, [_I, C.POINTER(_P)]),
    "zb_dac_destroy": (_I, [_P]),
    "zb_dac_last_error": (C.c_char_p, [_P]),
    "zb_dac_set_tensor": (_I, [_P, C.c_char_p, _P, _I64, _P]),
    "zb_dac_finalize": (_I, [_P, _P]),
    "zb_dac_decode": (_I, [_P, _P, _P, _I, _I, _I, _P]),
    "zb_dac_decode_static": (_I, [_P, _P, _P, _I, _I, _P]),
    "zb_dac_get_intermediate": (_I, [_P, C.c_char_p, _P, _I64, _P]),
    "zb_dac_set_option": (_I, [_P, C.c_char_p, _I]),
    "zb_dac_launch_count": (_I64, [_P]),
    "zb_dac_encode": (_I, [_P, _P, _P, _I, _I64, _P]),
    "zb_spk_conv2d": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _P]),
    "zb_spk_simam_add_relu": (_I, [_P, _P, _P, _I, _I, _I, _I, _F, _P]),
    "zb_spk_asp": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "zb_spk_linear": (_I, [_P, _P, _P, _P, _I, _I, _I, _P]),
    "zb_wav_to_pcm16": (_I, [_P, _P, _I64, _P]),
    "zb_pc_fourier": (_I, [_P, _P, _F, _F, _P, _I, _I, _I, _P]),
    "zb_pc_embed": (_I, [_P, _P, _P, _I, _I, _I64, _I64, _P]),
    "zb_pc_linear": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "zb_pc_layernorm": (_I, [_P, _P, _P, _F, _P, _I, _I, _P]),
    "zb_weight_bytes": (_I64, [_P]),
    "zb_launch_count": (_I64, [_P]),
    "zb_set_option": (_I, [_P, C.c_char_p, _I]),
    "zb_bench_kernel": (_I, [_P, C.c_char_p, _I, C.POINTER(_F), _P]),
    "zb_last_steps_ms": (_F, [_P]),
    "zb_last_trace": (_I, [_P, _P, _I]),
}

_lib = None


class ZbError(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load (building first if the sources are newer and nvcc is present) the native library."""
    global _lib
    if _lib is not None:
        return _lib
    override = os.environ.get("ZB_LIB")
    if override:
        # a library built some other way (scripts/build_checked.sh: the -DZB_BOUNDS build with index assertions); same ABI checks
        lib = C.CDLL(override)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        if lib.zb_abi_version() != 1:
            raise ZbError(f"{override}: ABI version mismatch")
        _lib = lib
        return lib
    from . import build as _build
    if _build._stale():
        # missing, or built from other sources than the ones next to it (content hash, build.py): rebuild when nvcc is
        # there; a stale library without a compiler is used as it is, loudly
        try:
            _build.build()
        except Exception as e:  # noqa: BLE001
            if not os.path.exists(LIB_PATH):
                raise ZbError(
                    f"libzonos_b200.so is missing and could not be built ({e}); the B200 path has no fallback") from e
            import warnings
            warnings.warn(f"libzonos_b200.so does not match csrc/ and could not be rebuilt ({e}); using it as it is")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the ABI lost a symbol
        fn.restype = res
        fn.argtypes = args
    if lib.zb_abi_version() != 1:
        raise ZbError("libzonos_b200.so ABI version mismatch")
    _lib = lib
    return lib


def check(rc: int, ctx=None, dac: bool = False) -> None:
    if rc == 0:
        return
    lib = load()
    msg = (lib.zb_dac_last_error(ctx) if dac else lib.zb_last_error(ctx)) or b""
    raise ZbError(f"libzonos_b200 error {rc}: {msg.decode(errors='replace')}")


def stream_ptr(device=None) -> int:
    import torch

    return torch.cuda.current_stream(device).cuda_stream


def make_sampling(temperature=1.0, top_p=0.0, top_k=0, min_p=0.0, linear=0.0, conf=0.0, quad=0.0,
                  repetition_penalty=3.0, repetition_penalty_window=2) -> ZbSampling:
    """kwargs of the reference's sample_from_logits (zonos/sampling.py:117-129), same defaults."""
    s = ZbSampling()
    s.temperature, s.top_p, s.top_k, s.min_p = float(temperature), float(top_p), int(top_k), float(min_p)
    s.linear, s.conf, s.quad = float(linear), float(conf), float(quad)
    s.repetition_penalty, s.repetition_penalty_window = float(repetition_penalty), int(repetition_penalty_window)
    return s
