"""Import the UNMODIFIED reference (BreakTheBeta/Zonos_Vibes) in THIS container.

TEST INFRASTRUCTURE ONLY.  Used by `oracle/make_golden.py` to freeze golden vectors
under `tests/golden/` and by the (non-gpu) tests that pin `oracle/zonos_oracle.py`
against the real reference when `/root/reference` is present.  `/root/reference` does
not exist on the GPU box, so nothing in the `-m gpu` tests, `smoke()` or `bench.py`
imports this module.

Recipe: the reference's `zonos/conditioning.py:60-65,171` imports
`inflect`, `kanjize`, `phonemizer`, `sudachipy` at module scope; none is installed and
none is on the hot path, so they are stubbed in `sys.modules`.  `DacModel.from_pretrained`
(`zonos/autoencoder.py:11`) needs the HF hub; it is patched to build the same 44 kHz
architecture from `DacConfig` defaults with random-init weights.
"""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("ZONOS_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "zonos", "model.py"))


def _install_stubs() -> None:
    def mod(name: str, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules.setdefault(name, m)
        return sys.modules[name]

    class _Engine:
        def number_to_words(self, *a, **k):
            return "number"

    mod("inflect", engine=lambda: _Engine())
    mod("kanjize", number2kanji=lambda n: str(n))

    class _Espeak:
        def __init__(self, *a, **k):
            pass

        def phonemize(self, texts, **k):
            return list(texts)

    ph = mod("phonemizer")
    be = mod("phonemizer.backend", EspeakBackend=_Espeak)
    ph.backend = be

    class _Tok:
        def tokenize(self, *a, **k):
            return []

    class _Dict:
        def __init__(self, *a, **k):
            pass

        def create(self, *a, **k):
            return _Tok()

    class _Split:
        A = 0
        B = 1
        C = 2

    mod("sudachipy", Dictionary=_Dict, SplitMode=_Split)


def load_reference():
    """Returns the imported reference package modules as a namespace."""
    if not reference_available():
        raise RuntimeError(f"reference not present at {REFERENCE_ROOT}")
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)

    from transformers.models.dac import DacConfig, DacModel

    if not getattr(DacModel, "_zb_patched", False):
        def _offline(cls_or_name, *a, **k):
            return DacModel(DacConfig(sampling_rate=44100))

        DacModel.from_pretrained = classmethod(lambda cls, *a, **k: _offline(cls, *a, **k))
        DacModel._zb_patched = True

    import zonos.model as zmodel
    import zonos.sampling as zsampling
    import zonos.codebook_pattern as zpattern
    import zonos.config as zconfig
    import zonos.autoencoder as zauto
    import zonos.backbone as zbackbone
    import zonos.backbone._torch as ztorch

    return types.SimpleNamespace(
        model=zmodel, sampling=zsampling, pattern=zpattern, config=zconfig,
        autoencoder=zauto, backbone=zbackbone, torch_backbone=ztorch,
    )


def build_reference_model(oracle_cfg, state_dict):
    """Instantiate the real reference `Zonos` (zonos/model.py:22-51) with the torch backbone on CPU in
    bf16 and load `state_dict` through `load_state_dict`, so the reference's own post-hook pads the heads."""
    import torch

    ref = load_reference()
    zc = ref.config.ZonosConfig.from_dict(oracle_cfg.to_reference_dict())
    model = ref.model.Zonos(zc, ref.backbone.BACKBONES["torch"]).to("cpu", torch.bfloat16).eval()
    sd = model.state_dict()
    for k, v in state_dict.items():
        assert k in sd and sd[k].shape == v.shape, (k, v.shape)
        sd[k] = v
    model.load_state_dict(sd)
    return ref, model
