"""zonos_vibes_b200 — B200-native (sm_100a) hot path for Zonos TTS generation.

Drop-in for the reference's generation path (BreakTheBeta/Zonos_Vibes): `Zonos.generate(conditioning)` and
`model.autoencoder.decode(codes)`; see INTEGRATION.md.  Python here is host plumbing over the C ABI of
libzonos_b200.so (include/zonos_b200.h); there is no CPU fallback.
"""
from .autoencoder import DACAutoencoder
from .backbone import BACKBONES, B200ZonosBackbone
from .codebook_pattern import apply_delay_pattern, revert_delay_pattern
from .conditioning import PrefixConditioner
from .config import BackboneConfig, InferenceParams, PrefixConditionerConfig, ZonosConfig
from .model import Zonos
from .sampling import sample_from_logits
from .serving import BatchingEngine

__all__ = [
    "Zonos", "DACAutoencoder", "BACKBONES", "B200ZonosBackbone", "apply_delay_pattern", "revert_delay_pattern",
    "sample_from_logits", "ZonosConfig", "BackboneConfig", "PrefixConditionerConfig", "InferenceParams", "PrefixConditioner", "BatchingEngine",
]
