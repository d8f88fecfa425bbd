"""Timings of the SURVEY §8 "next" rows on the device (CUDA events, best of 3): DAC encode of 10 s of audio (row f3) and the
speaker-embedding network at full depth on 10 s of 16 kHz audio (row f4).  Synthetic weights (oracle/ is used for them only)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dac_oracle as do            # noqa: E402  (synthetic weights / inputs only)
import speaker_oracle as so        # noqa: E402
import zonos_vibes_b200 as z       # noqa: E402
from zonos_vibes_b200 import speaker_cloning as sc   # noqa: E402

dev = torch.device("cuda", 0)


def best_ms(fn, n=3):
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


sd = dict(do.synth_dac_state_dict(2))
sd.update(do.synth_dac_encoder_state_dict(3))
auto = z.DACAutoencoder(sd, dev)
frames = 861
wav = do.synth_waveform(1, frames, 7).to(dev)
ms = best_ms(lambda: auto.encode(wav))
print(f"dac encode: {frames} frames (10.0 s of audio) in {ms:.1f} ms -> {0.71e9 * frames / ms / 1e9:.2f} TFLOP/s fp32 (0.71 GFLOP per frame)")

m = sc.ResNet293_based()
m.load_state_dict(so.synth_speaker_state_dict(5))
m = m.to(dev).eval()
feats = so.synth_features(1, 1000, 8).to(dev)
with torch.inference_mode():
    ms = best_ms(lambda: m.forward_features(feats))
print(f"speaker embedding: 1000 log-mel frames (10 s) through ResNet293-SimAM-ASP in {ms:.1f} ms -> {1.15e12 / ms / 1e9:.2f} TFLOP/s fp32 (~1.15 TFLOP)")
