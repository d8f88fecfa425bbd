"""Freeze the speaker-embedding golden vectors from the REAL reference classes (zonos/speaker_cloning.py) into
tests/golden/golden_spk.pt: the reference's ResNet293_based (full depth: 10 + 20 + 64 + 3 SimAM blocks) and LDA projection,
loaded with oracle.speaker_oracle's synthetic state_dict, on synthetic log-mel features.  Run in the build container only.

    python oracle/make_golden_spk.py
"""
import os
import sys
import time

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import speaker_oracle as so                       # noqa: E402
from ref_loader import load_reference           # noqa: E402

OUT = os.path.join(HERE, "..", "tests", "golden", "golden_spk.pt")


@torch.inference_mode()
def main():
    load_reference()                              # puts /root/reference on the path with the stubs it needs
    import zonos.speaker_cloning as sc
    g = {}
    for name, nb, frames in (("full", so.NUM_BLOCKS, 48), ("shallow", (2, 2, 3, 1), 101)):
        sd = so.synth_speaker_state_dict(5, nb)
        m = sc.ResNet293_based()
        if nb != so.NUM_BLOCKS:
            m.front = sc.ResNet(so.IN_PLANES, sc.SimAMBasicBlock, list(nb))
        m.load_state_dict(sd)
        m.eval()
        feats = so.synth_features(2, frames, 8)
        t0 = time.time()
        x = m.front(feats.unsqueeze(1))
        emb = m.bottleneck(m.pooling(x))
        lda = so.synth_lda_state_dict(6)
        want_o, want_lda = so.embedding_from_features(sd, feats, nb, lda)
        print(name, "reference forward", round(time.time() - t0, 1), "s; oracle vs reference max abs", (emb - want_o).abs().max().item(),
              "emb rms", emb.pow(2).mean().sqrt().item(), "front rms", x.pow(2).mean().sqrt().item())
        g[f"{name}.num_blocks"] = tuple(nb)
        g[f"{name}.frames"] = frames
        g[f"{name}.embedding"] = emb
        g[f"{name}.lda"] = torch.nn.functional.linear(emb, lda["weight"], lda["bias"])
        g[f"{name}.front_rms"] = x.pow(2).mean().sqrt()
    torch.save(g, OUT)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
