"""Row f2 measured: N concurrent /tts-style requests (full-size model, synthetic weights, 100-token conditioning, 3 s of audio each)
through serving.BatchingEngine (one batched generate + one batched DAC decode + PCM16 on the device), against the same requests
served one at a time like the reference's handler (server.py:120-152).  Wall clock around submit -> last result (host side included)."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dac_oracle as do            # noqa: E402  (synthetic weights only)
import zonos_oracle as zo          # noqa: E402
import zonos_vibes_b200 as z       # noqa: E402
from zonos_vibes_b200 import serving   # noqa: E402

dev = torch.device("cuda", 0)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 64
FRAMES = 258                       # 3 s of audio
cfg = zo.FULL_CONFIG
rd = cfg.to_reference_dict()
rd["eos_token_id"] = 1025          # fixed-length requests
m = z.Zonos.from_state_dict(z.ZonosConfig.from_dict(rd), zo.synth_state_dict(cfg, 0), dev)
m.autoencoder = z.DACAutoencoder(do.synth_dac_state_dict(2), dev)
prefixes = [zo.synth_prefix_conditioning(cfg, 1, 100, 100 + i).to(dev) for i in range(N)]
sp = dict(min_p=0.1)


def one_at_a_time(k):
    t0 = time.perf_counter()
    for p in prefixes[:k]:
        codes = m.generate(p, max_new_tokens=FRAMES, sampling_params=dict(sp), progress_bar=False)
        serving.pcm16(m.autoencoder.decode(codes)).cpu()
    return time.perf_counter() - t0


def batched():
    eng = serving.BatchingEngine(m, max_batch=N, max_wait_ms=200.0)
    t0 = time.perf_counter()
    futs = [eng.submit(p, FRAMES, sampling_params=dict(sp)) for p in prefixes]
    res = [f.result(timeout=600) for f in futs]
    for r in res:
        r.pcm.cpu()
    dt = time.perf_counter() - t0
    eng.close()
    return dt, eng.batches


one_at_a_time(2)                   # warm-up (graphs, allocations)
k = min(N, 8)
t1 = one_at_a_time(k)
batched()
tb, sizes = batched()
audio = FRAMES * 512 / 44100
print(f"one at a time: {k} requests of {audio:.2f} s audio in {t1:.2f} s = {k / t1:.2f} requests/s ({k * audio / t1:.1f} audio-s/s)")
print(f"BatchingEngine: {N} requests in {tb:.2f} s (batches {sizes}) = {N / tb:.2f} requests/s ({N * audio / tb:.1f} audio-s/s), "
      f"{(N / tb) / (k / t1):.1f} x")
