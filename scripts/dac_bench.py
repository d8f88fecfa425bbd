"""DAC decoder timing (BASELINE config 4: 32 x 2580 frames, and one 30 s utterance), CUDA events, best of N."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import zonos_vibes_b200 as z  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--frames", type=int, default=2580)
ap.add_argument("--runs", type=int, default=5)
ap.add_argument("--opt", action="append", default=[])
a = ap.parse_args()
dev = torch.device("cuda", 0)
dac = z.DACAutoencoder(bench.random_dac_state_dict(dev), dev)
for kv in a.opt:
    k, v = kv.split("=")
    dac.set_option(k, int(v))
codes = torch.randint(0, 1024, (a.batch, 9, a.frames), generator=torch.Generator().manual_seed(2)).to(dev)
dac.decode(codes)
torch.cuda.synchronize()
best = 1e9
for r in range(a.runs):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    dac.decode(codes)
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
fl = 1.608302592e9 * a.batch * a.frames
print(f"dac B={a.batch} T={a.frames} opts={a.opt}: best {best:.2f} ms -> {fl / best / 1e9:.1f} TFLOP/s", flush=True)
