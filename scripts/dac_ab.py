"""Interleaved A/B timing of DAC decoder tuning knobs on one box (box-to-box clock noise is ~10 %, so variants must share
a process).  Usage: python scripts/dac_ab.py --frames 2580 --reps 7 base ew_multi=8 ew_k1=16,ew_multi=8 impl=2 ..."""
import argparse, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from zonos_vibes_b200.autoencoder import DACAutoencoder

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=2580)
ap.add_argument("--reps", type=int, default=7)
ap.add_argument("--batch", type=int, default=1, help="utterances per call (BASELINE config 4 = 32 x 2580 frames)")
ap.add_argument("variants", nargs="*", default=["base"])
a = ap.parse_args()
dev = torch.device("cuda:0")
dac = DACAutoencoder(bench.random_dac_state_dict(dev), dev)
codes = torch.randint(0, 1024, (a.batch, 9, a.frames), device=dev)
DEFAULTS = {"ew_multi": 8, "ew_k1": 8, "wide_cout": 768, "mt2_min_waves": 8, "impl": 0}
times = {v: [] for v in a.variants}
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for rep in range(a.reps + 1):
    for v in a.variants:
        opts = dict(DEFAULTS)
        if v != "base":
            for kv in v.split(","):
                k, x = kv.split("=")
                opts[k] = int(x)
        for k, x in opts.items():
            if k == "impl":
                dac.impl = x
            else:
                dac.set_option(k, x)
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dac.decode(codes)
        e1.record()
        torch.cuda.synchronize()
        if rep:
            times[v].append(e0.elapsed_time(e1))
for v in a.variants:
    t = times[v]
    print(f"{v:40s} min {min(t):.3f} ms  median {statistics.median(t):.3f} ms  -> {1.608302592e-3 * a.frames * a.batch / (min(t) * 1e-3):.0f} TFLOP/s (best), {a.batch * a.frames / (min(t) * 1e-3):.0f} frames/s", flush=True)
