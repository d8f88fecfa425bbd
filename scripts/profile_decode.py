"""Short full-size generation used under ncu (launch list / --set full captures) and for quick timing probes."""
import argparse
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (random weights + config helpers)
import zonos_vibes_b200 as z  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=16)
ap.add_argument("--runs", type=int, default=2)
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--prefix-len", type=int, default=24, help="conditioning length (the decode steps then run at ctx >= this)")
ap.add_argument("--opt", action="append", default=[])
ap.add_argument("--dac", type=int, default=-1, help="-1: skip DAC, 0: tcgen05, 1: SIMT")
ap.add_argument("--dac-frames", type=int, default=256)
ap.add_argument("--hybrid", action="store_true", help="synthetic hybrid (Mamba2 + attention) architecture, BASELINE config 5")
ap.add_argument("--steps-per-poll", type=int, default=64, help="decode steps per native call (= per persistent-kernel launch)")
a = ap.parse_args()

dev = torch.device("cuda", 0)
cfgd = dict(backbone=dict(d_model=2048, d_intermediate=0, attn_mlp_d_intermediate=8192, n_layer=26, ssm_cfg={}, attn_layer_idx=[],
                          attn_cfg=dict(num_heads=16, num_heads_kv=4), rms_norm=False, residual_in_fp32=False, norm_epsilon=1e-5),
            prefix_conditioner=dict(conditioners=[], projection="none"), eos_token_id=1025)
if a.hybrid:
    cfgd = bench.hybrid_config_dict()
model = z.Zonos.from_state_dict(z.ZonosConfig.from_dict(cfgd), bench.random_hybrid_state_dict(dev) if a.hybrid else bench.random_backbone_state_dict(dev), dev,
                                dac_state_dict=bench.random_dac_state_dict(dev) if a.dac >= 0 else None)
ctx = model._native()
for kv in a.opt:
    k, v = kv.split("=")
    ctx.set_option(k, int(v))
prefix = torch.randn(2 * a.batch, a.prefix_len, 2048, device=dev).bfloat16()
for r in range(a.runs):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    codes = model.generate(prefix, max_new_tokens=a.frames, batch_size=a.batch, progress_bar=False, seed=r, steps_per_poll=a.steps_per_poll)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"run {r}: {a.frames} frames in {dt * 1e3:.1f} ms; device us/decode-step {model.last_step_ms * 1e3:.1f}  opts={a.opt}", flush=True)
if a.dac >= 0:
    model.autoencoder.impl = a.dac
    codes = torch.randint(0, 1024, (1, 9, a.dac_frames), device=dev)
    for r in range(a.runs):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        model.autoencoder.decode(codes)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(f"dac impl {a.dac}: {a.dac_frames} frames in {dt * 1e3:.2f} ms -> {a.dac_frames / dt:.0f} frames/s, "
              f"{1.608302592e-3 * a.dac_frames / dt:.1f} TFLOP/s", flush=True)
