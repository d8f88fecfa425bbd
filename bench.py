#!/usr/bin/env python
"""bench.py — headline benchmark of the Zonos generation hot path on B200 (contract: see the task's bench spec).

Workload (BASELINE.json configs[1]): Zonos-v0.1-transformer (26 layers, d=2048, 1.62 B params, random-init bf16),
batch 1 per GPU, synthetic conditioning [2, 24, 2048], 30 s of audio = 2580 frames, min-p 0.1 sampling, EOS disabled
(eos_token_id=1025) so every run has the same length, then DAC 44 kHz decode of the 2580 frames.
One "step" = one whole utterance: prefill + 2588 decode iterations + DAC decode -> 29.95 s of waveform.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch B] [--frames T]

N > 1 is launched by torchrun (one rank per GPU); every rank runs its own utterances (weak scaling, no data-path
collective); time = max over ranks; rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAME_RATE = 44100 / 512          # 86.1328 frames per second of audio
LC = 24                           # synthetic conditioning length
FULL = dict(d_model=2048, n_layer=26, num_heads=16, num_heads_kv=4, ffn=8192)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=1, help="utterances per GPU")
    ap.add_argument("--frames", type=int, default=86 * 30)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--dac-impl", type=int, default=0, help="0 = tcgen05 fp16 (default), 1 = SIMT fp32 cross-check path")
    ap.add_argument("--opt", action="append", default=[], help="native option key=value (pdl, l2_prefetch, graph)")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------------------------
# helpers
# ----------------------------------------------------------------------------------------------------------------
def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d["bf16_tflops"]), "measured"
    return 6650.0, 1590.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = "nccl" if (args.impl == "b200" and torch.cuda.is_available()) else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist.init_process_group(backend)
    return world, rank, local


def barrier(world, device=None):
    if world > 1:
        import torch.distributed as dist
        if device is not None and device.type == "cuda":
            dist.barrier(device_ids=[device.index])
        else:
            dist.barrier()


# ----------------------------------------------------------------------------------------------------------------
# synthetic random-init weights (product arm: generated on the device; no oracle import here)
# ----------------------------------------------------------------------------------------------------------------
def random_backbone_state_dict(dev, seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    d, f, nl = FULL["d_model"], FULL["ffn"], FULL["n_layer"]
    qkv = (FULL["num_heads"] + 2 * FULL["num_heads_kv"]) * 128

    def lin(o, i):
        return ((torch.rand(o, i, generator=g, device=dev) * 2 - 1) / math.sqrt(i)).bfloat16()

    def ln(sd, p):
        sd[p + ".weight"] = (1 + 0.1 * torch.randn(d, generator=g, device=dev)).bfloat16()
        sd[p + ".bias"] = (0.1 * torch.randn(d, generator=g, device=dev)).bfloat16()

    sd = {}
    for i in range(nl):
        p = f"backbone.layers.{i}"
        ln(sd, p + ".norm")
        sd[p + ".mixer.in_proj.weight"] = lin(qkv, d)
        sd[p + ".mixer.out_proj.weight"] = lin(d, d)
        ln(sd, p + ".norm2")
        sd[p + ".mlp.fc1.weight"] = lin(2 * f, d)
        sd[p + ".mlp.fc2.weight"] = lin(d, f)
    ln(sd, "backbone.norm_f")
    for k in range(9):
        sd[f"embeddings.{k}.weight"] = torch.randn(1026, d, generator=g, device=dev).bfloat16()
        sd[f"heads.{k}.weight"] = lin(1025, d)
    return sd


HYBRID = dict(n_layer=46, attn_layer_idx=[9, 18, 27, 36, 45], d_state=128, d_conv=4, expand=2, headdim=64, ngroups=1)
"""Synthetic hybrid architecture (BASELINE config 5).  The checkpoint's config.json is not in the reference repo and
mamba-ssm is not installable here (SURVEY App. C: dims UNVERIFIED): d_model 2048, Mamba2 defaults, 46 layers of which 5 are
attention + GatedMLP(8192) blocks."""


def hybrid_config_dict():
    return dict(backbone=dict(d_model=FULL["d_model"], d_intermediate=0, attn_mlp_d_intermediate=FULL["ffn"], n_layer=HYBRID["n_layer"],
                              ssm_cfg=dict(layer="Mamba2"), attn_layer_idx=HYBRID["attn_layer_idx"],
                              attn_cfg=dict(num_heads=FULL["num_heads"], num_heads_kv=FULL["num_heads_kv"]),
                              rms_norm=False, residual_in_fp32=False, norm_epsilon=1e-5),
                prefix_conditioner=dict(conditioners=[], projection="none"), eos_token_id=1025)


def random_hybrid_state_dict(dev, seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    d, f = FULL["d_model"], FULL["ffn"]
    qkv = (FULL["num_heads"] + 2 * FULL["num_heads_kv"]) * 128
    di = HYBRID["expand"] * d
    nh = di // HYBRID["headdim"]
    conv_dim = di + 2 * HYBRID["ngroups"] * HYBRID["d_state"]
    d_in_proj = 2 * di + 2 * HYBRID["ngroups"] * HYBRID["d_state"] + nh

    def lin(o, i):
        return ((torch.rand(o, i, generator=g, device=dev) * 2 - 1) / math.sqrt(i)).bfloat16()

    def ln(sd, p):
        sd[p + ".weight"] = (1 + 0.1 * torch.randn(d, generator=g, device=dev)).bfloat16()
        sd[p + ".bias"] = (0.1 * torch.randn(d, generator=g, device=dev)).bfloat16()

    sd = {}
    for i in range(HYBRID["n_layer"]):
        p = f"backbone.layers.{i}"
        ln(sd, p + ".norm")
        if i in HYBRID["attn_layer_idx"]:
            sd[p + ".mixer.in_proj.weight"] = lin(qkv, d)
            sd[p + ".mixer.out_proj.weight"] = lin(d, d)
            ln(sd, p + ".norm2")
            sd[p + ".mlp.fc1.weight"] = lin(2 * f, d)
            sd[p + ".mlp.fc2.weight"] = lin(d, f)
        else:
            sd[p + ".mixer.in_proj.weight"] = lin(d_in_proj, d)
            sd[p + ".mixer.conv1d.weight"] = ((torch.rand(conv_dim, 1, HYBRID["d_conv"], generator=g, device=dev) * 2 - 1) / 2).bfloat16()
            sd[p + ".mixer.conv1d.bias"] = (0.1 * torch.randn(conv_dim, generator=g, device=dev)).bfloat16()
            dt = torch.exp(torch.rand(nh, generator=g, device=dev) * (math.log(0.1) - math.log(0.001)) + math.log(0.001))
            sd[p + ".mixer.dt_bias"] = (dt + torch.log(-torch.expm1(-dt))).bfloat16()
            sd[p + ".mixer.A_log"] = torch.log(1 + 15 * torch.rand(nh, generator=g, device=dev)).bfloat16()
            sd[p + ".mixer.D"] = torch.ones(nh, device=dev).bfloat16()
            sd[p + ".mixer.norm.weight"] = (1 + 0.1 * torch.randn(di, generator=g, device=dev)).bfloat16()
            sd[p + ".mixer.out_proj.weight"] = lin(d, di)
    ln(sd, "backbone.norm_f")
    for k in range(9):
        sd[f"embeddings.{k}.weight"] = torch.randn(1026, d, generator=g, device=dev).bfloat16()
        sd[f"heads.{k}.weight"] = lin(1025, d)
    return sd


def random_dac_state_dict(dev, seed=2):
    g = torch.Generator(device=dev).manual_seed(seed)
    sd = {}

    def conv(p, cout, cin, k, gain=1.0, transposed=False):
        shape, fan = ((cin, cout, k), cin * 2) if transposed else ((cout, cin, k), cin * k)
        sd[p + ".weight"] = (torch.rand(shape, generator=g, device=dev) * 2 - 1) * gain * math.sqrt(3.0 / fan)
        sd[p + ".bias"] = 0.05 * torch.randn(cout, generator=g, device=dev)

    for k in range(9):
        q = f"quantizer.quantizers.{k}"
        sd[q + ".codebook.weight"] = torch.randn(1024, 8, generator=g, device=dev)
        conv(q + ".out_proj", 1024, 8, 1, gain=1 / 3)
    conv("decoder.conv1", 1536, 1024, 7)
    c = 1536
    for j, s in enumerate((8, 8, 4, 2)):
        p = f"decoder.block.{j}"
        sd[p + ".snake1.alpha"] = 0.5 + torch.rand(1, c, 1, generator=g, device=dev)
        conv(p + ".conv_t1", c // 2, c, 2 * s, transposed=True)
        c //= 2
        for u in (1, 2, 3):
            r = f"{p}.res_unit{u}"
            sd[r + ".snake1.alpha"] = 0.5 + torch.rand(1, c, 1, generator=g, device=dev)
            conv(r + ".conv1", c, c, 7, gain=0.7)
            sd[r + ".snake2.alpha"] = 0.5 + torch.rand(1, c, 1, generator=g, device=dev)
            conv(r + ".conv2", c, c, 1, gain=0.5)
    sd["decoder.snake1.alpha"] = 0.5 + torch.rand(1, c, 1, generator=g, device=dev)
    conv("decoder.conv2", 1, c, 7, gain=0.12)
    return sd


# ----------------------------------------------------------------------------------------------------------------
# CPU baseline (the oracle port of the reference's PyTorch path) — the ONLY place bench.py touches oracle/
# ----------------------------------------------------------------------------------------------------------------
def ncu_traffic_per_step(args):
    """dram__bytes_read.sum + dram__bytes_write.sum of the persistent decode kernel, per decode step, from the committed
    `ncu --set full` capture (profiles/r01_traffic.json); None when the run is not the configuration that was captured."""
    if args.batch != 1 or "mega=0" in args.opt:
        return None
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            return float(json.load(f)["dram_bytes_per_step"])
    except Exception:
        return None


def cpu_reference_sample(n_decode_steps=None, dac_frames=43, budget_s=12.0):
    """Times the reference algorithm (oracle port, torch CPU, all host threads) on a bounded sample of the workload:
    full-size model, batch 1 + CFG, prefill of Lc+1 tokens then decode steps at the start of the utterance, and a DAC
    decode of `dac_frames` frames.  Returns per-unit rates; no GPU involved."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import dac_oracle as do
    import zonos_oracle as zo

    # all host threads this process may use (torchrun exports OMP_NUM_THREADS=1, which would make this a 1-thread baseline)
    try:
        ncpu = len(os.sched_getaffinity(0))
    except AttributeError:
        ncpu = os.cpu_count() or 1
    torch.set_num_threads(max(1, ncpu))
    cfg = zo.OracleConfig(eos_token_id=1025)
    sd = zo.synth_state_dict(cfg, 0)
    o = zo.OracleZonos(cfg, sd)
    prefix = zo.synth_prefix_conditioning(cfg, 1, LC, 1)
    codes = torch.randint(0, 1024, (1, 9, 64), generator=torch.Generator().manual_seed(3))
    delayed = zo.apply_delay_pattern(codes, 1025)
    with torch.inference_mode():
        params = o.setup_cache(2, LC + 128)
        t0 = time.perf_counter()
        o.prefill(prefix, delayed[..., :1], params, 2.0)
        t_prefill = time.perf_counter() - t0
        params.seqlen_offset += LC + 1
        params.lengths_per_sample[:] += LC + 1
        times = []
        s = 0
        while s < 60 and (n_decode_steps is None or s < n_decode_steps):
            t0 = time.perf_counter()
            logits = o.decode_one_token(delayed[..., s + 1: s + 2], params, 2.0)
            zo.sample_from_logits(logits, min_p=0.1, generated_tokens=delayed[..., : s + 2])
            times.append(time.perf_counter() - t0)
            params.seqlen_offset += 1
            params.lengths_per_sample[:] += 1
            s += 1
            if n_decode_steps is None and sum(times) > budget_s and s >= 4:
                break
        step_s = sorted(times)[len(times) // 2]
        dsd = do.synth_dac_state_dict(2)
        dcodes = torch.randint(0, 1024, (1, 9, dac_frames), generator=torch.Generator().manual_seed(4))
        t0 = time.perf_counter()
        do.dac_decode(dsd, dcodes)
        t_dac = time.perf_counter() - t0
    dac_s_per_frame = t_dac / dac_frames
    s_per_frame = step_s + dac_s_per_frame
    return {
        "frames_per_s_decode": 1.0 / step_s, "ms_per_decode_step": step_s * 1e3, "prefill_s": t_prefill,
        "dac_frames_per_s": 1.0 / dac_s_per_frame, "audio_sec_per_sec": 1.0 / (s_per_frame * FRAME_RATE),
        "decode_steps_timed": len(times), "dac_frames_timed": dac_frames, "cores": torch.get_num_threads(),
    }


def run_reference(args, world, rank):
    """--impl reference: the reference's CPU implementation of the path (oracle port; the reference is a Python
    package that cannot travel to the GPU box), all host threads, bounded sample per step."""
    if rank != 0:
        return
    vals = []
    for i in range(args.warmup + args.steps):
        r = cpu_reference_sample(n_decode_steps=6, dac_frames=22)
        if i >= args.warmup:
            vals.append(r)
    v = sum(x["audio_sec_per_sec"] for x in vals) / len(vals)
    ms = sum(x["ms_per_decode_step"] for x in vals) / len(vals)
    sample = "full-size model B=1+CFG: prefill(25 tok) + 6 decode steps + DAC decode of 22 frames per step; per-frame rates extrapolated"
    line = {
        "impl": "reference", "metric": "audio_sec_per_sec", "value": v, "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": (args.frames / FRAME_RATE) / v * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "Zonos-v0.1-transformer bf16 batch 1, 30 s audio (2580 frames), min-p 0.1, + DAC 44 kHz decode",
                   "frames": args.frames, "batch_per_gpu": 1},
        "decode_frames_per_s": sum(x["frames_per_s_decode"] for x in vals) / len(vals), "us_per_decode_step": ms * 1e3,
        "cpu_baseline": {"value": v, "unit": "audio-s/s", "cores": vals[0]["cores"], "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------
# B200 arm
# ----------------------------------------------------------------------------------------------------------------
def run_b200(args, world, rank, local):
    import zonos_vibes_b200 as z

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a B200; there is no CPU fallback")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    B, T = args.batch, args.frames
    cfgd = dict(backbone=dict(d_model=FULL["d_model"], d_intermediate=0, attn_mlp_d_intermediate=FULL["ffn"], n_layer=FULL["n_layer"],
                              ssm_cfg={}, attn_layer_idx=[], attn_cfg=dict(num_heads=FULL["num_heads"], num_heads_kv=FULL["num_heads_kv"]),
                              rms_norm=False, residual_in_fp32=False, norm_epsilon=1e-5),
                prefix_conditioner=dict(conditioners=[], projection="none"), eos_token_id=1025)
    model = z.Zonos.from_state_dict(z.ZonosConfig.from_dict(cfgd), random_backbone_state_dict(dev), dev,
                                    dac_state_dict=random_dac_state_dict(dev))
    model.autoencoder.impl = args.dac_impl
    ctx = model._native()
    for kv in args.opt:
        k, v = kv.split("=")
        ctx.set_option(k, int(v))
    g = torch.Generator(device="cpu").manual_seed(100 + rank)
    prefix_host = torch.randn(2 * B, LC, FULL["d_model"], generator=g).bfloat16().pin_memory()
    prefix_dev = prefix_host.to(dev)
    wav_host = torch.empty((B, 1, 512 * T), dtype=torch.float32).pin_memory()
    sp = dict(min_p=0.1)

    def one_pass(e2e: bool, seed: int):
        p = prefix_host.to(dev, non_blocking=True) if e2e else prefix_dev
        codes = model.generate(p, max_new_tokens=T, batch_size=B, sampling_params=sp, progress_bar=False, seed=seed)
        wav = model.autoencoder.decode(codes)
        if e2e:
            wav_host[..., : wav.shape[-1]].copy_(wav, non_blocking=True)
        return codes, wav

    for i in range(args.warmup):
        one_pass(False, i)
    torch.cuda.synchronize()

    def timed(e2e: bool):
        launches0 = ctx.lib.zb_launch_count(ctx.handle) + model.autoencoder.launch_count()
        step_ms = []
        barrier(world, dev)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        frames = 0
        for i in range(args.steps):
            codes, _ = one_pass(e2e, 1000 + i)
            frames += codes.shape[0] * codes.shape[-1]
            step_ms.append(model.last_step_ms)
        e1.record()
        torch.cuda.synchronize()
        barrier(world, dev)
        ms = e0.elapsed_time(e1)
        launches = ctx.lib.zb_launch_count(ctx.handle) + model.autoencoder.launch_count() - launches0
        return ms, frames, sum(step_ms) / len(step_ms), launches

    sampler = ClockSampler(local) if rank == 0 else None
    ms_dev, frames_dev, step_ms_dev, launches = timed(False)
    ms_e2e, frames_e2e, _, _ = timed(True)
    clocks = sampler.stop() if sampler else None

    # isolated timing of the dominant kernel (fc1 GEMV: 67 MB of weights per launch) for the roofline object
    import ctypes as C
    ms_k = C.c_float(0)
    iters = 26 * 8
    ctx.check(ctx.lib.zb_bench_kernel(ctx.handle, b"fc1", 26, C.byref(ms_k), ctx.stream))       # warm-up
    ctx.check(ctx.lib.zb_bench_kernel(ctx.handle, b"fc1", iters, C.byref(ms_k), ctx.stream))
    fc1_bytes = 2 * FULL["ffn"] * FULL["d_model"] * 2
    fc1_gbs = fc1_bytes * iters / (ms_k.value * 1e-3) / 1e9

    from zonos_vibes_b200.replicas import max_over_ranks
    ms_dev = max_over_ranks(ms_dev, dev)
    ms_e2e = max_over_ranks(ms_e2e, dev)
    if rank != 0:
        return
    hbm_peak, _, peak_kind = measured_peaks()
    total_frames = frames_dev * world
    audio_s = total_frames / FRAME_RATE
    value = audio_s / (ms_dev * 1e-3)
    e2e_value = (frames_e2e * world / FRAME_RATE) / (ms_e2e * 1e-3)
    n_dec = T + 8
    mean_ctx = LC + 1 + n_dec / 2.0
    step_bytes = ctx.lib.zb_weight_bytes(ctx.handle) + 106496 * B * (mean_ctx + 1) + 36864 * B
    step_gbs = step_bytes / (step_ms_dev * 1e-3) / 1e9
    body = "embed + 26x[ln+qkv+rope+kv-append, split-kv attention, out_proj+resid, ln+fc1+swiglu, fc2+resid] + ln+heads+cfg + sample + state machine"
    if B * 2 > 8:
        kernel_name = f"decode step, batched tcgen05 path (one CUDA graph per step: {body})"
    elif "mega=0" in args.opt:
        kernel_name = f"decode step, per-op path (one CUDA graph per step: {body})"
    else:
        kernel_name = f"k_decode_mega<4> (persistent cooperative kernel, 64 decode steps per launch; bytes and time are per step: {body})"
    line = {
        "metric": "audio_sec_per_sec", "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": "Zonos-v0.1-transformer bf16 batch 1, 30 s audio (2580 frames), min-p 0.1, + DAC 44 kHz decode",
                   "frames": T, "batch_per_gpu": B, "l2": "inputs exceed L2 (3.2 GB of weights streamed per decode step)",
                   "options": args.opt, "dac_impl": "tcgen05-fp16" if args.dac_impl == 0 else "simt-fp32"},
        "decode_frames_per_s": B * world / (step_ms_dev * 1e-3), "us_per_decode_step": step_ms_dev * 1e3,
        "realtime_factor": value,
        "roofline": {"bound": "hbm", "kernel": kernel_name,
                     "achieved": step_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": step_gbs / hbm_peak, "frac_of_8TBs_nominal": step_gbs / 8000.0,
                     "peak_kind": peak_kind, "traffic": ncu_traffic_per_step(args), "bytes_per_launch": step_bytes,
                     "dominant_kernel": {"name": "k_gemv_mma<2,SWIGLU> (ln2+fc1+swiglu of the per-op path)", "achieved": fc1_gbs, "frac": fc1_gbs / hbm_peak,
                                         "bytes_per_launch": fc1_bytes, "timed": "isolated, back to back over the 26 layers"}},
        "e2e": {"value": e2e_value, "unit": "audio-s/s", "h2d_bytes_per_step": prefix_host.numel() * 2,
                "d2h_bytes_per_step": wav_host.numel() * 4},
        "gpu_launches": int(launches), "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        r = cpu_reference_sample()
        line["cpu_baseline"] = {
            "value": r["audio_sec_per_sec"], "unit": "audio-s/s", "cores": r["cores"], "kind": "port",
            "sample": f"full-size model B=1+CFG on the host: {r['decode_steps_timed']} decode steps (median {r['ms_per_decode_step']:.1f} ms) + "
                      f"DAC decode of {r['dac_frames_timed']} frames ({r['dac_frames_per_s']:.1f} frames/s); per-frame rates extrapolated to 2580 frames",
            "decode_frames_per_s": r["frames_per_s_decode"]}
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    world, rank, local = dist_setup(args)
    try:
        if args.impl == "reference":
            run_reference(args, world, rank)
        else:
            run_b200(args, world, rank, local)
    finally:
        if world > 1:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.destroy_process_group()


if __name__ == "__main__":
    main()
