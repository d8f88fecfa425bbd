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
WORKLOAD = "Zonos-v0.1-transformer bf16 batch 1, 30 s audio (2580 frames), min-p 0.1, + DAC 44 kHz decode"


def config_dict(args, dac_impl="tcgen05-fp16"):
    """`config` of the JSON line: key-identical for the B200 arm and the reference arm."""
    return {"workload": WORKLOAD, "frames": args.frames, "batch_per_gpu": args.batch,
            "l2": "inputs exceed L2 (3.2 GB of weights streamed per decode step)", "options": list(args.opt), "dac_impl": dac_impl}
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
    ap.add_argument("--no-extras", action="store_true", help="skip the other BASELINE configs / GPU baseline / phase table after the headline")
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
    `ncu --set full` capture (profiles/r02_traffic.json); None when the run is not the configuration that was captured."""
    if args.batch != 1 or "mega=0" in args.opt:
        return None
    try:
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
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
    package that cannot travel to the GPU box), all host threads.  One "step" here is a BOUNDED SAMPLE of the utterance
    (prefill + 6 decode steps + a 22-frame DAC decode): `ms_per_step` is the time such a sample really took; `value` is the
    audio-s/s that the sampled per-frame rates extrapolate to for the whole 2580-frame utterance (`extrapolated`)."""
    if rank != 0:
        return
    vals, walls = [], []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        r = cpu_reference_sample(n_decode_steps=6, dac_frames=22)
        if i >= args.warmup:
            vals.append(r)
            walls.append(time.perf_counter() - t0)
    v = sum(x["audio_sec_per_sec"] for x in vals) / len(vals)
    ms = sum(x["ms_per_decode_step"] for x in vals) / len(vals)
    sample = ("full-size model B=1+CFG: weights + prefill(25 tok) + 6 decode steps + DAC decode of 22 frames per step; "
              "per-frame rates extrapolated to 2580 frames")
    line = {
        "impl": "reference", "metric": "audio_sec_per_sec", "value": v, "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sum(walls) / len(walls) * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "extrapolated": True,
        "ms_per_utterance_extrapolated": (args.frames / FRAME_RATE) / v * 1e3,
        "config": config_dict(args, dac_impl="oracle-fp32"),
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

    from zonos_vibes_b200.replicas import max_over_ranks
    ms_dev = max_over_ranks(ms_dev, dev)
    ms_e2e = max_over_ranks(ms_e2e, dev)
    # ---- the other BASELINE configs, after (outside) the headline's timed region; every rank takes part in the sharded ones
    extra = {}
    if not args.no_extras and args.batch == 1 and not args.opt:
        extra["b64"] = extra_b64(model, ctx, world, rank, dev, args.frames, per_gpu=False, total=64)
        if world > 1:
            extra["b64_per_gpu"] = extra_b64(model, ctx, world, rank, dev, args.frames, per_gpu=True, total=64 * world)
        if rank == 0:
            extra["dac_b32x2580"] = extra_dac(model, dev)
            extra["phases_k_decode_mega"] = phase_table(model, ctx, dev, args)
        extra["hybrid_b16"] = extra_hybrid(dev, world, rank)
        if rank == 0 and world == 1:
            try:
                extra["gpu_baseline"] = gpu_baseline(dev)
            except Exception as e:  # noqa: BLE001  (a baseline leg must not take the bench line down)
                extra["gpu_baseline"] = {"unavailable": repr(e)[:200]}
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
        kernel_name = f"k_decode_mega<4,false,true> (persistent cooperative kernel, 64 decode steps per launch; bytes and time are per step: {body})"
    line = {
        "metric": "audio_sec_per_sec", "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": config_dict(args, "tcgen05-fp16" if args.dac_impl == 0 else "simt-fp32"),
        "decode_frames_per_s": B * world / (step_ms_dev * 1e-3), "us_per_decode_step": step_ms_dev * 1e3,
        "realtime_factor": value,
        "roofline": {"bound": "hbm", "kernel": kernel_name,
                     "achieved": step_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": step_gbs / hbm_peak, "frac_of_8TBs_nominal": step_gbs / 8000.0,
                     "peak_kind": peak_kind, "traffic": ncu_traffic_per_step(args), "bytes_per_launch": step_bytes},
        "e2e": {"value": e2e_value, "unit": "audio-s/s", "h2d_bytes_per_step": prefix_host.numel() * 2,
                "d2h_bytes_per_step": wav_host.numel() * 4},
        "gpu_launches": int(launches), "clocks": clocks,
    }
    if extra:
        line["configs"] = {k: v for k, v in extra.items() if k != "gpu_baseline" and v is not None}
        if "gpu_baseline" in extra:
            line["gpu_baseline"] = extra["gpu_baseline"]
            if "us_per_decode_step" in extra["gpu_baseline"]:
                gb = line["gpu_baseline"]
                gb["speedup_of_this_repo_vs_eager"] = gb["us_per_decode_step"] / (step_ms_dev * 1e3)
                if gb.get("us_per_decode_step_cuda_graph_replay"):
                    gb["speedup_of_this_repo_vs_graph_replay"] = gb["us_per_decode_step_cuda_graph_replay"] / (step_ms_dev * 1e3)
    if world == 1 and not args.no_cpu_baseline:
        r = cpu_reference_sample()
        line["cpu_baseline"] = {
            "value": r["audio_sec_per_sec"], "unit": "audio-s/s", "cores": r["cores"], "kind": "port",
            "sample": f"full-size model B=1+CFG on the host: {r['decode_steps_timed']} decode steps (median {r['ms_per_decode_step']:.1f} ms) + "
                      f"DAC decode of {r['dac_frames_timed']} frames ({r['dac_frames_per_s']:.1f} frames/s); per-frame rates extrapolated to 2580 frames",
            "decode_frames_per_s": r["frames_per_s_decode"]}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------
# the other BASELINE configs + same-box GPU baseline + per-phase table (all timed AFTER the headline's timed region)
# ----------------------------------------------------------------------------------------------------------------
def peaks_full():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d["bf16_tflops"]), float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


def extra_b64(model, ctx, world, rank, dev, T, per_gpu, total):
    """BASELINE config 3: `total` utterances of T frames sharded over the ranks (replicas.generate_sharded: batched
    tcgen05 decode path per rank + DAC + pinned-host gather).  Device events around the whole call, max over ranks."""
    from zonos_vibes_b200.replicas import generate_sharded, max_over_ranks
    g = torch.Generator(device="cpu").manual_seed(7)
    prefixes = [torch.randn(2, LC, FULL["d_model"], generator=g).bfloat16() for _ in range(total)]
    prefixes = [p.to(dev) for p in prefixes]
    st = {}
    barrier(world, dev)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = generate_sharded(model, prefixes, T, max_batch=64, seed=5, stats=st)
    e1.record()
    torch.cuda.synchronize()
    barrier(world, dev)
    ms = max_over_ranks(e0.elapsed_time(e1), dev)
    step_ms = max_over_ranks(st["decode_step_ms"] or 0.0, dev)
    if rank != 0:
        return None
    frames = sum(int(w.shape[-1]) // 512 for w in out)
    b_rank = (total + world - 1) // world
    n_dec = T + 8
    mean_ctx = LC + 1 + n_dec / 2.0
    step_bytes = ctx.lib.zb_weight_bytes(ctx.handle) + 106496 * b_rank * (mean_ctx + 1) + 36864 * b_rank
    hbm, _, _, kind = peaks_full()
    gbs = step_bytes / (step_ms * 1e-3) / 1e9
    return {"utterances_total": total, "utterances_per_gpu": b_rank, "mode": "64 per GPU (weak)" if per_gpu else "64 total (strong)",
            "frames": T, "ms": ms, "audio_sec_per_sec": frames / FRAME_RATE / (ms * 1e-3), "us_per_decode_step": step_ms * 1e3,
            "decode_frames_per_s": total / (step_ms * 1e-3),
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm, "peak_kind": kind,
                         "bytes_per_launch": step_bytes, "kernel": "decode step of the slowest rank, batched tcgen05 path (k_bt_gemm + finish kernels + k_attn_decode_tma)"}}


def extra_dac(model, dev, B=32, T=2580, runs=3):
    """BASELINE config 4: DAC decoder alone, 32 x 2580 frames -> waveform."""
    codes = torch.randint(0, 1024, (B, 9, T), generator=torch.Generator().manual_seed(2)).to(dev)
    model.autoencoder.decode(codes)
    torch.cuda.synchronize()
    ms = []
    for _ in range(runs):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        model.autoencoder.decode(codes)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    m = sum(ms) / len(ms)
    tf = 1.608302592e9 * B * T / (m * 1e-3) / 1e12
    _, burst, sus, kind = peaks_full()
    return {"batch": B, "frames": T, "ms": m, "ms_best": min(ms), "tflops": tf, "frames_per_s": B * T / (m * 1e-3),
            "roofline": {"bound": "tensor", "achieved": tf, "peak": sus, "unit": "TFLOP/s", "frac": tf / sus, "frac_of_burst_peak": tf / burst,
                         "peak_kind": kind + " bf16_tflops_sustained (a 0.2 s job under the power cap)", "flops_per_launch": 1.608302592e9 * B * T,
                         "kernel": "k_dac_conv_tc / k_dac_conv_tc2<8> (tcgen05 implicit-GEMM convs, 30 layers)"}}


def extra_hybrid(dev, world, rank, frames=96):
    """BASELINE config 5: synthetic Zonos-v0.1-hybrid (Mamba2 + attention), 16 utterances over the ranks."""
    import zonos_vibes_b200 as z
    from zonos_vibes_b200.replicas import max_over_ranks
    b = max(1, 16 // world)
    model = z.Zonos.from_state_dict(z.ZonosConfig.from_dict(hybrid_config_dict()), random_hybrid_state_dict(dev), dev)
    ctx = model._native()
    prefix = torch.randn(2 * b, LC, FULL["d_model"], device=dev).bfloat16()
    model.generate(prefix, max_new_tokens=8, batch_size=b, progress_bar=False, seed=1)
    model.generate(prefix, max_new_tokens=frames, batch_size=b, progress_bar=False, seed=2)
    step_ms = max_over_ranks(model.last_step_ms, dev)
    wb = int(ctx.lib.zb_weight_bytes(ctx.handle))
    n_attn, n_m = len(HYBRID["attn_layer_idx"]), HYBRID["n_layer"] - len(HYBRID["attn_layer_idx"])
    di = HYBRID["expand"] * FULL["d_model"]
    state_bytes = n_m * 2 * b * (2 * (di // HYBRID["headdim"]) * HYBRID["headdim"] * HYBRID["d_state"] * 4)      # ssm state read + write, 2 CFG rows
    kv_bytes = 2 * b * n_attn * 2 * FULL["num_heads_kv"] * 128 * 2 * (LC + frames / 2)
    step_bytes = wb + state_bytes + kv_bytes                        # algorithmic: every weight byte once per step (batched tcgen05 path, >= 9 rows)
    hbm, _, _, kind = peaks_full()
    gbs = step_bytes / (step_ms * 1e-3) / 1e9
    del model
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    return {"utterances_total": b * world, "utterances_per_gpu": b, "frames": frames, "us_per_decode_step": step_ms * 1e3,
            "decode_frames_per_s": b * world / (step_ms * 1e-3), "audio_sec_per_sec": b * world / (step_ms * 1e-3) / FRAME_RATE,
            "architecture": "synthetic: 46 layers, 5 attention + GatedMLP(8192), 41 Mamba2 (d_state 128, headdim 64, expand 2); dims unverified (SURVEY App. C)",
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm, "peak_kind": kind, "bytes_per_launch": step_bytes,
                         "kernel": "batched tcgen05 path (k_bt_gemm + finish kernels for the Mamba2 / attention projections, k_m2_conv / k_m2_ssm / k_m2_gated_norm, k_attn_decode_tma), one CUDA graph per step; bytes = weights once + SSM state read + write (fp32) + KV"}}


def gpu_baseline(dev, steps=48, prefix_frames=1290):
    """Secondary baseline: the reference ALGORITHM (oracle restatement of zonos/model.py + backbone/_torch.py,
    plain torch ops) on this same B200 with stock torch kernels (cuBLAS GEMVs + SDPA, eager - the torch backbone of the
    reference does not use CUDA graphs: model.py `can_use_cudagraphs` is true for the mamba-ssm backbone only), full-size
    model, decode steps at ctx ~ 1300 incl. sampling.  Not the product path: oracle/ is imported here as a baseline only."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import zonos_oracle as zo
    cfg = zo.OracleConfig(eos_token_id=1025)
    sd = {k: v.to(dev) for k, v in zo.synth_state_dict(cfg, 0).items()}
    prefix = zo.synth_prefix_conditioning(cfg, 1, LC, 1).to(dev)
    codes = torch.randint(0, 1024, (1, 9, prefix_frames + steps + 16), generator=torch.Generator().manual_seed(3))
    delayed = zo.apply_delay_pattern(codes, 1025).to(dev)
    with torch.inference_mode(), torch.device(dev):         # factory functions inside the oracle default to the GPU
        o = zo.OracleZonos(cfg, sd)
        params = o.setup_cache(2, LC + delayed.shape[2] + 9)
        o.freqs_cis = o.freqs_cis.to(dev)
        for k in params.key_value_memory_dict:
            params.key_value_memory_dict[k] = params.key_value_memory_dict[k].to(dev)
        params.lengths_per_sample = params.lengths_per_sample.to(dev)
        o.prefill(prefix, delayed[..., : prefix_frames + 1], params, 2.0)
        n0 = LC + prefix_frames + 1
        params.seqlen_offset += n0
        params.lengths_per_sample[:] += n0

        def step(s):
            col = prefix_frames + 1 + s
            logits = o.decode_one_token(delayed[..., col: col + 1], params, 2.0)
            zo.sample_from_logits(logits, min_p=0.1, generated_tokens=delayed[..., : col + 1])
            params.seqlen_offset += 1
            params.lengths_per_sample[:] += 1

        for s in range(8):
            step(s)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in range(8, 8 + steps):
            step(s)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / steps
        # the same step (fixed context) captured in a CUDA graph and replayed: the GPU time of the stock kernels without the
        # eager host overhead (an upper bound on what graph capture could give the reference; the reference itself runs
        # this backbone eagerly)
        us_graph = None
        try:
            col = prefix_frames + 1 + 8 + steps
            stream = torch.cuda.Stream()
            stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(stream):
                for _ in range(2):
                    zo.sample_from_logits(o.decode_one_token(delayed[..., col: col + 1], params, 2.0), min_p=0.1, generated_tokens=delayed[..., : col + 1])
            torch.cuda.current_stream().wait_stream(stream)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                zo.sample_from_logits(o.decode_one_token(delayed[..., col: col + 1], params, 2.0), min_p=0.1, generated_tokens=delayed[..., : col + 1])
            for _ in range(3):
                graph.replay()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(steps):
                graph.replay()
            e1.record()
            torch.cuda.synchronize()
            us_graph = e0.elapsed_time(e1) * 1e3 / steps
        except Exception as e:  # noqa: BLE001
            us_graph = None
            graph_err = repr(e)[:160]
    del o, sd, params
    torch.cuda.empty_cache()
    out = {"us_per_decode_step": us, "decode_frames_per_s": 1e6 / us, "steps_timed": steps, "ctx": [n0 + 8, n0 + 8 + steps],
           "what": "reference algorithm (oracle port) with stock torch " + torch.__version__ + " CUDA kernels (cuBLAS + SDPA), eager like the "
                   "reference's torch backbone, same B200, batch 1 + CFG, min-p sampling",
           "us_per_decode_step_cuda_graph_replay": us_graph}
    if us_graph is None:
        out["cuda_graph_error"] = graph_err
    return out


def phase_table(model, ctx, dev, args, T=128, ctx_len=1290):
    """Per-phase breakdown of one decode step of k_decode_mega from one TRACED launch (slower instantiation with
    %globaltimer stamps on CTA 0), outside every timed region, at the mean context of the headline workload (a conditioning
    prefix of `ctx_len` tokens): us and algorithmic bytes per phase, layers 1-2 averaged."""
    import ctypes as C
    ctx.set_option("trace", 1)
    ctx.set_option("trace_quiet", 1)
    try:
        prefix = torch.randn(2, ctx_len, FULL["d_model"], device=dev).bfloat16()
        model.generate(prefix, max_new_tokens=T, batch_size=1, progress_bar=False, seed=3)
        buf = (C.c_uint64 * 64)()
        n = ctx.lib.zb_last_trace(ctx.handle, buf, 64)
    finally:
        ctx.set_option("trace", 0)
        ctx.set_option("trace_quiet", 0)
    # stamp order per layer (mega_kernel.cuh, MG_STAMP): dependencies that are grid barriers in the ll < 2 / ll < 3 modes are
    # polls of epoch-tagged words inside the following staging in the default mode (ll = 3), where their slots are ~0
    ll = 3
    for kv in args.opt:
        if kv.startswith("ll="):
            ll = int(kv[3:])
    names = ["ln1_stage", "in_proj", "dep_in_proj_to_attention", "attention", "dep_attention_to_out_proj", "y_stage", "out_proj", "ln2_stage", "fc1"]
    names += (["dep_fc1_to_fc2"] if ll < 3 else []) + ["h_stage", "fc2"]
    per = len(names)
    if n < 2 + 3 * per:
        return None
    t = [int(buf[i]) for i in range(n)]
    d, f = FULL["d_model"], FULL["ffn"]
    qkv = (FULL["num_heads"] + 2 * FULL["num_heads_kv"]) * 128
    by = {"in_proj": qkv * d * 2, "out_proj": d * d * 2, "fc1": 2 * f * d * 2, "fc2": d * f * 2,
          "attention": 2 * FULL["num_heads_kv"] * 2 * 128 * 2 * (ctx_len + T // 2)}
    acc = {}
    for layer in (1, 2):
        base = 2 + per * layer
        for j, nm in enumerate(names):
            acc[nm] = acc.get(nm, 0.0) + (t[base + j] - t[base + j - 1]) / 2e3
    rows = [{"phase": nm, "us": round(us, 2), "bytes": by.get(nm, 0), "gbs": round(by[nm] / us / 1e3, 1) if nm in by and us > 0 else None} for nm, us in acc.items()]
    return {"per_layer_us": round(sum(acc.values()), 2), "ctx": ctx_len + T // 2, "dependency_mode_ll": ll, "phases": rows,
            "note": "traced instantiation (a few % slower than the production kernel); stamps of CTA 0; *_stage = polling the previous phase's "
                    "epoch-tagged output + LayerNorm where there is one; dep_* = grid barriers (zero when the data carries the dependency)"}


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
