"""Freeze golden vectors from the REAL reference (imported from /root/reference) into tests/golden/.

Run in the build container only (`python oracle/make_golden.py`); the GPU box has no /root/reference.  Every
fixture is produced by the reference's own code path (zonos/model.py, sampling.py, codebook_pattern.py, and the
installed transformers DacModel behind zonos/autoencoder.py) on deterministic synthetic weights that
`oracle/zonos_oracle.synth_state_dict` / `oracle/dac_oracle.synth_dac_state_dict` regenerate bit-identically anywhere.
"""
from __future__ import annotations

import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import dac_oracle as do          # noqa: E402
import zonos_oracle as zo        # noqa: E402
from ref_loader import build_reference_model, load_reference   # noqa: E402

OUT = os.path.join(HERE, "..", "tests", "golden")


def eos_biased(sd: dict, cfg, alpha: float = 0.05) -> dict:
    """Make EOS likely for codebook 0 (head row 1024 aligned with norm_f.bias) so the wind-down path is exercised."""
    sd = dict(sd)
    w = sd["heads.0.weight"].clone()
    w[1024] = (alpha * torch.sign(sd["backbone.norm_f.bias"].float())).to(w.dtype)
    sd["heads.0.weight"] = w
    return sd


def ref_teacher_forced(ref, model, prefix, delayed, n_steps, cfg_scale=2.0):
    """Per-step logits from the reference's own _prefill/_decode_one_token (+ logit_bias, model.py:266-267,280)."""
    params = model.setup_cache(2, prefix.shape[1] + delayed.shape[2] + 9)
    out = [model._prefill(prefix, delayed[..., :1], params, cfg_scale)]
    plen = prefix.shape[1] + 1
    params.seqlen_offset += plen
    params.lengths_per_sample[:] += plen
    bias = torch.zeros_like(out[0])
    bias[:, 1:, model.eos_token_id] = -torch.inf
    for s in range(n_steps):
        out.append(model._decode_one_token(delayed[..., s + 1: s + 2], params, torch.tensor(cfg_scale), allow_cudagraphs=False) + bias)
        params.seqlen_offset += 1
        params.lengths_per_sample[:] += 1
    return torch.cat(out, 0)


@torch.inference_mode()
def main():
    os.makedirs(OUT, exist_ok=True)
    ref = load_reference()
    g = {}

    # ---- delay pattern (codebook_pattern.py) incl. ragged/empty cases
    for name, (b, t) in {"small": (2, 17), "one": (1, 1), "empty": (1, 0), "big": (3, 300)}.items():
        codes = torch.randint(0, 1024, (b, 9, t), generator=torch.Generator().manual_seed(10 + t))
        d = ref.pattern.apply_delay_pattern(codes, 1025)
        g[f"delay.{name}.codes"] = codes
        g[f"delay.{name}.delayed"] = d
        g[f"delay.{name}.reverted"] = ref.pattern.revert_delay_pattern(d)

    # ---- sampling: filtered distributions for fixed logits (masks must be bit-exact)
    gen = torch.Generator().manual_seed(7)
    logits = torch.randn(2, 9, 1026, generator=gen) * 3.0
    logits[..., 1025:] = -torch.inf
    logits[:, 1:, 1024] = -torch.inf
    hist = torch.randint(0, 1026, (2, 9, 5), generator=gen)
    hist[0, 0, -1] = hist[0, 0, -2]          # a token twice in the window -> penalty squared
    g["sample.logits"], g["sample.history"] = logits, hist
    cases = {
        "minp": dict(min_p=0.1), "topk": dict(top_k=50), "topp": dict(top_p=0.9), "topk_topp": dict(top_k=40, top_p=0.8),
        "unified": dict(linear=0.5, conf=0.4, quad=0.0), "unified_q": dict(linear=0.3, conf=-0.1, quad=0.25, min_p=0.05),
        "temp": dict(temperature=0.7, min_p=0.02), "all": dict(temperature=1.3, top_p=0.95, top_k=200, min_p=0.01),
    }
    sys.path.insert(0, HERE)
    for name, kw in cases.items():
        for use_hist in (0, 1):
            lg, pr = zo.filtered_probs(logits.clone(), generated_tokens=hist if use_hist else None, **kw)
            # cross-check the restated chain against the reference's own functions
            l2 = logits.clone()
            if use_hist:
                l2 = ref.sampling.modify_logit_for_repetition_penalty(l2, hist, 3.0, 2)
            p2 = torch.softmax(l2 / kw.get("temperature", 1.0), dim=-1)
            if kw.get("linear", 0) > 0:
                p2 = ref.sampling.apply_unified(p2, kw["linear"], kw.get("conf", 0.0), kw.get("quad", 0.0))
            if kw.get("top_p", 0) > 0:
                p2 = ref.sampling.apply_top_p(p2, kw["top_p"])
            if kw.get("top_k", 0) > 0:
                p2 = ref.sampling.apply_top_k(p2, kw["top_k"])
            if kw.get("min_p", 0) > 0:
                p2 = ref.sampling.apply_min_p(p2, kw["min_p"])
            assert torch.equal(pr, p2), name
            g[f"sample.{name}.h{use_hist}.probs"] = pr
    g["sample.greedy.h1.tokens"] = ref.sampling.sample_from_logits(logits.clone(), temperature=0, generated_tokens=hist)
    g["sample.greedy.h0.tokens"] = ref.sampling.sample_from_logits(logits.clone(), temperature=0)
    g["sample.cases"] = {k: dict(v) for k, v in cases.items()}

    # ---- generation: greedy + sampled streams, per-step logits, EOS wind-down (tiny/small configs)
    for cname, cfg in {"tiny": zo.TINY_CONFIG, "small": zo.SMALL_CONFIG}.items():
        sd = zo.synth_state_dict(cfg, 0)
        _, model = build_reference_model(cfg, sd)
        prefix = zo.synth_prefix_conditioning(cfg, 1, 6, 1)
        for sname, sp in {"greedy": dict(temperature=0), "minp": dict(min_p=0.1)}.items():
            torch.manual_seed(5)
            out = model.generate(prefix, max_new_tokens=48, sampling_params=sp, progress_bar=False, disable_torch_compile=True)
            g[f"gen.{cname}.{sname}.codes"] = out
        # teacher-forced logits on the greedy stream
        oracle = zo.OracleZonos(cfg, sd)
        trace = {}
        oracle.generate(prefix, max_new_tokens=48, sampling_params=dict(temperature=0), trace=trace)
        delayed = trace["delayed"]
        g[f"gen.{cname}.greedy.delayed"] = delayed
        tf = ref_teacher_forced(ref, model, prefix, delayed, 40)
        assert torch.equal(tf, torch.cat(trace["logits"][:41], 0))
        g[f"gen.{cname}.greedy.logits"] = tf
        # EOS wind-down
        sd_e = eos_biased(sd, cfg)
        _, model_e = build_reference_model(cfg, sd_e)
        for seed in (1, 2, 3):
            torch.manual_seed(seed)
            out = model_e.generate(prefix, max_new_tokens=64, sampling_params=dict(min_p=0.1), progress_bar=False,
                                   disable_torch_compile=True)
            g[f"gen.{cname}.eos.seed{seed}.codes"] = out
        torch.manual_seed(0)
        g[f"gen.{cname}.eos.greedy.codes"] = model_e.generate(prefix, max_new_tokens=64, sampling_params=dict(temperature=0),
                                                              progress_bar=False, disable_torch_compile=True)
        tr = {}
        out_e = zo.OracleZonos(cfg, sd_e).generate(prefix, max_new_tokens=64, sampling_params=dict(temperature=0), trace=tr)
        assert torch.equal(out_e, g[f"gen.{cname}.eos.greedy.codes"])
        g[f"gen.{cname}.eos.greedy.delayed"] = tr["delayed"]
        g[f"gen.{cname}.eos.greedy.logits"] = torch.cat(tr["logits"], 0)
        # backbone plugin boundary: forward() on a [2, S, d] block, prefill then one decode token
        params = model.setup_cache(2, 32)
        x = zo.synth_prefix_conditioning(cfg, 1, 9, 3)
        y0 = model.backbone(x[:, :8], params)
        params.seqlen_offset += 8
        params.lengths_per_sample[:] += 8
        y1 = model.backbone(x[:, 8:9], params)
        g[f"backbone.{cname}.x"], g[f"backbone.{cname}.y_prefill"], g[f"backbone.{cname}.y_decode"] = x, y0, y1

    # ---- full-size model (26 layers, 1.6 B params): teacher-forced logits for a few steps
    cfg = zo.FULL_CONFIG
    sd = zo.synth_state_dict(cfg, 0)
    _, model = build_reference_model(cfg, sd)
    prefix = zo.synth_prefix_conditioning(cfg, 1, 24, 1)
    codes = torch.randint(0, 1024, (1, 9, 12), generator=torch.Generator().manual_seed(3))
    delayed = ref.pattern.apply_delay_pattern(codes, 1025)
    g["full.delayed"] = delayed
    g["full.logits"] = ref_teacher_forced(ref, model, prefix, delayed, 6)
    # greedy stream of the first 256 frames from the real reference + the oracle's per-step top-2 (after the repetition
    # penalty) so that GPU tests can tell a genuine mismatch from a bf16 coin flip
    torch.manual_seed(0)
    codes256 = model.generate(prefix, max_new_tokens=256, sampling_params=dict(temperature=0), progress_bar=False,
                              disable_torch_compile=True)
    tr = {}
    o_full = zo.OracleZonos(cfg, sd)
    assert torch.equal(o_full.generate(prefix, max_new_tokens=256, sampling_params=dict(temperature=0), trace=tr), codes256)
    g["full.greedy.codes"] = codes256
    g["full.greedy.delayed"] = tr["delayed"]
    top_v, top_i = [], []
    for s, lg in enumerate(tr["logits"]):
        if s > 0:
            lg = zo.repetition_penalty(lg, tr["delayed"][..., : s + 1], 3.0, 2)
        v, i = lg[0].topk(2, dim=-1)
        top_v.append(v)
        top_i.append(i)
    g["full.greedy.top2_values"] = torch.stack(top_v)          # [steps, 9, 2]
    g["full.greedy.top2_indices"] = torch.stack(top_i)
    g["full.greedy.logit_rms"] = torch.stack([lg[torch.isfinite(lg)].pow(2).mean().sqrt() for lg in tr["logits"]])
    del model, sd, o_full

    # ---- DAC decode through the reference's DACAutoencoder (transformers DacModel, fp32 on CPU)
    dsd = do.synth_dac_state_dict(2)
    auto = ref.autoencoder.DACAutoencoder()
    full = auto.dac.state_dict()
    full.update(dsd)
    auto.dac.load_state_dict(full)
    for name, (b, t) in {"b1": (1, 40), "b2": (2, 23)}.items():
        codes = torch.randint(0, 1024, (b, 9, t), generator=torch.Generator().manual_seed(30 + t))
        g[f"dac.{name}.codes"] = codes
        g[f"dac.{name}.wav"] = auto.decode(codes)

    torch.save(g, os.path.join(OUT, "golden.pt"))
    print("wrote", len(g), "entries,", os.path.getsize(os.path.join(OUT, "golden.pt")) / 1e6, "MB")


if __name__ == "__main__":
    main()
