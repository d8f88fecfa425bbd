"""CPU: the oracle (oracle/*.py) against the golden vectors frozen from the real reference (oracle/make_golden.py),
and live against /root/reference + the installed transformers DacModel when they are present."""
import pytest
import torch

import dac_oracle as do
import zonos_oracle as zo
from ref_loader import reference_available

CFGS = {"tiny": zo.TINY_CONFIG, "small": zo.SMALL_CONFIG}


@pytest.mark.parametrize("name", ["small", "one", "empty", "big"])
def test_delay_pattern_golden(golden, name):
    codes = golden[f"delay.{name}.codes"]
    d = zo.apply_delay_pattern(codes, 1025)
    assert torch.equal(d, golden[f"delay.{name}.delayed"])
    assert torch.equal(zo.revert_delay_pattern(d), golden[f"delay.{name}.reverted"])
    assert torch.equal(zo.revert_delay_pattern(d), codes)          # round trip


def test_sampling_filters_golden(golden):
    logits, hist = golden["sample.logits"], golden["sample.history"]
    for name, kw in golden["sample.cases"].items():
        for h in (0, 1):
            _, pr = zo.filtered_probs(logits.clone(), generated_tokens=hist if h else None, **kw)
            assert torch.equal(pr, golden[f"sample.{name}.h{h}.probs"]), (name, h)
    assert torch.equal(zo.sample_from_logits(logits.clone(), temperature=0, generated_tokens=hist), golden["sample.greedy.h1.tokens"])
    assert torch.equal(zo.sample_from_logits(logits.clone(), temperature=0), golden["sample.greedy.h0.tokens"])


@pytest.mark.parametrize("cname", ["tiny", "small"])
def test_generate_streams_golden(golden, cname):
    cfg = CFGS[cname]
    sd = zo.synth_state_dict(cfg, 0)
    o = zo.OracleZonos(cfg, sd)
    prefix = zo.synth_prefix_conditioning(cfg, 1, 6, 1)
    for sname, sp in {"greedy": dict(temperature=0), "minp": dict(min_p=0.1)}.items():
        torch.manual_seed(5)
        out = o.generate(prefix, max_new_tokens=48, sampling_params=sp)
        assert torch.equal(out, golden[f"gen.{cname}.{sname}.codes"]), sname
    tf = o.teacher_forced_logits(prefix, golden[f"gen.{cname}.greedy.delayed"], 40)
    assert torch.equal(tf, golden[f"gen.{cname}.greedy.logits"])


@pytest.mark.parametrize("cname", ["tiny", "small"])
def test_generate_eos_winddown_golden(golden, cname):
    from make_golden import eos_biased
    cfg = CFGS[cname]
    o = zo.OracleZonos(cfg, eos_biased(zo.synth_state_dict(cfg, 0), cfg))
    prefix = zo.synth_prefix_conditioning(cfg, 1, 6, 1)
    for seed in (1, 2, 3):
        torch.manual_seed(seed)
        out = o.generate(prefix, max_new_tokens=64, sampling_params=dict(min_p=0.1))
        assert torch.equal(out, golden[f"gen.{cname}.eos.seed{seed}.codes"]), seed
    torch.manual_seed(0)
    out = o.generate(prefix, max_new_tokens=64, sampling_params=dict(temperature=0))
    assert torch.equal(out, golden[f"gen.{cname}.eos.greedy.codes"])


@pytest.mark.parametrize("cname", ["tiny", "small"])
def test_backbone_boundary_golden(golden, cname):
    cfg = CFGS[cname]
    o = zo.OracleZonos(cfg, zo.synth_state_dict(cfg, 0))
    x = golden[f"backbone.{cname}.x"]
    params = o.setup_cache(2, 32)
    y0 = o.backbone(x[:, :8], params)
    params.seqlen_offset += 8
    params.lengths_per_sample[:] += 8
    y1 = o.backbone(x[:, 8:9], params)
    assert torch.equal(y0, golden[f"backbone.{cname}.y_prefill"])
    assert torch.equal(y1, golden[f"backbone.{cname}.y_decode"])


def test_full_size_logits_golden(golden):
    cfg = zo.FULL_CONFIG
    o = zo.OracleZonos(cfg, zo.synth_state_dict(cfg, 0))
    prefix = zo.synth_prefix_conditioning(cfg, 1, 24, 1)
    tf = o.teacher_forced_logits(prefix, golden["full.delayed"], 2)
    assert torch.equal(tf, golden["full.logits"][:3])


@pytest.mark.parametrize("name", ["b1", "b2"])
def test_dac_decode_golden(golden, name):
    wav = do.dac_decode(do.synth_dac_state_dict(2), golden[f"dac.{name}.codes"])
    assert wav.shape == golden[f"dac.{name}.wav"].shape
    assert torch.equal(wav, golden[f"dac.{name}.wav"])
    assert wav.abs().max() > 0.2          # the fixture is not a trivial near-zero signal


def test_dac_against_installed_transformers():
    sd = do.synth_dac_state_dict(4)
    codes = torch.randint(0, 1024, (1, 9, 12), generator=torch.Generator().manual_seed(9))
    hf = do.load_into_hf(sd)
    with torch.inference_mode():
        want = hf.decode(audio_codes=codes).audio_values.unsqueeze(1).float()
    assert torch.equal(do.dac_decode(sd, codes), want)


@pytest.mark.skipif(not reference_available(), reason="/root/reference not present (GPU box)")
def test_oracle_against_live_reference():
    from ref_loader import build_reference_model
    cfg = zo.TINY_CONFIG
    sd = zo.synth_state_dict(cfg, 3)
    ref, model = build_reference_model(cfg, sd)
    o = zo.OracleZonos(cfg, sd)
    prefix = zo.synth_prefix_conditioning(cfg, 1, 5, 2)
    for sp in (dict(temperature=0), dict(top_k=30, top_p=0.9), dict(linear=0.5, conf=0.4, quad=0.0)):
        torch.manual_seed(11)
        a = model.generate(prefix, max_new_tokens=24, sampling_params=sp, progress_bar=False, disable_torch_compile=True)
        torch.manual_seed(11)
        b = o.generate(prefix, max_new_tokens=24, sampling_params=sp)
        assert torch.equal(a, b), sp
    codes = torch.randint(0, 1024, (2, 9, 11))
    assert torch.equal(ref.pattern.apply_delay_pattern(codes, 1025), zo.apply_delay_pattern(codes, 1025))


def test_mamba2_mixer_matches_hf_port():
    """Row a8: the Mamba2 mixer arithmetic of oracle/hybrid_oracle.py (stepped recurrence) against the pure-torch port of
    mamba-ssm's Mamba2 that ships with transformers (`Mamba2Mixer.torch_forward`, chunked SSD over the whole sequence;
    same parameter layout as mamba_ssm.modules.mamba2.Mamba2).  mamba-ssm itself is not installable here."""
    import hybrid_oracle as ho
    mm = pytest.importorskip("transformers.models.mamba2.modeling_mamba2")
    from transformers.models.mamba2.configuration_mamba2 import Mamba2Config

    cfg = ho.HybridConfig(d_model=256, headdim=64, d_state=128)
    hc = Mamba2Config(hidden_size=cfg.d_model, state_size=cfg.d_state, conv_kernel=cfg.d_conv, expand=cfg.expand, head_dim=cfg.headdim,
                      n_groups=cfg.ngroups, num_heads=cfg.nheads, use_conv_bias=True, use_bias=False, rms_norm=True, chunk_size=16,
                      layer_norm_epsilon=1e-5, num_hidden_layers=1)
    mix = mm.Mamba2Mixer(hc, layer_idx=0).float().eval()
    p = ho.synth_mamba2_params(cfg, 3, dtype=torch.float32)
    with torch.no_grad():
        for name, dst in (("in_proj.weight", mix.in_proj.weight), ("conv1d.weight", mix.conv1d.weight), ("conv1d.bias", mix.conv1d.bias),
                          ("dt_bias", mix.dt_bias), ("A_log", mix.A_log), ("D", mix.D), ("norm.weight", mix.norm.weight),
                          ("out_proj.weight", mix.out_proj.weight)):
            dst.copy_(p[name])
        u = torch.randn(2, 37, cfg.d_model, generator=torch.Generator().manual_seed(5))
        want = mix.torch_forward(u)
        got = ho.mamba2_mixer_sequence(cfg, p, ho.mamba2_alloc_state(cfg, 2), u)
    assert got.shape == want.shape
    assert (got - want).abs().max().item() < 1e-4 * want.abs().max().item()


# ---- round 2: long-context / audio-prefix / long-DAC fixtures (oracle/make_golden_long.py, frozen from the imported reference)
@pytest.fixture(scope="module")
def golden_long():
    import os
    return torch.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_long.pt"))


@pytest.mark.parametrize("p", [5, 300])
def test_generate_with_audio_prefix_golden(golden_long, p):
    """model.py:246-264: the oracle's audio-prefix path against the reference's generate(audio_prefix_codes=...)."""
    cfg = zo.SMALL_CONFIG
    o = zo.OracleZonos(cfg, zo.synth_state_dict(cfg, 0))
    prefix = zo.synth_prefix_conditioning(cfg, 1, 6, 1)
    tr = {}
    out = o.generate(prefix, audio_prefix_codes=golden_long[f"prefix.p{p}.audio_prefix_codes"], max_new_tokens=24,
                     sampling_params=dict(temperature=0), trace=tr)
    assert torch.equal(out, golden_long[f"prefix.p{p}.codes"])
    assert torch.equal(tr["delayed"], golden_long[f"prefix.p{p}.delayed"])
    assert torch.equal(torch.cat(tr["logits"], 0), golden_long[f"prefix.p{p}.logits"])


def test_dac_long_windows_golden(golden_long):
    """T = 2580 fixture: the oracle decodes each 32-frame window with a 16-frame halo (window decode equals the
    full decode in the interior) and must reproduce the reference waveform there."""
    dsd = do.synth_dac_state_dict(2)
    codes = golden_long["dac.long.codes"]
    T = codes.shape[-1]
    for f, want in list(zip(golden_long["dac.long.window_frames"].tolist(), golden_long["dac.long.windows"]))[:3]:
        lo, hi = max(0, f - 16), min(T, f + 32 + 16)
        wav = do.dac_decode(dsd, codes[..., lo:hi])
        got = wav[0, 0, 512 * (f - lo): 512 * (f - lo + 32)]
        assert (got - want).abs().max().item() <= 1e-4, f


def test_long_fixture_shapes(golden_long):
    for p in (1290, 2560):
        assert golden_long[f"long.p{p}.logits"].shape == (9, 9, 1026)
        assert golden_long[f"long.p{p}.delayed"].shape == (1, 9, p + 12 + 9)
        assert bool(torch.isfinite(golden_long[f"long.p{p}.logits"][..., :1024]).all())
    assert golden_long["full.truth_logits"].shape == (7, 9, 1026)


def test_dac_encode_oracle_matches_hf_model():
    """Row f3 pin: the encoder + RVQ restatement (oracle/dac_oracle.py dac_encode) against the installed transformers
    DacModel.encode (what zonos/autoencoder.py:22-23 calls) on the same synthetic weights and waveform: codes identical."""
    import dac_oracle as do
    sd = dict(do.synth_dac_state_dict(2))
    sd.update(do.synth_dac_encoder_state_dict(3))
    wav = do.synth_waveform(2, 20, 7)
    codes, gaps = do.dac_encode(sd, wav, with_gaps=True)
    assert codes.shape == (2, 9, 20) and codes.dtype == torch.int64 and int(codes.min()) >= 0 and int(codes.max()) < 1024
    m = do.load_into_hf(sd)
    with torch.inference_mode():
        ref = m.encode(wav).audio_codes
    assert torch.equal(ref, codes)
    assert bool((gaps >= 0).all())


@pytest.mark.parametrize("name", ["shallow", "full"])
def test_speaker_embedding_oracle_golden(name):
    """Row f4 pin: oracle/speaker_oracle.py against vectors frozen from the reference's own ResNet293_based + LDA classes
    (oracle/make_golden_spk.py, full depth 10 + 20 + 64 + 3 SimAM blocks and a shallow stack)."""
    import os
    import speaker_oracle as so
    g = torch.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_spk.pt"))
    nb = g[f"{name}.num_blocks"]
    sd = so.synth_speaker_state_dict(5, nb)
    emb, lda = so.embedding_from_features(sd, so.synth_features(2, g[f"{name}.frames"], 8), nb, so.synth_lda_state_dict(6))
    for got, want in ((emb, g[f"{name}.embedding"]), (lda, g[f"{name}.lda"])):
        assert got.shape == want.shape
        assert (got - want).abs().max().item() <= 1e-4 * want.abs().max().item()
