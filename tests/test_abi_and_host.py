"""CPU: the C-ABI library loads and exports every symbol include/zonos_b200.h declares; host-side logic; the
multi-replica plumbing over gloo with world_size 2.  No compute calls (no GPU here)."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "zonos_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(zb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from zonos_vibes_b200 import _lib
    lib = _lib.load()
    syms = _header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in zonos_b200.h but not exported"
    assert set(_lib.SIGNATURES) == set(syms), set(_lib.SIGNATURES) ^ set(syms)
    assert lib.zb_abi_version() == 1


def test_struct_layout_matches_header():
    from zonos_vibes_b200 import _lib
    assert C.sizeof(_lib.ZbConfig) == 16 * 4
    assert C.sizeof(_lib.ZbSampling) == 12 * 4


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful without a GPU")
def test_no_cpu_fallback_fails_loudly():
    from zonos_vibes_b200 import _lib
    lib = _lib.load()
    zc = _lib.ZbConfig()
    zc.d_model, zc.n_layer, zc.num_heads, zc.num_heads_kv, zc.ffn = 256, 1, 2, 1, 512
    zc.n_codebooks, zc.vocab, zc.rope_table_len = 9, 1026, 16384
    h = C.c_void_p()
    rc = lib.zb_ctx_create(C.byref(zc), 0, C.byref(h))
    assert rc == -2 and b"no CPU fallback" in lib.zb_last_error(None)
    d = C.c_void_p()
    assert lib.zb_dac_create(0, C.byref(d)) == -2
    import zonos_vibes_b200 as z
    with pytest.raises(_lib.ZbError):
        z.apply_delay_pattern(torch.zeros(1, 9, 4, dtype=torch.int64), 1025)
    with pytest.raises(_lib.ZbError):
        z.sample_from_logits(torch.zeros(1, 9, 1026))


def test_bad_config_rejected():
    from zonos_vibes_b200 import _lib
    lib = _lib.load()
    zc = _lib.ZbConfig()
    zc.d_model, zc.n_layer, zc.num_heads, zc.num_heads_kv, zc.ffn = 250, 1, 2, 1, 512      # not a multiple of 256
    zc.n_codebooks, zc.vocab, zc.rope_table_len = 9, 1026, 16384
    h = C.c_void_p()
    assert lib.zb_ctx_create(C.byref(zc), 0, C.byref(h)) == -1
    assert lib.zb_delay_apply(None, None, 1, 9, 4, 1025, None) == -1
    assert lib.zb_delay_apply(None, None, 0, 9, 4, 1025, None) == 0          # empty batch: nothing to do


def test_config_mirror_and_backbone_state_dict_keys():
    import zonos_oracle as zo
    import zonos_vibes_b200 as z
    cfg = zo.TINY_CONFIG
    zc = z.ZonosConfig.from_dict(cfg.to_reference_dict())
    assert zc.eos_token_id == 1024 and zc.masked_token_id == 1025 and zc.backbone.d_model == 256
    sd = zo.synth_state_dict(cfg, 0)
    m = z.Zonos(zc)
    own = set(m.state_dict())
    assert own == set(sd), (own ^ set(sd))                      # same keys as the reference
    for k, v in m.state_dict().items():
        assert tuple(v.shape) == tuple(sd[k].shape), k
    assert "b200" in z.BACKBONES and z.BACKBONES["b200"].supported_architectures == ["transformer", "hybrid"]   # like MambaSSMZonosBackbone (_mamba_ssm.py:9)


def test_shard_indices():
    from zonos_vibes_b200.replicas import shard_indices
    for n in (0, 1, 7, 64):
        for w in (1, 2, 8):
            parts = [shard_indices(n, r, w) for r in range(w)]
            assert sorted(i for p in parts for i in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from zonos_vibes_b200.replicas import shard_indices, gather_ragged, max_over_ranks
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 5
local = {i: torch.full((9, 3 + i), i, dtype=torch.int64) for i in shard_indices(n, rank, world)}
out = gather_ragged(local, n)
m = max_over_ranks(1.0 + rank)
assert m == float(world)
if rank == 0:
    assert [tuple(t.shape) for t in out] == [(9, 3 + i) for i in range(n)]
    assert all(int(t[0, 0]) == i for i, t in enumerate(out))
    print("GLOO_OK")
else:
    assert out is None
dist.barrier()
dist.destroy_process_group()
"""


def test_replica_gather_gloo_world2(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    r = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
         "--master-port", "29533", str(script), ROOT],
        capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "GLOO_OK" in r.stdout


def test_hybrid_backbone_state_dict_keys():
    """Hybrid (Mamba2) holder keeps mamba-ssm's Block / Mamba2 parameter names and shapes."""
    import hybrid_oracle as ho
    import zonos_vibes_b200 as z
    cfg = ho.HYBRID_TINY
    zc = z.ZonosConfig.from_dict(cfg.to_reference_dict())
    bb = z.BACKBONES["b200"](zc.backbone)
    sd = {k[len("backbone."):]: v for k, v in ho.synth_hybrid_state_dict(cfg, 0).items()}
    own = bb.state_dict()
    assert set(own) == set(sd), set(own) ^ set(sd)
    for k, v in own.items():
        assert tuple(v.shape) == tuple(sd[k].shape), k
    assert "layers.0.mixer.conv1d.weight" in own and "layers.2.mlp.fc1.weight" in own and "layers.0.mlp.fc1.weight" not in own


def test_bench_reports_ncu_traffic_only_for_the_captured_configuration():
    """bench.py's roofline.traffic comes from the committed ncu capture of the persistent kernel (batch 1); any other
    configuration reports null rather than a number that was not measured for it."""
    import argparse
    import importlib
    import json
    import os
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    bench = importlib.import_module("bench")
    t = bench.ncu_traffic_per_step(argparse.Namespace(batch=1, opt=[]))
    with open(os.path.join(root, "profiles", "r02_traffic.json")) as f:
        rec = json.load(f)
    assert t == float(rec["dram_bytes_per_step"])
    # traffic must be close to (and not below) the algorithmic bytes: every weight byte read once per step
    assert 0.99 * rec["algorithmic_bytes_per_step_at_this_ctx"] <= t <= 1.05 * rec["algorithmic_bytes_per_step_at_this_ctx"]
    assert bench.ncu_traffic_per_step(argparse.Namespace(batch=64, opt=[])) is None
    assert bench.ncu_traffic_per_step(argparse.Namespace(batch=1, opt=["mega=0"])) is None


def test_from_state_dict_accepts_padded_heads():
    """a state_dict taken from a LOADED reference model has 1026-row heads (post-load pad hook,
    model.py:46-51); from_state_dict must accept it (device 'cpu' only builds the holder: no native call here)."""
    import torch.nn.functional as F
    import zonos_oracle as zo
    import zonos_vibes_b200 as z
    cfg = zo.TINY_CONFIG
    sd = zo.synth_state_dict(cfg, 0)
    padded = {k: (F.pad(v, (0, 0, 0, 1)) if k.startswith("heads.") else v) for k, v in sd.items()}
    assert padded["heads.0.weight"].shape[0] == 1026
    m = z.Zonos.from_state_dict(z.ZonosConfig.from_dict(cfg.to_reference_dict()), padded, "cpu")
    assert m.heads[0].weight.shape[0] == 1025
    assert torch.equal(m.heads[3].weight.detach(), sd["heads.3.weight"])


def test_hybrid_rejects_fp32_residual_and_reads_rotary_cfg():
    import hybrid_oracle as ho
    d = ho.HYBRID_TINY.to_reference_dict()
    assert d["backbone"]["attn_cfg"]["rotary_emb_interleaved"] is False and d["backbone"]["attn_cfg"]["rotary_emb_dim"] == 128


_SHARD_WORKER = r"""
import sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from zonos_vibes_b200.replicas import generate_sharded
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()

class FakeDAC:
    def decode(self, codes):                       # [B, 9, T] -> [B, 1, 512 T]
        return codes.float().sum(1, keepdim=True).repeat_interleave(512, -1)

class FakeModel:                                   # the interface generate_sharded drives (Zonos.generate / autoencoder.decode)
    autoencoder, last_step_ms = FakeDAC(), 1.0
    def generate(self, prefix, max_new_tokens, cfg_scale, batch_size, sampling_params, progress_bar, seed):
        assert prefix.shape[0] == 2 * batch_size
        ids = prefix[:batch_size, 0, 0].long()     # utterance id carried in the conditional row
        assert torch.equal(prefix[batch_size:, 0, 0].long(), -ids - 1)      # cond rows first, then the uncond rows, same order
        self.last_lengths = [max_new_tokens - int(i) % 3 for i in ids]
        return ids[:, None, None].expand(batch_size, 9, max_new_tokens).clone()

n = 7
prefixes = []
for i in range(n):
    p = torch.zeros(2, 3 + i % 2, 4)
    p[0, 0, 0], p[1, 0, 0] = i, -i - 1
    prefixes.append(p)
stats = {}
out = generate_sharded(FakeModel(), prefixes, 10, max_batch=2, stats=stats)
assert stats["frames"] == sum(10 - i % 3 for i in range(rank, n, world))
if rank == 0:
    assert len(out) == n
    for i, w in enumerate(out):
        assert tuple(w.shape) == (1, 512 * (10 - i % 3)) and float(w[0, 0]) == 9.0 * i, (i, w.shape)
    print("SHARD_OK")
else:
    assert out is None
dist.barrier()
dist.destroy_process_group()
"""


def test_generate_sharded_gloo_world2(tmp_path):
    """Config 3's multi-utterance driver: utterance i -> rank i mod world, batched per rank, gathered in
    utterance order on rank 0, no data-path collective - with a stand-in model on CPU (the real one needs a B200)."""
    script = tmp_path / "s.py"
    script.write_text(_SHARD_WORKER)
    r = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
         "--master-port", "29541", str(script), ROOT],
        capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "SHARD_OK" in r.stdout


def test_prefix_conditioner_mirror_keys_and_cond_dict():
    """Row f1 host side: the conditioner mirror takes the reference's state_dict by key (frozen from the reference in
    golden_long.pt `cond.state_dict`), make_cond_dict reproduces conditioning.py:331-405's keys / shapes, and the ops fail
    loudly on CPU tensors (no fallback)."""
    import zonos_vibes_b200 as z
    from zonos_vibes_b200 import _lib
    from zonos_vibes_b200.conditioning import make_cond_dict, supported_language_codes
    g = torch.load(os.path.join(ROOT, "tests", "golden", "golden_long.pt"))
    pc = z.PrefixConditioner(z.PrefixConditionerConfig(**g["cond.config"]), 256)
    assert set(pc.state_dict()) == set(g["cond.state_dict"])
    for k, v in pc.state_dict().items():
        assert v.shape == g["cond.state_dict"][k].shape, k
    assert pc.required_keys == {"espeak"}
    with pytest.raises(ValueError, match="Missing required keys"):
        pc({})
    pc.load_state_dict(g["cond.state_dict"])
    with pytest.raises(_lib.ZbError):
        pc({"espeak": ("ids", g["cond.phoneme_ids"])})

    cd = make_cond_dict(text="hi", language="en-us", device="cpu")
    assert set(cd) == {"espeak", "speaker", "emotion", "fmax", "pitch_std", "speaking_rate", "language_id", "ctc_loss", "speaker_noised"}
    assert cd["espeak"] == (["hi"], ["en-us"]) and cd["speaker"] is None
    assert cd["emotion"].shape == (1, 1, 8) and abs(cd["emotion"].sum().item() - 1.0) < 1e-6
    assert cd["language_id"].item() == 24 == supported_language_codes.index("en-us") and len(supported_language_codes) == 109
    assert make_cond_dict(device="cpu", phoneme_ids=torch.zeros(1, 3, dtype=torch.int64))["espeak"][0] == "ids"
    mcd = g["cond.mcd_inputs"]
    for k in ("emotion", "fmax", "pitch_std", "speaking_rate", "language_id"):
        assert cd[k].shape == mcd[k].shape and torch.allclose(cd[k].float(), mcd[k].float(), atol=1e-6), k


def test_library_staleness_is_by_content(tmp_path, monkeypatch):
    """_lib.load() rebuilds a library that was built from other sources: staleness is a content hash of csrc/ +
    the header recorded next to the .so, so a copied tree (the GPU box's snapshot) with new timestamps stays current."""
    from zonos_vibes_b200 import build
    assert os.path.exists(build.LIB) and not build._stale()
    fake = tmp_path / "stamp"
    fake.write_text("0" * 64 + "\n")
    monkeypatch.setattr(build, "STAMP", str(fake))
    assert build._stale()
    fake.write_text(build._source_hash() + "\n")
    assert not build._stale()


def test_library_override_by_environment(monkeypatch):
    """ZB_LIB names another build of the library (scripts/build_checked.sh); it goes through the same symbol / ABI checks."""
    from zonos_vibes_b200 import _lib
    prod = _lib.load()
    monkeypatch.setenv("ZB_LIB", _lib.LIB_PATH)
    monkeypatch.setattr(_lib, "_lib", None)
    other = _lib.load()
    assert other is not prod and other.zb_abi_version() == 1
    monkeypatch.setenv("ZB_LIB", "/nonexistent/libzonos_b200_checked.so")
    monkeypatch.setattr(_lib, "_lib", None)
    with pytest.raises(OSError):
        _lib.load()


def test_serving_shell_batches_and_wav_container():
    """Row f2 host side: the batching engine groups compatible requests (same conditioning length and sampling parameters) into
    one generate call of at most max_batch utterances, keeps arrival order, and resolves every future; the WAV container is
    byte-identical to scipy.io.wavfile.write on int16 mono (server.py:149-152); tts_request keeps the reference's validation."""
    import io
    import numpy as np
    import scipy.io.wavfile
    from zonos_vibes_b200 import serving

    pcm = torch.randint(-32768, 32767, (5000,), dtype=torch.int16)
    buf = io.BytesIO()
    scipy.io.wavfile.write(buf, 44100, pcm.numpy())
    assert serving.wav_bytes(pcm, 44100) == buf.getvalue()

    class FakeModel:
        autoencoder = None
        device = torch.device("cpu")

        def __init__(self):
            self.calls = []

        def generate(self, prefix, max_new_tokens, cfg_scale, batch_size, sampling_params, progress_bar, seed):
            assert prefix.shape[0] == 2 * batch_size
            self.calls.append((batch_size, prefix.shape[1], max_new_tokens))
            # utterance i's codes carry its conditional row's first value, so that routing can be checked
            ids = prefix[:batch_size, 0, 0].to(torch.int64)
            self.last_lengths = [3 + int(i) % 2 for i in ids]
            return ids.view(-1, 1, 1).expand(batch_size, 9, 5).contiguous()

    m = FakeModel()
    eng = serving.BatchingEngine(m, max_batch=4, max_wait_ms=200.0)
    futs = []
    for i in range(6):                                   # six requests with Lc = 7, two with Lc = 9, one with other sampling
        futs.append(eng.submit(torch.full((2, 7, 8), float(i)), 10))
    for i in range(6, 8):
        futs.append(eng.submit(torch.full((2, 9, 8), float(i)), 10))
    futs.append(eng.submit(torch.full((2, 7, 8), 8.0), 10, sampling_params=dict(min_p=0.1)))
    res = [f.result(timeout=30) for f in futs]
    eng.close()
    for i, r in enumerate(res):
        assert r.codes.shape == (9, 3 + i % 2) and int(r.codes[0, 0]) == i and r.pcm is None
    assert sorted(eng.batches) == [1, 2, 2, 4] and max(c[0] for c in m.calls) == 4
    assert [r.batch_size for r in res[:4]] == [4, 4, 4, 4]
    with pytest.raises(ValueError):
        eng.submit(torch.zeros(3, 7, 8))

    eng2 = serving.BatchingEngine(m, max_batch=2, max_wait_ms=1.0)
    spk = lambda p: (_ for _ in ()).throw(FileNotFoundError(p))     # noqa: E731
    assert serving.tts_request(eng2, None, spk)[0] == 400
    assert serving.tts_request(eng2, {"speaker_audio_path": "a", "speaking_rate": 1}, spk) == (400, {"error": "Missing 'text' parameter"})
    assert serving.tts_request(eng2, {"text": "t", "speaking_rate": 1}, spk)[1] == {"error": "Missing 'speaker_audio_path' parameter"}
    assert serving.tts_request(eng2, {"text": "t", "speaker_audio_path": "a"}, spk)[1] == {"error": "Missing 'speaking_rate' parameter"}
    assert serving.tts_request(eng2, {"text": "t", "speaker_audio_path": "a", "speaking_rate": "x"}, spk)[1] == {"error": "'speaking_rate' must be a number"}
    assert serving.tts_request(eng2, {"text": "t", "speaker_audio_path": "missing.wav", "speaking_rate": 15}, spk)[0] == 404
    eng2.close()


def test_speaker_mirror_state_dict_keys():
    """Row f4 host side: the mirror module takes the reference's ResNet293_SimAM_ASP state_dict by key (the oracle's synthetic
    state_dict has exactly the keys of the reference's ResNet293_based, see oracle/make_golden_spk.py) and refuses CPU tensors."""
    import speaker_oracle as so
    from zonos_vibes_b200 import _lib, speaker_cloning as sc
    m = sc.ResNet293_based(num_blocks=(2, 2, 3, 1))
    sd = so.synth_speaker_state_dict(5, (2, 2, 3, 1))
    assert set(m.state_dict()) == set(sd)
    for k, v in m.state_dict().items():
        assert v.shape == sd[k].shape, k
    m.load_state_dict(sd)
    with pytest.raises(_lib.ZbError):
        m.forward_features(so.synth_features(1, 16, 8))
    full = sc.ResNet293_based()
    assert sum(len(layer) for layer in (full.front.layer1, full.front.layer2, full.front.layer3, full.front.layer4)) == 97
