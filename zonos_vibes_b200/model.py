"""`Zonos` — the generation driver, API-compatible with the reference's zonos/model.py:22-315.

`generate` keeps the reference's signature and return value, but the whole loop body (embedding gather-sum,
26 transformer layers, 9 heads + CFG, logit bias, sampling, EOS wind-down, frame write-back) runs on the device as
one CUDA graph per step inside libzonos_b200; the host only polls a done flag every `steps_per_poll` steps.
"""
from __future__ import annotations

import ctypes as C
import json
from typing import Callable

import torch
import torch.nn as nn

from . import _lib
from .autoencoder import DACAutoencoder
from .backbone import BACKBONES, N_CODEBOOKS, VOCAB, B200ZonosBackbone
from .codebook_pattern import apply_delay_pattern, revert_delay_pattern
from .config import InferenceParams, ZonosConfig


def find_multiple(n: int, k: int) -> int:
    """zonos/utils.py:6-9."""
    return n if (k == 0 or n % k == 0) else n + k - (n % k)


class Zonos(nn.Module):
    def __init__(self, config: ZonosConfig, backbone_cls=B200ZonosBackbone, autoencoder: DACAutoencoder | None = None,
                 prefix_conditioner: nn.Module | None = None):
        super().__init__()
        self.config = config
        dim = config.backbone.d_model
        self.eos_token_id = config.eos_token_id
        self.masked_token_id = config.masked_token_id
        self.autoencoder = autoencoder            # not an nn.Module, like the reference (model.py:30)
        self.backbone = backbone_cls(config.backbone)
        self.backbone.eos_token_id, self.backbone.masked_token_id = config.eos_token_id, config.masked_token_id
        # the device-side mirror of the reference's PrefixConditioner (conditioning.py) is built whenever the
        # config lists conditioners, so a checkpoint's `prefix_conditioner.*` entries load; any module with the same
        # interface (conditioning.py:297-313) can be passed in instead.
        if prefix_conditioner is None and config.prefix_conditioner.conditioners:
            from .conditioning import PrefixConditioner
            prefix_conditioner = PrefixConditioner(config.prefix_conditioner, dim)
        self.prefix_conditioner = prefix_conditioner
        self.spk_clone_model = None
        self.embeddings = nn.ModuleList([nn.Embedding(1026, dim) for _ in range(N_CODEBOOKS)])
        self.heads = nn.ModuleList([nn.Linear(dim, 1025, bias=False) for _ in range(N_CODEBOOKS)])
        self._heads_synced = False
        self.last_lengths: list[int] | None = None
        self.last_step_ms: float | None = None
        self.last_delayed_codes: torch.Tensor | None = None

    # ---- construction ---------------------------------------------------------------------------
    @property
    def device(self) -> torch.device:
        return next(self.parameters()).device

    @classmethod
    def from_state_dict(cls, config: ZonosConfig, state_dict: dict, device="cuda", dac_state_dict: dict | None = None,
                        prefix_conditioner: nn.Module | None = None) -> "Zonos":
        """Build from a reference state_dict (SURVEY App. B keys).  `prefix_conditioner.*` entries load into the conditioner
        (built from config.prefix_conditioner, or the module supplied) and may be absent from a backbone-only state_dict."""
        dac = DACAutoencoder(dac_state_dict, device) if dac_state_dict is not None else None
        model = cls(config, autoencoder=dac, prefix_conditioner=prefix_conditioner).to(device, torch.bfloat16)
        own = model.state_dict()
        sd = {k: v for k, v in state_dict.items() if k in own}
        for k in range(N_CODEBOOKS):
            # a state_dict taken from a LOADED reference model has 1026-row heads (post-load pad hook, model.py:46-51 /
            # utils.py:12-27: one zero row appended); the holder keeps the 1025 real rows and the library re-pads
            w = sd.get(f"heads.{k}.weight")
            if w is not None and w.shape[0] == VOCAB:
                sd[f"heads.{k}.weight"] = w[: VOCAB - 1]
        missing = [k for k in own if k not in sd and not k.startswith("prefix_conditioner.")]
        if missing:
            raise KeyError(f"state_dict lacks {missing[:4]}{'...' if len(missing) > 4 else ''}")
        model.load_state_dict(sd, strict=False)
        model.backbone.invalidate()
        model._heads_synced = False
        return model.eval()

    @classmethod
    def from_local(cls, config_path: str, model_path: str, device: str = "cuda", backbone: str | None = None) -> "Zonos":
        """model.py:66-88.  `backbone` selects from this package's BACKBONES (only "b200")."""
        import safetensors

        config = ZonosConfig.from_dict(json.load(open(config_path)))
        if backbone is not None and backbone not in BACKBONES:
            raise KeyError(f"unknown backbone {backbone!r}; available: {list(BACKBONES)}")
        sd = {}
        with safetensors.safe_open(model_path, framework="pt") as f:
            for k in f.keys():
                sd[k] = f.get_tensor(k)
        return cls.from_state_dict(config, sd, device)

    @classmethod
    def from_pretrained(cls, repo_id: str, revision: str | None = None, device: str = "cuda", **kwargs) -> "Zonos":
        """model.py:57-64."""
        from huggingface_hub import hf_hub_download

        config_path = hf_hub_download(repo_id=repo_id, filename="config.json", revision=revision)
        model_path = hf_hub_download(repo_id=repo_id, filename="model.safetensors", revision=revision)
        return cls.from_local(config_path, model_path, device, **kwargs)

    def _native(self):
        """Native context with embeddings + heads (padded 1025 -> 1026 rows like model.py:46-51) pushed in."""
        if not self._heads_synced or self.backbone._ctx is None:
            extra = {}
            for k in range(N_CODEBOOKS):
                extra[f"embeddings.{k}.weight"] = self.embeddings[k].weight
                extra[f"heads.{k}.weight"] = self.heads[k].weight       # 1025 rows; the library zero-pads row 1025
            self.backbone._extra_weights = extra
            self.backbone.invalidate()
            self._heads_synced = True
        return self.backbone.native()

    # ---- reference API ---------------------------------------------------------------------------
    def setup_cache(self, batch_size: int, max_seqlen: int, dtype: torch.dtype = torch.bfloat16) -> InferenceParams:
        """model.py:198-202."""
        max_seqlen = find_multiple(max_seqlen, 8)
        self._native()
        kv = self.backbone.allocate_inference_cache(batch_size, max_seqlen, dtype=dtype)
        lengths = torch.zeros((batch_size,), dtype=torch.int32, device=self.device)
        return InferenceParams(max_seqlen, batch_size, 0, 0, kv, lengths)

    def make_speaker_embedding(self, wav: torch.Tensor, sr: int) -> torch.Tensor:
        """model.py:90-95: speaker embedding of an audio clip ([1, 1, 128] bf16) from the ResNet293-SimAM-ASP + LDA network
        (zonos_vibes_b200/speaker_cloning.py).  The checkpoints come from the hub like the reference's, or
        assign `model.spk_clone_model = SpeakerEmbeddingLDA(device, spk_model_path=..., lda_spk_model_path=...)` first."""
        if self.spk_clone_model is None:
            from .speaker_cloning import SpeakerEmbeddingLDA
            self.spk_clone_model = SpeakerEmbeddingLDA(str(self.device))
        _, spk_embedding = self.spk_clone_model(wav.to(self.spk_clone_model.device), sr)
        return spk_embedding.unsqueeze(0).bfloat16()

    def prepare_conditioning(self, cond_dict: dict, uncond_dict: dict | None = None) -> torch.Tensor:
        """model.py:204-212."""
        if self.prefix_conditioner is None:
            raise _lib.ZbError("no prefix_conditioner attached: pass prefix_conditioning tensors to generate() directly, "
                               "or attach the reference's PrefixConditioner module (INTEGRATION.md)")
        if uncond_dict is None:
            uncond_dict = {k: cond_dict[k] for k in self.prefix_conditioner.required_keys}
        return torch.cat([self.prefix_conditioner(cond_dict), self.prefix_conditioner(uncond_dict)])

    def can_use_cudagraphs(self) -> bool:
        return True

    @torch.inference_mode()
    def compute_logits_teacher_forced(self, prefix_conditioning: torch.Tensor, delayed_codes: torch.Tensor, n_steps: int,
                                      cfg_scale: float = 2.0, skip: int = 0) -> torch.Tensor:
        """Parity helper: logits of the prefill and of `n_steps` decode steps when the model is fed the given,
        fully known delayed codes [B, 9, L] (= what `_prefill` / `_decode_one_token` + logit_bias return in
        model.py:253-280).  Returns fp32 [1 + n_steps, B, 9, 1026].  `skip` = number of leading audio-prefix columns
        that are part of the prefill (model.py:253-255 with audio_prefix_codes of `skip` frames): they are fed through
        the same decode path, their logits are dropped, and entry 0 of the result is the logits of column `skip`."""
        ctx = self._native()
        b = delayed_codes.shape[0]
        delayed = delayed_codes.to(ctx.device, torch.int64).contiguous()
        prefix = prefix_conditioning.to(ctx.device, torch.bfloat16).contiguous()
        lc = prefix.shape[1]
        self.setup_cache(2 * b, lc + delayed.shape[2] + 9)
        out = torch.empty((1 + n_steps, b, N_CODEBOOKS, VOCAB), dtype=torch.float32, device=ctx.device)
        scratch = torch.empty((b, N_CODEBOOKS, VOCAB), dtype=torch.float32, device=ctx.device)
        with torch.cuda.device(ctx.device):
            ctx.check(ctx.lib.zb_backbone_forward(ctx.handle, prefix.data_ptr(), None, 2 * b, lc, 0, None, ctx.stream))
            for s in range(skip + 1 + n_steps):
                dst = out[s - skip] if s >= skip else scratch
                ctx.check(ctx.lib.zb_decode_logits(ctx.handle, delayed.data_ptr(), b, delayed.shape[2], s, float(cfg_scale),
                                                   1 if s > skip else 0, dst.data_ptr(), ctx.stream))
        return out

    @torch.inference_mode()
    def generate(
        self,
        prefix_conditioning: torch.Tensor,  # [2*bsz, cond_seq_len, d_model]
        audio_prefix_codes: torch.Tensor | None = None,  # [bsz, 9, prefix_audio_seq_len]
        max_new_tokens: int = 86 * 30,
        cfg_scale: float = 2.0,
        batch_size: int = 1,
        sampling_params: dict = dict(min_p=0.1),
        progress_bar: bool = True,
        disable_torch_compile: bool = False,   # accepted for signature compatibility; nothing is traced here
        callback: Callable[[torch.Tensor, int, int], bool] | None = None,
        *,
        seed: int | None = None,
        compat_scatter: bool = True,
        steps_per_poll: int = 64,
    ) -> torch.Tensor:
        """model.py:218-315.  Returns int64 [bsz, 9, T'].  For bsz > 1 (which the reference cannot run)
        utterances are independent batch-1 runs: T' is the longest, shorter ones are zero-padded and their lengths
        are left in `self.last_lengths`."""
        assert cfg_scale != 1, "TODO: add support for cfg_scale=1"
        ctx = self._native()
        lib = ctx.lib
        device = ctx.device
        b = batch_size
        prefix = prefix_conditioning.to(device, torch.bfloat16).contiguous()
        if prefix.shape[0] != 2 * b:
            raise _lib.ZbError(f"prefix_conditioning has {prefix.shape[0]} rows, expected 2*batch_size = {2 * b}")
        lc = prefix.shape[1]
        prefix_audio_len = 0 if audio_prefix_codes is None else audio_prefix_codes.shape[2]
        audio_seq_len = prefix_audio_len + max_new_tokens
        seq_len = lc + audio_seq_len + 9

        self.setup_cache(batch_size=2 * b, max_seqlen=seq_len)
        codes = torch.full((b, 9, audio_seq_len), -1, dtype=torch.int64, device=device)
        if audio_prefix_codes is not None:
            codes[..., :prefix_audio_len] = audio_prefix_codes.to(device)
        delayed_codes = apply_delay_pattern(codes, self.masked_token_id)
        length = delayed_codes.shape[2]
        offset0 = prefix_audio_len + 1
        max_steps = length - offset0

        sp = _lib.make_sampling(**sampling_params)
        if seed is None:
            seed = int(torch.randint(0, 2**62, (1,)).item())
        done, col = C.c_int(0), C.c_int(0)
        out_len = (C.c_int32 * b)()
        try:
            from tqdm import tqdm
            progress = tqdm(total=max_steps, desc="Generating", disable=not progress_bar)
        except ImportError:  # pragma: no cover
            progress = None
        with torch.cuda.device(device):
            stream = ctx.stream
            ctx.check(lib.zb_gen_begin(ctx.handle, prefix.data_ptr(), b, lc, delayed_codes.data_ptr(), length, offset0,
                                       float(cfg_scale), C.byref(sp), int(seed), 1 if compat_scatter else 0, stream))
            # Two zb_gen_steps calls are kept in flight: the host waits for call k (pinned snapshot of the loop state)
            # while call k+1 already runs, so the device never idles between persistent-kernel launches.  A call issued
            # after the utterances have finished returns at once (device-side done flag).
            depth = 1 if callback is not None else 2
            step, aborted, total_ms, issued, in_flight = 0, False, 0.0, 0, 0
            while True:
                while in_flight < depth and (issued < max_steps or in_flight == 0):
                    n = 1 if callback is not None else max(1, min(steps_per_poll, max_steps - issued))
                    ctx.check(lib.zb_gen_steps(ctx.handle, n, stream))
                    issued += n
                    in_flight += 1
                ctx.check(lib.zb_gen_poll(ctx.handle, C.byref(done), C.byref(col), out_len, stream))
                in_flight -= 1
                total_ms += max(0.0, float(lib.zb_last_steps_ms(ctx.handle)))
                new_step = col.value - offset0
                if progress is not None:
                    progress.update(new_step - step)
                step = new_step
                if callback is not None:
                    frame = delayed_codes[..., col.value: col.value + 1]
                    if not callback(frame, step, max_steps):
                        aborted = True
                        break
                if done.value or step >= max_steps:
                    break
            while in_flight > 0:                  # drain (the leftover call saw the done flag and did nothing)
                ctx.check(lib.zb_gen_poll(ctx.handle, C.byref(done), C.byref(col), out_len, stream))
                in_flight -= 1
        if progress is not None:
            progress.close()
        self.last_step_ms = total_ms / max(step, 1)
        self.last_delayed_codes = delayed_codes          # debug / parity: the raw delayed buffer of this run

        out_codes = revert_delay_pattern(delayed_codes)
        out_codes.masked_fill_(out_codes >= 1024, 0)
        out_codes.masked_fill_(out_codes < 0, 0)
        lens = [int(out_len[i]) if (out_len[i] > 0 or not aborted) else 0 for i in range(b)]
        if aborted:
            lens = [l if l > 0 else max(col.value - 9, 0) for l in lens]
        self.last_lengths = lens
        t_out = max(lens) if lens else 0
        out_codes = out_codes[..., :t_out].contiguous()
        for i, l in enumerate(lens):
            if l < t_out:
                out_codes[i, :, l:] = 0
        return out_codes
