"""DAC 44 kHz codec — mirror of the reference's zonos/autoencoder.py:8-27 `DACAutoencoder` (decode: the hot path; preprocess /
encode: SURVEY §8 row f3, the audio-prefix input side)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


class DACAutoencoder:
    """`decode(codes)` runs the RVQ gather-sum and the 30 conv layers + 29 Snake activations in libzonos_b200
    (tcgen05 implicit-GEMM, fp16 operands like the reference's fp16 autocast, fp32 accumulate).

    Construction differs from the reference only in where the weights come from: pass an HF `DacModel`
    state_dict (`descript/dac_44khz`; the reference loads it from the hub, autoencoder.py:11) or call
    `from_hf_model(dac_model)`.
    """

    codebook_size = 1024      # dac.config.codebook_size
    num_codebooks = 9         # dac.quantizer.n_codebooks
    sampling_rate = 44100     # dac.config.sampling_rate
    hop_length = 512

    def __init__(self, state_dict: dict, device: torch.device | str = "cuda"):
        self.lib = _lib.load()
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _lib.ZbError("zonos_vibes_b200 needs a CUDA (sm_100) device; there is no CPU fallback")
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        self.handle = C.c_void_p()
        _lib.check(self.lib.zb_dac_create(idx, C.byref(self.handle)), None, dac=True)
        self.impl = 0          # 0 = tcgen05 fp16, 1 = SIMT fp32 cross-check
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device)
            for name, t in state_dict.items():
                if not (name.startswith("decoder.") or name.startswith("quantizer.") or name.startswith("encoder.")):
                    continue           # (encoder.* and quantizer.*.in_proj.* are optional: without them encode() raises)
                t = t.detach().to(self.device, torch.float32).contiguous()
                self._check(self.lib.zb_dac_set_tensor(self.handle, name.encode(), t.data_ptr(), t.numel(), stream.cuda_stream))
                stream.synchronize()
            self._check(self.lib.zb_dac_finalize(self.handle, stream.cuda_stream))

    @classmethod
    def from_hf_model(cls, dac_model, device="cuda") -> "DACAutoencoder":
        return cls(dac_model.state_dict(), device)

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self.lib.zb_dac_destroy(self.handle)
                self.handle = None
        except Exception:  # noqa: BLE001
            pass

    def _check(self, rc: int) -> None:
        _lib.check(rc, self.handle, dac=True)

    def set_option(self, key: str, value: int) -> None:
        self._check(self.lib.zb_dac_set_option(self.handle, key.encode(), int(value)))

    def preprocess(self, wav: torch.Tensor, sr: int) -> torch.Tensor:
        """autoencoder.py:17-20: resample to 44.1 kHz (torchaudio, like the reference) and zero-pad on the right to a multiple of 512."""
        import math

        import torchaudio
        wav = torchaudio.functional.resample(wav, sr, 44_100)
        right_pad = math.ceil(wav.shape[-1] / 512) * 512 - wav.shape[-1]
        return torch.nn.functional.pad(wav, (0, right_pad))

    @torch.inference_mode()
    def encode(self, wav: torch.Tensor) -> torch.Tensor:
        """wav fp32 [B, 1, S] (S % 512 == 0, see preprocess) -> codes int64 [B, 9, S / 512]   (autoencoder.py:22-23)."""
        if wav.dim() != 3 or wav.shape[1] != 1:
            raise _lib.ZbError(f"expected a waveform of shape [B, 1, S], got {tuple(wav.shape)}")
        wav = wav.to(self.device, torch.float32).contiguous()
        b, _, n = wav.shape
        codes = torch.empty((b, self.num_codebooks, n // self.hop_length), dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            self._check(self.lib.zb_dac_encode(self.handle, wav.data_ptr(), codes.data_ptr(), b, n,
                                               torch.cuda.current_stream(self.device).cuda_stream))
        return codes

    @torch.inference_mode()
    def decode(self, codes: torch.Tensor) -> torch.Tensor:
        """codes int64 [B, 9, T] in [0, 1023] -> fp32 [B, 1, 512*T]   (autoencoder.py:25-27)."""
        if not codes.is_cuda:
            codes = codes.to(self.device)
        codes = codes.to(torch.int64).contiguous()
        b, nq, t = codes.shape
        if nq != self.num_codebooks:
            raise _lib.ZbError(f"expected {self.num_codebooks} codebooks, got {nq}")
        wav = torch.empty((b, 1, self.hop_length * t), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self._check(self.lib.zb_dac_decode(self.handle, codes.data_ptr(), wav.data_ptr(), b, t, int(self.impl),
                                               torch.cuda.current_stream(self.device).cuda_stream))
        return wav

    def intermediate(self, name: str, channels: int, length: int) -> torch.Tensor:
        """fp32 [C, T_l] copy of an intermediate of the LAST utterance decoded with keep_intermediates=1."""
        out = torch.empty((channels, length), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self._check(self.lib.zb_dac_get_intermediate(self.handle, name.encode(), out.data_ptr(), out.numel(),
                                                         torch.cuda.current_stream(self.device).cuda_stream))
        return out

    def launch_count(self) -> int:
        return int(self.lib.zb_dac_launch_count(self.handle))
