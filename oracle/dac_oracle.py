"""CPU oracle for the DAC 44 kHz decode path (`model.autoencoder.decode(codes)`).

TEST INFRASTRUCTURE — NOT PRODUCT CODE (same rules as `zonos_oracle.py`).

The reference's `DACAutoencoder.decode` (zonos/autoencoder.py:25-27) delegates to the third-party
`transformers` package (pinned 4.48.3 in the reference's uv.lock:1937-1938; 5.5.0 in this image),
`DacModel.decode` — models/dac/modeling_dac.py:608-638 → `DacResidualVectorQuantizer.from_codes`
(:345-369) → `DacDecoder.forward` (:405-439) → `DacDecoderBlock` (:234-262) → `DacResidualUnit`
(:173-207) → `Snake1d` (:85-99).  That arithmetic is restated here functionally, fp32 on CPU (the
reference disables autocast on CPU), over an explicit HF-keyed state_dict.  It is pinned live against
the installed `transformers.DacModel` in `tests/test_oracle_pinned.py` (transformers is part of the
image on both boxes) and against golden vectors in `tests/golden/`.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

N_CODEBOOKS = 9
CODEBOOK_SIZE = 1024
CODEBOOK_DIM = 8
HIDDEN = 1024                     # DacConfig.hidden_size
DECODER_HIDDEN = 1536             # DacConfig.decoder_hidden_size
UPSAMPLING = (8, 8, 4, 2)         # DacConfig.upsampling_ratios
HOP = 512
SAMPLING_RATE = 44100
DILATIONS = (1, 3, 9)


def synth_dac_state_dict(seed: int = 2) -> dict:
    """Deterministic random decoder/quantizer weights under HF `DacModel` keys.
    Scales are chosen so activations stay O(1) through the 30 conv layers and the waveform has
    O(0.3) amplitude — a 1e-2 max-abs gate is then a meaningful test (default HF init gives 1e-2 peaks)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    sd: dict = {}

    def conv(prefix, cout, cin, k, gain=1.0, transposed=False):
        fan_in = cin * k
        if transposed:      # ConvTranspose1d weight is [Cin, Cout, k]; each output sees k/stride taps
            shape, fan_in = (cin, cout, k), cin * 2
        else:
            shape = (cout, cin, k)
        b = gain * math.sqrt(3.0 / fan_in)
        sd[prefix + ".weight"] = (torch.rand(shape, generator=g) * 2 - 1) * b
        sd[prefix + ".bias"] = 0.05 * torch.randn(cout, generator=g)

    def snake(prefix, c):
        sd[prefix + ".alpha"] = 0.5 + torch.rand(1, c, 1, generator=g)

    for k in range(N_CODEBOOKS):
        q = f"quantizer.quantizers.{k}"
        sd[q + ".codebook.weight"] = torch.randn(CODEBOOK_SIZE, CODEBOOK_DIM, generator=g)
        conv(q + ".out_proj", HIDDEN, CODEBOOK_DIM, 1, gain=1.0 / 3.0)
    conv("decoder.conv1", DECODER_HIDDEN, HIDDEN, 7)
    c = DECODER_HIDDEN
    for j, s in enumerate(UPSAMPLING):
        p = f"decoder.block.{j}"
        snake(p + ".snake1", c)
        conv(p + ".conv_t1", c // 2, c, 2 * s, transposed=True)
        c //= 2
        for u in (1, 2, 3):
            r = f"{p}.res_unit{u}"
            snake(r + ".snake1", c)
            conv(r + ".conv1", c, c, 7, gain=0.7)
            snake(r + ".snake2", c)
            conv(r + ".conv2", c, c, 1, gain=0.5)
    snake("decoder.snake1", c)
    conv("decoder.conv2", 1, c, 7, gain=0.12)
    return sd


def snake1d(x, alpha):
    """modeling_dac.py:85-99: x + (alpha + 1e-9)^-1 * sin(alpha x)^2, alpha [1,C,1]."""
    return x + (alpha + 1e-9).reciprocal() * torch.sin(alpha * x).pow(2)


def rvq_from_codes(sd: dict, codes: torch.Tensor) -> torch.Tensor:
    """modeling_dac.py:345-369: sum_k out_proj_k(codebook_k[codes[:,k]]^T) -> [B, 1024, T]."""
    z = 0.0
    for k in range(codes.shape[1]):
        q = f"quantizer.quantizers.{k}"
        lat = F.embedding(codes[:, k, :], sd[q + ".codebook.weight"]).transpose(1, 2)
        z = z + F.conv1d(lat, sd[q + ".out_proj.weight"], sd[q + ".out_proj.bias"])
    return z


def residual_unit(sd: dict, prefix: str, x, dilation: int):
    """modeling_dac.py:173-207."""
    y = F.conv1d(snake1d(x, sd[prefix + ".snake1.alpha"]), sd[prefix + ".conv1.weight"], sd[prefix + ".conv1.bias"],
                 dilation=dilation, padding=3 * dilation)
    y = F.conv1d(snake1d(y, sd[prefix + ".snake2.alpha"]), sd[prefix + ".conv2.weight"], sd[prefix + ".conv2.bias"])
    return x + y


def decoder_forward(sd: dict, z: torch.Tensor, collect: dict | None = None) -> torch.Tensor:
    """modeling_dac.py:405-439 + :234-262."""
    x = F.conv1d(z, sd["decoder.conv1.weight"], sd["decoder.conv1.bias"], padding=3)
    if collect is not None:
        collect["conv1"] = x
    for j, s in enumerate(UPSAMPLING):
        p = f"decoder.block.{j}"
        x = snake1d(x, sd[p + ".snake1.alpha"])
        x = F.conv_transpose1d(x, sd[p + ".conv_t1.weight"], sd[p + ".conv_t1.bias"], stride=s,
                               padding=math.ceil(s / 2))
        if collect is not None:
            collect[f"block{j}.conv_t1"] = x
        for u, dil in zip((1, 2, 3), DILATIONS):
            x = residual_unit(sd, f"{p}.res_unit{u}", x, dil)
        if collect is not None:
            collect[f"block{j}"] = x
    x = snake1d(x, sd["decoder.snake1.alpha"])
    x = F.conv1d(x, sd["decoder.conv2.weight"], sd["decoder.conv2.bias"], padding=3)
    return torch.tanh(x)


@torch.inference_mode()
def dac_decode(sd: dict, codes: torch.Tensor, collect: dict | None = None) -> torch.Tensor:
    """zonos/autoencoder.py:25-27 on CPU: codes int64 [B,9,T] -> fp32 [B,1,512*T]."""
    z = rvq_from_codes(sd, codes)
    if collect is not None:
        collect["latent"] = z
    return decoder_forward(sd, z, collect).float()


def load_into_hf(sd: dict):
    """Build the installed transformers `DacModel` with these weights (pinning only)."""
    from transformers.models.dac import DacConfig, DacModel

    m = DacModel(DacConfig(sampling_rate=SAMPLING_RATE)).eval().requires_grad_(False)
    full = m.state_dict()
    for k, v in sd.items():
        assert k in full and full[k].shape == v.shape, (k, tuple(v.shape))
        full[k] = v
    m.load_state_dict(full)
    return m
