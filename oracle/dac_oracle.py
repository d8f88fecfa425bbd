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


# ---- encoder side: DACAutoencoder.encode, zonos/autoencoder.py:22-23 ------------------------------------
ENCODER_HIDDEN = 64               # DacConfig.encoder_hidden_size
DOWNSAMPLING = (2, 4, 8, 8)       # DacConfig.downsampling_ratios


def synth_dac_encoder_state_dict(seed: int = 3) -> dict:
    """Deterministic random encoder + quantizer in_proj weights under HF `DacModel` keys; same scale rules as
    synth_dac_state_dict (activations stay O(1) through the 29 conv layers)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    sd: dict = {}

    def conv(prefix, cout, cin, k, gain=1.0):
        b = gain * math.sqrt(3.0 / (cin * k))
        sd[prefix + ".weight"] = (torch.rand((cout, cin, k), generator=g) * 2 - 1) * b
        sd[prefix + ".bias"] = 0.05 * torch.randn(cout, generator=g)

    def snake(prefix, c):
        sd[prefix + ".alpha"] = 0.5 + torch.rand(1, c, 1, generator=g)

    conv("encoder.conv1", ENCODER_HIDDEN, 1, 7, gain=3.0)
    c = ENCODER_HIDDEN
    for j, s in enumerate(DOWNSAMPLING):
        p = f"encoder.block.{j}"
        for u in (1, 2, 3):
            r = f"{p}.res_unit{u}"
            snake(r + ".snake1", c)
            conv(r + ".conv1", c, c, 7, gain=0.7)
            snake(r + ".snake2", c)
            conv(r + ".conv2", c, c, 1, gain=0.5)
        snake(p + ".snake1", c)
        conv(p + ".conv1", 2 * c, c, 2 * s)
        c *= 2
    snake("encoder.snake1", c)
    conv("encoder.conv2", HIDDEN, c, 3)
    for k in range(N_CODEBOOKS):
        conv(f"quantizer.quantizers.{k}.in_proj", CODEBOOK_DIM, HIDDEN, 1)
    return sd


def encoder_forward(sd: dict, wav: torch.Tensor) -> torch.Tensor:
    """modeling_dac.py DacEncoder.forward / DacEncoderBlock.forward: wav [B, 1, S] -> latents [B, 1024, S / 512]."""
    x = F.conv1d(wav, sd["encoder.conv1.weight"], sd["encoder.conv1.bias"], padding=3)
    for j, s in enumerate(DOWNSAMPLING):
        p = f"encoder.block.{j}"
        for u, dil in zip((1, 2, 3), DILATIONS):
            x = residual_unit(sd, f"{p}.res_unit{u}", x, dil)
        x = snake1d(x, sd[p + ".snake1.alpha"])
        x = F.conv1d(x, sd[p + ".conv1.weight"], sd[p + ".conv1.bias"], stride=s, padding=math.ceil(s / 2))
    x = snake1d(x, sd["encoder.snake1.alpha"])
    return F.conv1d(x, sd["encoder.conv2.weight"], sd["encoder.conv2.bias"], padding=1)


def rvq_encode(sd: dict, z: torch.Tensor):
    """DacResidualVectorQuantizer.forward (eval) + DacVectorQuantize.decode_latents: latents [B, 1024, T] ->
    (codes int64 [B, 9, T], gap fp32 [B, 9, T] = best distance minus runner-up: how close each decision was)."""
    residual = z
    codes, gaps = [], []
    for k in range(N_CODEBOOKS):
        q = f"quantizer.quantizers.{k}"
        proj = F.conv1d(residual, sd[q + ".in_proj.weight"], sd[q + ".in_proj.bias"])          # [B, 8, T]
        b, dcb, t = proj.shape
        enc = F.normalize(proj.permute(0, 2, 1).reshape(b * t, dcb))
        cb = F.normalize(sd[q + ".codebook.weight"])
        dist = -(enc.pow(2).sum(1, keepdim=True) - 2 * enc @ cb.t()) + cb.pow(2).sum(1, keepdim=True).t()
        top2 = dist.topk(2, dim=1)
        idx = dist.max(1)[1].reshape(b, t)
        codes.append(idx)
        gaps.append((top2.values[:, 0] - top2.values[:, 1]).reshape(b, t))
        quant = F.embedding(idx, sd[q + ".codebook.weight"]).transpose(1, 2)
        residual = residual - F.conv1d(quant, sd[q + ".out_proj.weight"], sd[q + ".out_proj.bias"])
    return torch.stack(codes, 1), torch.stack(gaps, 1)


@torch.inference_mode()
def dac_encode(sd: dict, wav: torch.Tensor, with_gaps: bool = False):
    """zonos/autoencoder.py:22-23 on CPU: wav fp32 [B, 1, S] (S % 512 == 0) -> codes int64 [B, 9, S / 512]."""
    codes, gaps = rvq_encode(sd, encoder_forward(sd, wav.float()))
    return (codes, gaps) if with_gaps else codes


def synth_waveform(batch: int, frames: int, seed: int = 7) -> torch.Tensor:
    """Audio-like test input [B, 1, 512 * frames]: a few partials with slow amplitude modulation + a little noise, |x| < 0.6."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    n = HOP * frames
    t = torch.arange(n, dtype=torch.float32) / SAMPLING_RATE
    out = []
    for _ in range(batch):
        f0 = 90.0 + 200.0 * torch.rand(1, generator=g).item()
        x = sum((0.3 / (h + 1)) * torch.sin(2 * math.pi * f0 * (h + 1) * t + 6.28 * torch.rand(1, generator=g).item()) for h in range(5))
        x = x * (0.6 + 0.4 * torch.sin(2 * math.pi * 3.1 * t)) + 0.02 * torch.randn(n, generator=g)
        out.append(x)
    return torch.stack(out).unsqueeze(1)
