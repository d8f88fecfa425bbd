"""CPU oracle for the speaker-embedding network: zonos/speaker_cloning.py `ResNet293_based`
(:199-224 = ResNet293-SimAM front :64-97,153-196 + attentive statistics pooling :38-61 + bottleneck) and the LDA projection
(:388-412), restated functionally in fp32 torch over an explicit state_dict with the reference's keys.

TEST INFRASTRUCTURE - NOT PRODUCT CODE: only tests/ may import this.  Pinned against the reference's own classes (imported from
/root/reference in the build container) by oracle/make_golden_spk.py -> tests/golden/golden_spk.pt.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

NUM_BLOCKS = (10, 20, 64, 3)      # ResNet293 (speaker_cloning.py:195-196)
IN_PLANES = 64
N_MELS = 80
EMBD_DIM = 256
LDA_DIM = 128


def synth_speaker_state_dict(seed: int = 5, num_blocks=NUM_BLOCKS) -> dict:
    """Deterministic random parameters under the keys of `ResNet293_based.state_dict()`.  BatchNorm gains are small on the
    second conv of a block so that 190 residual blocks keep the activations O(1)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    sd: dict = {}

    def conv(key, cout, cin, k):
        sd[key + ".weight"] = torch.randn(cout, cin, k, k, generator=g) * math.sqrt(2.0 / (cin * k * k))

    def bn(key, c, gain):
        sd[key + ".weight"] = gain * (0.8 + 0.4 * torch.rand(c, generator=g))
        sd[key + ".bias"] = 0.05 * torch.randn(c, generator=g)
        sd[key + ".running_mean"] = 0.05 * torch.randn(c, generator=g)
        sd[key + ".running_var"] = 0.8 + 0.4 * torch.rand(c, generator=g)
        sd[key + ".num_batches_tracked"] = torch.tensor(100)

    conv("front.conv1", IN_PLANES, 1, 3)
    bn("front.bn1", IN_PLANES, 1.0)
    cin = IN_PLANES
    for li, (nb, mult, stride) in enumerate(zip(num_blocks, (1, 2, 4, 8), (1, 2, 2, 2)), start=1):
        planes = IN_PLANES * mult
        for bi in range(nb):
            p = f"front.layer{li}.{bi}"
            s = stride if bi == 0 else 1
            conv(p + ".conv1", planes, cin, 3)
            bn(p + ".bn1", planes, 1.0)
            conv(p + ".conv2", planes, planes, 3)
            bn(p + ".bn2", planes, 0.3)
            if s != 1 or cin != planes:
                conv(p + ".downsample.0", planes, cin, 1)
                bn(p + ".downsample.1", planes, 1.0)
            cin = planes
    c = IN_PLANES * 8 * (N_MELS // 8)
    sd["pooling.attention.0.weight"] = torch.randn(128, c, 1, generator=g) / math.sqrt(c)
    sd["pooling.attention.0.bias"] = 0.05 * torch.randn(128, generator=g)
    bn("pooling.attention.2", 128, 1.0)
    sd["pooling.attention.3.weight"] = torch.randn(c, 128, 1, generator=g) * (2.0 / math.sqrt(128))
    sd["pooling.attention.3.bias"] = 0.05 * torch.randn(c, generator=g)
    sd["bottleneck.weight"] = torch.randn(EMBD_DIM, 2 * c, generator=g) / math.sqrt(2 * c)
    sd["bottleneck.bias"] = 0.05 * torch.randn(EMBD_DIM, generator=g)
    return sd


def synth_lda_state_dict(seed: int = 6) -> dict:
    g = torch.Generator(device="cpu").manual_seed(seed)
    return {"weight": torch.randn(LDA_DIM, EMBD_DIM, generator=g) / math.sqrt(EMBD_DIM), "bias": 0.05 * torch.randn(LDA_DIM, generator=g)}


def _bn(sd, key, x):
    return F.batch_norm(x, sd[key + ".running_mean"], sd[key + ".running_var"], sd[key + ".weight"], sd[key + ".bias"], False, 0.0, 1e-5)


def simam(x, lambda_p=1e-4):
    """speaker_cloning.py:91-97."""
    n = x.shape[2] * x.shape[3] - 1
    d = (x - x.mean(dim=[2, 3], keepdim=True)).pow(2)
    v = d.sum(dim=[2, 3], keepdim=True) / n
    return x * torch.sigmoid(d / (4 * (v + lambda_p)) + 0.5)


def simam_block(sd, p, x, stride):
    """SimAMBasicBlock.forward, speaker_cloning.py:83-89."""
    out = F.relu(_bn(sd, p + ".bn1", F.conv2d(x, sd[p + ".conv1.weight"], stride=stride, padding=1)))
    out = _bn(sd, p + ".bn2", F.conv2d(out, sd[p + ".conv2.weight"], padding=1))
    out = simam(out)
    if p + ".downsample.0.weight" in sd:
        x = _bn(sd, p + ".downsample.1", F.conv2d(x, sd[p + ".downsample.0.weight"], stride=stride))
    return F.relu(out + x)


def front(sd, feats, num_blocks=NUM_BLOCKS):
    """ResNet.forward, speaker_cloning.py:186-192: feats [B, 80, frames] -> [B, 512, 10, frames / 8]."""
    x = F.relu(_bn(sd, "front.bn1", F.conv2d(feats.unsqueeze(1), sd["front.conv1.weight"], padding=1)))
    for li, (nb, stride) in enumerate(zip(num_blocks, (1, 2, 2, 2)), start=1):
        for bi in range(nb):
            x = simam_block(sd, f"front.layer{li}.{bi}", x, stride if bi == 0 else 1)
    return x


def asp(sd, x):
    """ASP.forward, speaker_cloning.py:53-61."""
    x = x.reshape(x.size(0), -1, x.size(-1))
    h = F.relu(F.conv1d(x, sd["pooling.attention.0.weight"], sd["pooling.attention.0.bias"]))
    h = F.batch_norm(h, sd["pooling.attention.2.running_mean"], sd["pooling.attention.2.running_var"], sd["pooling.attention.2.weight"],
                     sd["pooling.attention.2.bias"], False, 0.0, 1e-5)
    w = torch.softmax(F.conv1d(h, sd["pooling.attention.3.weight"], sd["pooling.attention.3.bias"]), dim=2)
    mu = torch.sum(x * w, dim=2)
    sg = torch.sqrt((torch.sum((x ** 2) * w, dim=2) - mu ** 2).clamp(min=1e-5))
    return torch.cat((mu, sg), 1)


@torch.inference_mode()
def embedding_from_features(sd, feats, num_blocks=NUM_BLOCKS, lda=None):
    """ResNet293_based.forward after featCal (speaker_cloning.py:218-224) [+ SpeakerEmbeddingLDA.forward :410-412]."""
    emb = F.linear(asp(sd, front(sd, feats.float(), num_blocks)), sd["bottleneck.weight"], sd["bottleneck.bias"])
    if lda is None:
        return emb
    return emb, F.linear(emb, lda["weight"], lda["bias"])


def synth_features(batch: int, frames: int, seed: int = 8) -> torch.Tensor:
    """Log-mel-like input [B, 80, frames]: smooth in time and frequency, mean-normalised over time like featCal's output."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(batch, N_MELS, frames + 8, generator=g)
    x = F.avg_pool1d(x, 5, stride=1)[..., :frames] * 3.0 + torch.linspace(2.0, -2.0, N_MELS).view(1, -1, 1)
    return x - x.mean(dim=2, keepdim=True)
