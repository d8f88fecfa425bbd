"""Speaker-embedding network on the device — mirror of the reference's zonos/speaker_cloning.py:
`logFbankCal`, `ASP`, `SimAMBasicBlock`, `ResNet`, `ResNet293`, `ResNet293_based`, `SpeakerEmbedding`, `SpeakerEmbeddingLDA`.

Same class names, constructor arguments and `state_dict` keys, so `ResNet293_SimAM_ASP_base.pt` and the LDA file load by
key.  The modules only HOLD the parameters; the arithmetic (190 SimAM blocks = 381 convolutions with folded eval-mode
BatchNorm, SimAM gating, attentive statistics pooling, bottleneck, LDA) runs in libzonos_b200 (`zb_spk_*`, csrc/spk_api.cu) in
fp32 like the reference.  The log-mel front end is the same torchaudio `MelSpectrogram` call as the reference's
(speaker_cloning.py:13-35), on the device: an FFT library call, not part of the network.  The ECAPA-TDNN variant of the
reference file (:227-353) is not used by Zonos (`SpeakerEmbedding` builds `ResNet293_based`, :361) and is not mirrored.
"""
from __future__ import annotations

from functools import cache

import torch
import torch.nn as nn

from . import _lib


def _s(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _f32(t: torch.Tensor) -> torch.Tensor:
    if not t.is_cuda:
        raise _lib.ZbError("the zonos_vibes_b200 speaker-embedding network runs on a CUDA (sm_100) device; there is no CPU fallback")
    return t.detach().to(torch.float32).contiguous()


def _fold_bn(bn: nn.modules.batchnorm._BatchNorm) -> tuple[torch.Tensor, torch.Tensor]:
    """eval-mode BatchNorm as y = x * scale + shift."""
    scale = bn.weight.detach().float() / torch.sqrt(bn.running_var.detach().float() + bn.eps)
    shift = bn.bias.detach().float() - bn.running_mean.detach().float() * scale
    return scale.contiguous(), shift.contiguous()


def _conv_bn(x: torch.Tensor, conv: nn.Conv2d, bn: nn.BatchNorm2d, relu: bool, cache_: dict) -> torch.Tensor:
    key = id(conv)
    if key not in cache_:
        cache_[key] = (_f32(conv.weight), *_fold_bn(bn))
    w, scale, shift = cache_[key]
    b, cin, h, wd = x.shape
    k, stride = conv.kernel_size[0], conv.stride[0]
    pad = 1 if k == 3 else 0
    ho, wo = (h + 2 * pad - k) // stride + 1, (wd + 2 * pad - k) // stride + 1
    out = torch.empty((b, conv.out_channels, ho, wo), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().zb_spk_conv2d(x.data_ptr(), w.data_ptr(), scale.data_ptr(), shift.data_ptr(), out.data_ptr(), b, cin, h, wd,
                                             conv.out_channels, k, stride, 1 if relu else 0, _s(x)))
    return out


class logFbankCal(nn.Module):
    """speaker_cloning.py:13-35 (torchaudio MelSpectrogram -> log -> mean normalisation over time)."""

    def __init__(self, sample_rate: int = 16_000, n_fft: int = 512, win_length: float = 0.025, hop_length: float = 0.01, n_mels: int = 80):
        super().__init__()
        import torchaudio
        self.fbankCal = torchaudio.transforms.MelSpectrogram(sample_rate=sample_rate, n_fft=n_fft, win_length=int(win_length * sample_rate),
                                                             hop_length=int(hop_length * sample_rate), n_mels=n_mels)

    def forward(self, x):
        out = self.fbankCal(x)
        out = torch.log(out + 1e-6)
        return out - out.mean(axis=2).unsqueeze(dim=2)


class ASP(nn.Module):
    """Attentive statistics pooling, speaker_cloning.py:38-61."""

    def __init__(self, in_planes, acoustic_dim):
        super().__init__()
        outmap_size = int(acoustic_dim / 8)
        self.out_dim = in_planes * 8 * outmap_size * 2
        c = in_planes * 8 * outmap_size
        self.attention = nn.Sequential(nn.Conv1d(c, 128, kernel_size=1), nn.ReLU(), nn.BatchNorm1d(128), nn.Conv1d(128, c, kernel_size=1),
                                       nn.Softmax(dim=2))

    def forward(self, x):
        x = _f32(x.reshape(x.size()[0], -1, x.size()[-1]))
        b, c, t = x.shape
        conv1, bn, conv2 = self.attention[0], self.attention[2], self.attention[3]
        scale, shift = _fold_bn(bn)
        hidden = torch.empty((b, conv1.out_channels, t), dtype=torch.float32, device=x.device)
        out = torch.empty((b, 2 * c), dtype=torch.float32, device=x.device)
        w1, b1, w2, b2 = _f32(conv1.weight), _f32(conv1.bias), _f32(conv2.weight), _f32(conv2.bias)
        with torch.cuda.device(x.device):
            _lib.check(_lib.load().zb_spk_asp(x.data_ptr(), w1.data_ptr(), b1.data_ptr(), scale.data_ptr(), shift.data_ptr(), w2.data_ptr(),
                                              b2.data_ptr(), hidden.data_ptr(), out.data_ptr(), b, c, t, conv1.out_channels, _s(x)))
        return out


class SimAMBasicBlock(nn.Module):
    """speaker_cloning.py:64-97."""
    expansion = 1

    def __init__(self, ConvLayer, NormLayer, in_planes, planes, stride=1, block_id=1):
        super().__init__()
        self.conv1 = ConvLayer(in_planes, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn1 = NormLayer(planes)
        self.conv2 = ConvLayer(planes, planes, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn2 = NormLayer(planes)
        self.relu = nn.ReLU(inplace=True)
        self.sigmoid = nn.Sigmoid()
        self.downsample = nn.Sequential()
        if stride != 1 or in_planes != self.expansion * planes:
            self.downsample = nn.Sequential(ConvLayer(in_planes, self.expansion * planes, kernel_size=1, stride=stride, bias=False),
                                            NormLayer(self.expansion * planes))

    def run(self, x: torch.Tensor, cache_: dict, lambda_p: float = 1e-4) -> torch.Tensor:
        out = _conv_bn(x, self.conv1, self.bn1, True, cache_)
        out = _conv_bn(out, self.conv2, self.bn2, False, cache_)
        shortcut = _conv_bn(x, self.downsample[0], self.downsample[1], False, cache_) if len(self.downsample) else x
        b, c, h, w = out.shape
        res = torch.empty_like(out)
        with torch.cuda.device(x.device):
            _lib.check(_lib.load().zb_spk_simam_add_relu(out.data_ptr(), shortcut.data_ptr(), res.data_ptr(), b, c, h, w, float(lambda_p), _s(x)))
        return res


class ResNet(nn.Module):
    """speaker_cloning.py:153-192 (feat_dim "2d" only: the variant Zonos uses)."""

    def __init__(self, in_planes, block, num_blocks, in_ch=1, feat_dim="2d", **kwargs):
        super().__init__()
        if feat_dim != "2d":
            raise _lib.ZbError("zonos_vibes_b200 mirrors the 2d ResNet front (the one ResNet293_based builds)")
        self.NormLayer, self.ConvLayer = nn.BatchNorm2d, nn.Conv2d
        self.in_planes = in_planes
        self.conv1 = self.ConvLayer(in_ch, in_planes, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = self.NormLayer(in_planes)
        self.relu = nn.ReLU(inplace=True)
        self.layer1 = self._make_layer(block, in_planes, num_blocks[0], stride=1, block_id=1)
        self.layer2 = self._make_layer(block, in_planes * 2, num_blocks[1], stride=2, block_id=2)
        self.layer3 = self._make_layer(block, in_planes * 4, num_blocks[2], stride=2, block_id=3)
        self.layer4 = self._make_layer(block, in_planes * 8, num_blocks[3], stride=2, block_id=4)
        self._cache: dict = {}

    def _make_layer(self, block, planes, num_blocks, stride, block_id=1):
        layers = []
        for s in [stride] + [1] * (num_blocks - 1):
            layers.append(block(self.ConvLayer, self.NormLayer, self.in_planes, planes, s, block_id))
            self.in_planes = planes * block.expansion
        return nn.Sequential(*layers)

    def invalidate(self) -> None:
        """Forget the folded BatchNorm / fp32 weight copies (call after changing parameters; load_state_dict does)."""
        self._cache.clear()

    def load_state_dict(self, *a, **k):
        self.invalidate()
        return super().load_state_dict(*a, **k)

    def forward(self, x):
        x = _conv_bn(_f32(x), self.conv1, self.bn1, True, self._cache)
        for layer in (self.layer1, self.layer2, self.layer3, self.layer4):
            for blk in layer:
                x = blk.run(x, self._cache)
        return x


def ResNet293(in_planes: int, **kwargs):
    return ResNet(in_planes, SimAMBasicBlock, [10, 20, 64, 3], **kwargs)


def _linear(x: torch.Tensor, lin: nn.Linear) -> torch.Tensor:
    x = _f32(x)
    w = _f32(lin.weight)
    b = _f32(lin.bias) if lin.bias is not None else None
    out = torch.empty((x.shape[0], lin.out_features), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().zb_spk_linear(x.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None, out.data_ptr(), x.shape[0],
                                             lin.in_features, lin.out_features, _s(x)))
    return out


class ResNet293_based(nn.Module):
    """speaker_cloning.py:199-224.  `num_blocks` (not in the reference) lets tests build a shallower front with the same blocks."""

    def __init__(self, in_planes: int = 64, embd_dim: int = 256, acoustic_dim: int = 80, featCal=None, dropout: float = 0, num_blocks=None, **kwargs):
        super().__init__()
        self.featCal = featCal
        self.front = ResNet293(in_planes) if num_blocks is None else ResNet(in_planes, SimAMBasicBlock, list(num_blocks))
        self.pooling = ASP(in_planes * SimAMBasicBlock.expansion, acoustic_dim)
        self.bottleneck = nn.Linear(self.pooling.out_dim, embd_dim)
        self.drop = nn.Dropout(dropout) if dropout else None

    def load_state_dict(self, *a, **k):
        self.front.invalidate()
        return super().load_state_dict(*a, **k)

    def forward_features(self, feats: torch.Tensor) -> torch.Tensor:
        """From log-mel features [B, 80, frames] (the output of featCal) to the embedding [B, embd_dim]."""
        x = self.front(feats.unsqueeze(dim=1))
        x = self.pooling(x)
        return _linear(x, self.bottleneck)          # dropout is the identity in eval mode (and 0 in the reference's construction)

    @torch.inference_mode()
    def forward(self, x):
        return self.forward_features(self.featCal(x))


class SpeakerEmbedding(nn.Module):
    """speaker_cloning.py:356-385."""

    def __init__(self, ckpt_path: str = "ResNet293_SimAM_ASP_base.pt", device: str = "cuda"):
        super().__init__()
        self.device = device
        with torch.device(device):
            self.model = ResNet293_based()
            state_dict = torch.load(ckpt_path, weights_only=True, mmap=True, map_location="cpu")
            self.model.load_state_dict(state_dict)
            self.model.featCal = logFbankCal()
        self.requires_grad_(False).eval()

    @property
    def dtype(self):
        return next(self.parameters()).dtype

    @cache
    def _get_resampler(self, orig_sample_rate: int):
        import torchaudio
        return torchaudio.transforms.Resample(orig_sample_rate, 16_000).to(self.device)

    def prepare_input(self, wav: torch.Tensor, sample_rate: int) -> torch.Tensor:
        assert wav.ndim < 3
        if wav.ndim == 2:
            wav = wav.mean(0, keepdim=True)
        return self._get_resampler(sample_rate)(wav)

    def forward(self, wav: torch.Tensor, sample_rate: int):
        wav = self.prepare_input(wav.to(self.device), sample_rate).to(self.device, self.dtype)
        return self.model(wav).to(wav.device)


class SpeakerEmbeddingLDA(nn.Module):
    """speaker_cloning.py:388-412: `(embedding [B, 256], lda(embedding) [B, 128])`.  Checkpoints come from the same hub repo as the
    reference's, or from local paths (`spk_model_path`, `lda_spk_model_path`) when there is no network."""

    def __init__(self, device: str = "cuda", spk_model_path: str | None = None, lda_spk_model_path: str | None = None):
        super().__init__()
        if spk_model_path is None or lda_spk_model_path is None:
            from huggingface_hub import hf_hub_download
            spk_model_path = spk_model_path or hf_hub_download(repo_id="Zyphra/Zonos-v0.1-speaker-embedding", filename="ResNet293_SimAM_ASP_base.pt")
            lda_spk_model_path = lda_spk_model_path or hf_hub_download(repo_id="Zyphra/Zonos-v0.1-speaker-embedding",
                                                                       filename="ResNet293_SimAM_ASP_base_LDA-128.pt")
        self.device = device
        with torch.device(device):
            self.model = SpeakerEmbedding(spk_model_path, device)
            lda_sd = torch.load(lda_spk_model_path, weights_only=True)
            out_features, in_features = lda_sd["weight"].shape
            self.lda = nn.Linear(in_features, out_features, bias=True, dtype=torch.float32)
            self.lda.load_state_dict(lda_sd)
        self.requires_grad_(False).eval()

    def forward(self, wav: torch.Tensor, sample_rate: int):
        emb = self.model(wav, sample_rate).to(torch.float32)
        return emb, _linear(emb, self.lda)
