"""CPU restatement of the BigVGAN trunk + anti-aliased activation (TEST INFRASTRUCTURE ONLY).

Follows (all under /root/reference/cosyvoice/BigVGAN):
  bigvgan.py:406-436                              BigVGAN.forward trunk        -> trunk_forward
  bigvgan.py:128-137                              AMPBlock1.forward            -> _ampblock1
  alias_free_activation/torch/act.py:26-31        Activation1d.forward         -> anti_alias_act
  alias_free_activation/torch/resample.py:29-38   UpSample1d.forward           -> _upsample2x
  alias_free_activation/torch/filter.py:31-63     kaiser_sinc_filter1d         -> kaiser_sinc_filter
  alias_free_activation/torch/filter.py:95-102    LowPassFilter1d.forward      -> _downsample2x
  nnet/activations.py:109-122                     SnakeBeta.forward (logscale) -> snakebeta
  alias_free_activation/cuda/anti_alias_activation_cuda.cu:44-181  the fused native kernel (same math)
"""
import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

from .hift_ref import wn

SD = Dict[str, torch.Tensor]


def arch(topology: str) -> dict:
    if topology == "24k100":     # BASELINE config 4: NVIDIA 24 kHz / 100-band topology
        return dict(in_ch=100, rates=[4, 4, 2, 2, 2, 2], ks=[8, 8, 4, 4, 4, 4], ch0=1536, spk=512)
    if topology == "fork":       # examples/tts_vc/cosyvoice2/conf/cosyvoice_bigvgan_tts.yaml:43-61
        return dict(in_ch=1024, rates=[8, 5, 3, 2], ks=[16, 11, 7, 2], ch0=1536, spk=512)
    raise ValueError(topology)


def kaiser_sinc_filter(cutoff: float = 0.25, half_width: float = 0.3, kernel_size: int = 12) -> torch.Tensor:
    """filter.py:31-63 (even kernel): Kaiser-windowed sinc normalised to unit sum.  Returns [K]."""
    half = kernel_size // 2
    delta_f = 4 * half_width
    A = 2.285 * (half - 1) * math.pi * delta_f + 7.95
    if A > 50.0:
        beta = 0.1102 * (A - 8.7)
    elif A >= 21.0:
        beta = 0.5842 * (A - 21) ** 0.4 + 0.07886 * (A - 21.0)
    else:
        beta = 0.0
    window = torch.kaiser_window(kernel_size, beta=beta, periodic=False)
    time = torch.arange(-half, half) + 0.5
    f = 2 * cutoff * window * torch.sinc(2 * cutoff * time)
    return f / f.sum()


def snakebeta(x, alpha, beta):
    """nnet/activations.py:116-120 with alpha_logscale=True."""
    a = torch.exp(alpha).reshape(1, -1, 1)
    b = torch.exp(beta).reshape(1, -1, 1)
    return x + (1.0 / (b + 1e-9)) * torch.sin(x * a) ** 2


def _upsample2x(x, filt):
    C = x.shape[1]
    x = F.pad(x, (5, 5), mode="replicate")
    x = 2 * F.conv_transpose1d(x, filt.reshape(1, 1, -1).expand(C, -1, -1), stride=2, groups=C)
    return x[..., 15:-15]


def _downsample2x(x, filt):
    C = x.shape[1]
    x = F.pad(x, (5, 6), mode="replicate")
    return F.conv1d(x, filt.reshape(1, 1, -1).expand(C, -1, -1), stride=2, groups=C)


def anti_alias_act(x, alpha, beta, filt: Optional[torch.Tensor] = None):
    """Activation1d(SnakeBeta): 2x kaiser-sinc up, SnakeBeta, 2x low-pass down.  x [B,C,T]."""
    if filt is None:
        filt = kaiser_sinc_filter().to(x.dtype)
    return _downsample2x(snakebeta(_upsample2x(x, filt), alpha, beta), filt)


def _ampblock1(sd: SD, p: str, x, k: int, filt, dils=(1, 3, 5)):
    for i, d in enumerate(dils):
        xt = anti_alias_act(x, sd[f"{p}activations.{2 * i}.act.alpha"], sd[f"{p}activations.{2 * i}.act.beta"], filt)
        xt = F.conv1d(xt, wn(sd, f"{p}convs1.{i}."), sd[f"{p}convs1.{i}.bias"], dilation=d, padding=(k * d - d) // 2)
        xt = anti_alias_act(xt, sd[f"{p}activations.{2 * i + 1}.act.alpha"], sd[f"{p}activations.{2 * i + 1}.act.beta"], filt)
        xt = F.conv1d(xt, wn(sd, f"{p}convs2.{i}."), sd[f"{p}convs2.{i}.bias"], padding=(k - 1) // 2)
        x = xt + x
    return x


def _ampblock2(sd: SD, p: str, x, k: int, filt, dils=(1, 3, 5)):
    """AMPBlock2.forward, BigVGAN/bigvgan.py:223-228: one dilated conv per layer, x = conv(act(x)) + x."""
    for i, d in enumerate(dils):
        xt = anti_alias_act(x, sd[f"{p}activations.{i}.act.alpha"], sd[f"{p}activations.{i}.act.beta"], filt)
        xt = F.conv1d(xt, wn(sd, f"{p}convs.{i}."), sd[f"{p}convs.{i}.bias"], dilation=d, padding=(k * d - d) // 2)
        x = xt + x
    return x


def trunk_forward(sd: SD, feat, spk, topology: str = "24k100", taps: Optional[dict] = None):
    """bigvgan.py:406-436.  feat [B,C_in,T] (the encoder_proj output), spk [B,512,1] -> wav [B,L]."""
    a = arch(topology)
    filt = kaiser_sinc_filter().to(feat.dtype)
    x = F.conv1d(feat, wn(sd, "conv_pre."), sd["conv_pre.bias"], padding=3)
    x = x + F.conv1d(spk, sd["cond_layer.weight"], sd["cond_layer.bias"])
    for i, (u, k) in enumerate(zip(a["rates"], a["ks"])):
        x = F.conv_transpose1d(x, wn(sd, f"ups.{i}.0."), sd[f"ups.{i}.0.bias"], stride=u, padding=(k - u) // 2)
        x = x + F.conv1d(spk, sd[f"conds.{i}.weight"], sd[f"conds.{i}.bias"])
        xs = None
        for j, rk in enumerate((3, 7, 11)):
            p = f"resblocks.{i * 3 + j}."
            r = (_ampblock2 if f"{p}convs.0.bias" in sd else _ampblock1)(sd, p, x, rk, filt)
            xs = r if xs is None else xs + r
        x = xs / 3
        if taps is not None:
            taps[f"stage{i}"] = x
    x = anti_alias_act(x, sd["activation_post.act.alpha"], sd["activation_post.act.beta"], filt)
    x = F.conv1d(x, wn(sd, "conv_post."), sd["conv_post.bias"], padding=3)
    return torch.tanh(x).squeeze(1)
