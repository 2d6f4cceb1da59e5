"""CPU restatement of the HiFT vocoder (TEST INFRASTRUCTURE ONLY).

Functional torch code over a state_dict with the reference's key names (old-style
`weight_g` / `weight_v` weight-norm pairs, cosyvoice/hifigan/generator.py:25-26).

Follows (all under /root/reference/cosyvoice):
  hifigan/generator.py:399-411   HiFTGenerator.inference     -> hift_inference
  hifigan/generator.py:349-381   HiFTGenerator.decode        -> decode
  hifigan/generator.py:333-347   _stft / _istft              -> stft / istft
  hifigan/generator.py:91-98     ResBlock.forward            -> _resblock
  hifigan/generator.py:138-168   SineGen.forward             -> sine_source
  hifigan/generator.py:204-220   SourceModuleHnNSF.forward   -> sine_source (merge part)
  hifigan/f0_predictor.py:52-55  ConvRNNF0Predictor.forward  -> f0_predictor
  transformer/activation.py:73-84 Snake                      -> snake
The three RNG draws of the source module (generator.py:150,163,219) are INJECTED.
"""
import math
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]

N_FFT, HOP = 16, 4
SR = 24000
NB_HARM = 8
SINE_AMP, NOISE_STD, VOICED_THR = 0.1, 0.003, 10.0
LRELU = 0.1
AUDIO_LIMIT = 0.99


def arch(v2: bool) -> dict:
    """examples/tts_vc/cosyvoice/conf/cosyvoice.yaml:114-136 and cosyvoice2/conf/cosyvoice.yaml."""
    if v2:
        return dict(up_rates=[8, 5, 3], up_k=[16, 11, 7], src_k=[7, 7, 11], hop=480)
    return dict(up_rates=[8, 8], up_k=[16, 16], src_k=[7, 11], hop=256)


def wn(sd: SD, p: str) -> torch.Tensor:
    """Fold old-style torch.nn.utils.weight_norm (dim=0): w = g * v / ||v||, norm over all dims but 0."""
    if p + "weight" in sd:
        return sd[p + "weight"]
    v, g = sd[p + "weight_v"], sd[p + "weight_g"]
    n = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
    return v * (g / n)


def hann_window(dtype=torch.float32) -> torch.Tensor:
    """scipy.signal.get_window('hann', 16, fftbins=True) (generator.py:316) = periodic Hann."""
    n = torch.arange(N_FFT, dtype=torch.float64)
    return (0.5 - 0.5 * torch.cos(2 * math.pi * n / N_FFT)).to(dtype)


def snake(x, alpha):
    """transformer/activation.py:79-82, alpha_logscale=False."""
    a = alpha.reshape(1, -1, 1)
    return x + (1.0 / (a + 1e-9)) * torch.sin(x * a) ** 2


def f0_predictor(sd: SD, mel):
    """f0_predictor.py:52-55: 5x (weight-normed Conv1d k3 + ELU), Linear 512->1, abs."""
    x = mel
    for i in range(5):
        p = f"f0_predictor.condnet.{2 * i}."
        x = F.elu(F.conv1d(x, wn(sd, p), sd[p + "bias"], padding=1))
    y = F.linear(x.transpose(1, 2), sd["f0_predictor.classifier.weight"], sd["f0_predictor.classifier.bias"])
    return torch.abs(y.squeeze(-1))


def sine_source(sd: SD, f0_up, phase_vec, noise, fp64_phase: bool = False):
    """SineGen + SourceModuleHnNSF merge.  f0_up [B,1,L] (already nearest-upsampled),
    phase_vec [B,9,1] (row 0 zeroed, generator.py:150-151), noise [B,9,L] = randn_like(sine_waves)."""
    B, _, L = f0_up.shape
    mult = torch.arange(1, NB_HARM + 2, dtype=f0_up.dtype).reshape(1, -1, 1)
    F_mat = f0_up * mult / SR                                                 # :144-146
    if fp64_phase:
        theta = (2 * np.pi * (torch.cumsum(F_mat.double(), dim=-1) % 1)).to(f0_up.dtype)
    else:
        theta = 2 * np.pi * (torch.cumsum(F_mat, dim=-1) % 1)                 # :148
    sine = SINE_AMP * torch.sin(theta + phase_vec)                            # :154
    uv = (f0_up > VOICED_THR).float()                                         # :134
    noise_amp = uv * NOISE_STD + (1 - uv) * SINE_AMP / 3                      # :162
    sine = sine * uv + noise_amp * noise                                      # :163-167
    merged = torch.tanh(F.linear(sine.transpose(1, 2), sd["m_source.l_linear.weight"], sd["m_source.l_linear.bias"]))
    return merged.transpose(1, 2)                                             # [B,1,L]


def stft(s):
    """generator.py:333-339: torch.stft(n_fft 16, hop 4, hann, center=True/reflect) -> real || imag [B,18,F]."""
    spec = torch.stft(s, N_FFT, HOP, N_FFT, window=hann_window(s.dtype), return_complex=True)
    return torch.cat([spec.real, spec.imag], dim=1)


def istft(mag, phase):
    """generator.py:341-347."""
    mag = torch.clip(mag, max=1e2)
    spec = torch.complex(mag * torch.cos(phase), mag * torch.sin(phase))
    return torch.istft(spec, N_FFT, HOP, N_FFT, window=hann_window(mag.dtype))


def _resblock(sd: SD, p: str, x, k: int, dils=(1, 3, 5)):
    for i, d in enumerate(dils):
        xt = snake(x, sd[f"{p}activations1.{i}.alpha"])
        xt = F.conv1d(xt, wn(sd, f"{p}convs1.{i}."), sd[f"{p}convs1.{i}.bias"], dilation=d, padding=(k * d - d) // 2)
        xt = snake(xt, sd[f"{p}activations2.{i}.alpha"])
        xt = F.conv1d(xt, wn(sd, f"{p}convs2.{i}."), sd[f"{p}convs2.{i}.bias"], padding=(k - 1) // 2)
        x = xt + x
    return x


def decode(sd: SD, mel, s, v2: bool = False, taps: Optional[dict] = None):
    """generator.py:349-381.  mel [B,80,T], s [B,1,L] -> wav [B,L]."""
    a = arch(v2)
    rates, ks = a["up_rates"], a["up_k"]
    n_up = len(rates)
    s_stft = stft(s.squeeze(1))
    if taps is not None:
        taps["s_stft"] = s_stft
    x = F.conv1d(mel, wn(sd, "conv_pre."), sd["conv_pre.bias"], padding=3)
    # source_downs strides: downsample_cum_rates[::-1] (generator.py:290-300)
    down = [1] + rates[::-1][:-1]
    cum = list(np.cumprod(down))[::-1]
    for i in range(n_up):
        x = F.leaky_relu(x, LRELU)
        x = F.conv_transpose1d(x, wn(sd, f"ups.{i}."), sd[f"ups.{i}.bias"], stride=rates[i], padding=(ks[i] - rates[i]) // 2)
        if i == n_up - 1:
            x = F.pad(x, (1, 0), mode="reflect")
        u = int(cum[i])
        w, b = sd[f"source_downs.{i}.weight"], sd[f"source_downs.{i}.bias"]
        si = F.conv1d(s_stft, w, b) if u == 1 else F.conv1d(s_stft, w, b, stride=u, padding=u // 2)
        si = _resblock(sd, f"source_resblocks.{i}.", si, a["src_k"][i])
        x = x + si
        if taps is not None:
            taps[f"fuse{i}"] = x
        xs = None
        for j, k in enumerate((3, 7, 11)):
            r = _resblock(sd, f"resblocks.{i * 3 + j}.", x, k)
            xs = r if xs is None else xs + r
        x = xs / 3
        if taps is not None:
            taps[f"stage{i}"] = x
    x = F.leaky_relu(x)                       # default slope 0.01, generator.py:374
    x = F.conv1d(x, wn(sd, "conv_post."), sd["conv_post.bias"], padding=3)
    if taps is not None:
        taps["conv_post"] = x
    mag = torch.exp(x[:, : N_FFT // 2 + 1])
    ph = torch.sin(x[:, N_FFT // 2 + 1:])
    wav = istft(mag, ph)
    return torch.clamp(wav, -AUDIO_LIMIT, AUDIO_LIMIT)


def hift_inference(sd: SD, mel, phase_vec, noise, cache_source=None, v2: bool = False,
                   fp64_phase: bool = False, taps: Optional[dict] = None):
    """generator.py:399-411 with injected RNG.  Returns (wav [B,L], s [B,1,L])."""
    hop = arch(v2)["hop"]
    f0 = f0_predictor(sd, mel)
    f0_up = f0[:, None].repeat_interleave(hop, dim=-1)          # nn.Upsample(scale_factor=hop), nearest
    s = sine_source(sd, f0_up, phase_vec, noise, fp64_phase=fp64_phase)
    if cache_source is not None and cache_source.shape[2] != 0:
        s = s.clone()
        s[:, :, : cache_source.shape[2]] = cache_source
    if taps is not None:
        taps["f0"] = f0
        taps["s"] = s
    return decode(sd, mel, s, v2=v2, taps=taps), s


def draw_source_rng(B: int, L: int, generator: Optional[torch.Generator] = None, device="cpu"):
    """Replays the reference's RNG call order (generator.py:150,163,219) so that a given
    torch seed yields the tensors the reference would have consumed."""
    u = torch.rand(B, NB_HARM + 1, 1, generator=generator, device=device)
    phase_vec = (-np.pi) + u * (2 * np.pi)      # Uniform(-pi, pi).sample == low + rand*(high-low)
    phase_vec[:, 0, :] = 0
    noise = torch.randn(B, NB_HARM + 1, L, generator=generator, device=device)
    _unused = torch.randn(B, L, 1, generator=generator, device=device)      # :219, consumed but discarded
    return phase_vec, noise
