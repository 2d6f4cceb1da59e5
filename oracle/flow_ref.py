"""CPU restatement of the flow-matching decoder (TEST INFRASTRUCTURE ONLY).

Functional torch code over a plain state_dict that uses the reference's own
parameter names (`decoder.estimator.*` stripped to the estimator's keys), so the
same dict drives the reference module, this oracle and the CUDA weight packer.

Follows (reference file:line, all under /root/reference/cosyvoice):
  flow/decoder.py:222-334                 ConditionalDecoder.forward      -> estimator_forward
  flow/components/decoder.py:12-27        SinusoidalPosEmb                -> sinusoidal_pos_emb
  flow/components/decoder.py:71-115       TimestepEmbedding               -> time_mlp
  flow/components/decoder.py:30-59        Block1D / ResnetBlock1D         -> _block1d / _resnet
  flow/decoder.py:36-85                   CausalBlock1D / CausalConv1d    -> causal=True branches
  flow/components/transformer.py:243-316  BasicTransformerBlock.forward   -> _transformer_block
  diffusers==0.27.2 (un-vendored)         Attention / AttnProcessor2_0 / GELU -> _attention, F.gelu
  flow/flow_matching.py:37-124,209-240    ConditionalCFM.forward / solve_euler -> cfm_forward / solve_euler
"""
import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]

CFG_RATE = 0.7          # inference_cfg_rate, examples/tts_vc/cosyvoice/conf/cosyvoice.yaml
N_HEADS = 8
HEAD_DIM = 64
FLOW_CACHE_TAIL = 34    # hard-coded at flow/flow_matching.py:63-64


def arch(causal: bool) -> dict:
    """Stage structure derived from flow/decoder.py:126-205 for the two shipped yamls."""
    if causal:   # CosyVoice2: channels=[256]
        return dict(causal=True, n_down=1, n_mid=12, n_up=1, n_blocks=4)
    return dict(causal=False, n_down=2, n_mid=12, n_up=2, n_blocks=4)


# ----------------------------------------------------------------------------- time path
def sinusoidal_pos_emb(t: torch.Tensor, dim: int = 320, scale: float = 1000.0) -> torch.Tensor:
    """components/decoder.py:18-27."""
    half = dim // 2
    k = math.log(10000) / (half - 1)
    freq = torch.exp(torch.arange(half, dtype=torch.float32) * -k)
    e = scale * t.reshape(-1, 1).float() * freq.reshape(1, -1)
    return torch.cat((e.sin(), e.cos()), dim=-1)


def time_mlp(sd: SD, t: torch.Tensor) -> torch.Tensor:
    """flow/decoder.py:240-241 + components/decoder.py:103-115 (SiLU between two Linears)."""
    e = sinusoidal_pos_emb(t).to(t.dtype)
    e = F.linear(e, sd["time_mlp.linear_1.weight"], sd["time_mlp.linear_1.bias"])
    e = F.silu(e)
    return F.linear(e, sd["time_mlp.linear_2.weight"], sd["time_mlp.linear_2.bias"])


# ----------------------------------------------------------------------------- conv blocks
def _conv_k3(x, w, b, causal):
    if causal:      # flow/decoder.py:80-84: left pad k-1, no right context
        return F.conv1d(F.pad(x, (2, 0)), w, b)
    return F.conv1d(x, w, b, padding=1)


def _block1d(sd: SD, p: str, x, mask, causal):
    """components/decoder.py:39-41 (GroupNorm(8)) / flow/decoder.py:39-49 (per-frame LayerNorm)."""
    h = _conv_k3(x * mask, sd[p + "block.0.weight"], sd[p + "block.0.bias"], causal)
    if causal:      # Sequential(conv, T, LayerNorm, T, Mish): norm params live at index 2
        h = F.layer_norm(h.transpose(1, 2), (h.shape[1],), sd[p + "block.2.weight"], sd[p + "block.2.bias"]).transpose(1, 2)
    else:
        h = F.group_norm(h, 8, sd[p + "block.1.weight"], sd[p + "block.1.bias"])
    return F.mish(h) * mask


def _resnet(sd: SD, p: str, x, mask, temb, causal):
    """components/decoder.py:54-59."""
    h = _block1d(sd, p + "block1.", x, mask, causal)
    h = h + F.linear(F.mish(temb), sd[p + "mlp.1.weight"], sd[p + "mlp.1.bias"]).unsqueeze(-1)
    h = _block1d(sd, p + "block2.", h, mask, causal)
    return h + F.conv1d(x * mask, sd[p + "res_conv.weight"], sd[p + "res_conv.bias"])


# ----------------------------------------------------------------------------- transformer
def _attention(sd: SD, p: str, x, attn_mask):
    """diffusers 0.27.2 Attention + AttnProcessor2_0 as the reference reaches it (components/transformer.py:196-204,
    266-271): prepare_attention_mask repeats the float mask per head, then ONE F.scaled_dot_product_attention call
    (softmax(QK^T * d^-1/2 + M) V; the float mask M is ADDED, flow/decoder.py:258).  Calling SDPA -- not an explicit
    matmul/softmax/matmul -- matters for the CPU baseline: it is the reference's own cost (round-1 review: the explicit
    form was 1.55x slower on the flow)."""
    B, S, _ = x.shape
    q = F.linear(x, sd[p + "to_q.weight"]).view(B, S, N_HEADS, HEAD_DIM).transpose(1, 2)
    k = F.linear(x, sd[p + "to_k.weight"]).view(B, S, N_HEADS, HEAD_DIM).transpose(1, 2)
    v = F.linear(x, sd[p + "to_v.weight"]).view(B, S, N_HEADS, HEAD_DIM).transpose(1, 2)
    if attn_mask is not None:
        attn_mask = attn_mask.repeat_interleave(N_HEADS, dim=0).view(B, N_HEADS, -1, attn_mask.shape[-1])
    o = F.scaled_dot_product_attention(q, k, v, attn_mask=attn_mask, dropout_p=0.0, is_causal=False)
    o = o.transpose(1, 2).reshape(B, S, N_HEADS * HEAD_DIM)
    return F.linear(o, sd[p + "to_out.0.weight"], sd[p + "to_out.0.bias"])


def _attention_explicit(sd: SD, p: str, x, attn_mask):
    """The same attention written out (independent cross-check of the SDPA call, SURVEY.md 8c last row)."""
    B, S, _ = x.shape
    q = F.linear(x, sd[p + "to_q.weight"]).view(B, S, N_HEADS, HEAD_DIM).transpose(1, 2)
    k = F.linear(x, sd[p + "to_k.weight"]).view(B, S, N_HEADS, HEAD_DIM).transpose(1, 2)
    v = F.linear(x, sd[p + "to_v.weight"]).view(B, S, N_HEADS, HEAD_DIM).transpose(1, 2)
    s = torch.matmul(q, k.transpose(-1, -2)) * (HEAD_DIM ** -0.5)
    if attn_mask is not None:
        s = s + attn_mask.unsqueeze(1)
    o = torch.matmul(torch.softmax(s, dim=-1), v)
    o = o.transpose(1, 2).reshape(B, S, N_HEADS * HEAD_DIM)
    return F.linear(o, sd[p + "to_out.0.weight"], sd[p + "to_out.0.bias"])


def _transformer_block(sd: SD, p: str, x, attn_mask):
    """components/transformer.py:262-316 with norm_type='layer_norm', no cross-attention."""
    C = x.shape[-1]
    n = F.layer_norm(x, (C,), sd[p + "norm1.weight"], sd[p + "norm1.bias"])
    x = _attention(sd, p + "attn1.", n, attn_mask) + x
    n = F.layer_norm(x, (C,), sd[p + "norm3.weight"], sd[p + "norm3.bias"])
    h = F.gelu(F.linear(n, sd[p + "ff.net.0.proj.weight"], sd[p + "ff.net.0.proj.bias"]))   # exact erf GELU
    h = F.linear(h, sd[p + "ff.net.2.weight"], sd[p + "ff.net.2.bias"])
    return h + x


def _stage_transformers(sd, p, x, mask, n_blocks):
    """flow/decoder.py:257-276: NCT -> NTC, additive {0,1} mask = mask^T mask, 4 blocks, back."""
    x = x.transpose(1, 2).contiguous()
    attn_mask = torch.matmul(mask.transpose(1, 2), mask)
    for i in range(n_blocks):
        x = _transformer_block(sd, f"{p}{i}.", x, attn_mask)
    return x.transpose(1, 2).contiguous()


# ----------------------------------------------------------------------------- estimator
def estimator_forward(sd: SD, x, mask, mu, t, spks, cond, causal: bool = False,
                      taps: Optional[dict] = None) -> torch.Tensor:
    """flow/decoder.py:222-334.  x,mu,cond [N,80,T]; mask [N,1,T]; t [N]; spks [N,80] -> [N,80,T].
    `taps`, if given, receives named intermediate tensors (used by per-layer CUDA tests)."""
    a = arch(causal)
    temb = time_mlp(sd, t)
    T = x.shape[-1]
    h = torch.cat([x, mu, spks.unsqueeze(-1).expand(-1, -1, T), cond], dim=1)      # 320 ch, :243-249
    hiddens: List[torch.Tensor] = []
    masks = [mask]
    for i in range(a["n_down"]):
        m = masks[-1]
        h = _resnet(sd, f"down_blocks.{i}.0.", h, m, temb, causal)
        if taps is not None:
            taps[f"down{i}.resnet"] = h
        h = _stage_transformers(sd, f"down_blocks.{i}.1.", h, m, a["n_blocks"])
        if taps is not None:
            taps[f"down{i}.tb"] = h
        hiddens.append(h)
        w, b = sd[f"down_blocks.{i}.2.conv.weight"] if i < a["n_down"] - 1 else sd[f"down_blocks.{i}.2.weight"], \
            sd[f"down_blocks.{i}.2.conv.bias"] if i < a["n_down"] - 1 else sd[f"down_blocks.{i}.2.bias"]
        if i < a["n_down"] - 1:            # Downsample1D: Conv1d(k3, stride 2, pad 1), components/decoder.py:65
            h = F.conv1d(h * m, w, b, stride=2, padding=1)
        else:
            h = _conv_k3(h * m, w, b, causal)
        masks.append(m[:, :, ::2])
    masks = masks[:-1]
    m_mid = masks[-1]
    for i in range(a["n_mid"]):
        h = _resnet(sd, f"mid_blocks.{i}.0.", h, m_mid, temb, causal)
        h = _stage_transformers(sd, f"mid_blocks.{i}.1.", h, m_mid, a["n_blocks"])
    if taps is not None:
        taps["mid"] = h
    for i in range(a["n_up"]):
        m = masks.pop()
        skip = hiddens.pop()
        h = torch.cat([h[:, :, :skip.shape[-1]], skip], dim=1)                       # :309
        h = _resnet(sd, f"up_blocks.{i}.0.", h, m, temb, causal)
        h = _stage_transformers(sd, f"up_blocks.{i}.1.", h, m, a["n_blocks"])
        if i < a["n_up"] - 1:              # Upsample1D: ConvTranspose1d(k4, s2, p1), components/decoder.py:142
            h = F.conv_transpose1d(h * m, sd[f"up_blocks.{i}.2.conv.weight"], sd[f"up_blocks.{i}.2.conv.bias"],
                                   stride=2, padding=1)
        else:
            h = _conv_k3(h * m, sd[f"up_blocks.{i}.2.weight"], sd[f"up_blocks.{i}.2.bias"], causal)
        if taps is not None:
            taps[f"up{i}"] = h
    h = _block1d(sd, "final_block.", h, m, causal)
    out = F.conv1d(h * m, sd["final_proj.weight"], sd["final_proj.bias"])
    return out * mask


# ----------------------------------------------------------------------------- CFM solver
def cosine_t_span(n_timesteps: int, dtype=torch.float32) -> torch.Tensor:
    """flow_matching.py:67-69."""
    s = torch.linspace(0, 1, n_timesteps + 1, dtype=dtype)
    return 1 - torch.cos(s * 0.5 * torch.pi)


def step_times(t_span: torch.Tensor) -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
    """The (t, dt) pairs solve_euler actually uses: t is a running fp32 sum and
    dt = t_span[k+1] - t (flow_matching.py:87,118-122), NOT t_span[k] / diff(t_span)."""
    t, dt = t_span[0], t_span[1] - t_span[0]
    ts, dts = [], []
    for step in range(1, len(t_span)):
        ts.append(t.clone())
        dts.append(dt.clone())
        t = t + dt
        if step < len(t_span) - 1:
            dt = t_span[step + 1] - t
    return ts, dts


def solve_euler(sd: SD, z, t_span, mu, mask, spks, cond, causal: bool = False, cfg_rate: float = CFG_RATE):
    """flow_matching.py:72-124 for one utterance (B=1): CFG pair [cond, uncond(zeros)]."""
    x = z
    ts, dts = step_times(t_span)
    zeros_mu, zeros_sp, zeros_c = torch.zeros_like(mu), torch.zeros_like(spks), torch.zeros_like(cond)
    for t, dt in zip(ts, dts):
        x_in = torch.cat([x, x], 0)
        v = estimator_forward(sd, x_in, torch.cat([mask, mask], 0), torch.cat([mu, zeros_mu], 0),
                              t.reshape(1).expand(2).to(x.dtype), torch.cat([spks, zeros_sp], 0),
                              torch.cat([cond, zeros_c], 0), causal=causal)
        v = (1.0 + cfg_rate) * v[:1] - cfg_rate * v[1:]
        x = x + dt * v
    return x.float()


def cfm_forward(sd: SD, mu, mask, spks, cond, z, n_timesteps: int = 10, temperature: float = 1.0,
                prompt_len: int = 0, flow_cache: Optional[torch.Tensor] = None, causal: bool = False):
    """ConditionalCFM.forward (flow_matching.py:37-70) with the noise INJECTED as `z`
    (the tensor randn_like / rand_noise would have produced).  Returns (mel, flow_cache);
    flow_cache is None for the causal variant (:240).  Note: like the reference (:62) this
    overwrites the first cache_size frames of mu."""
    z = z * temperature
    new_cache = None
    if not causal:
        if flow_cache is None:
            flow_cache = torch.zeros(1, 80, 0, 2)
        cache_size = min(flow_cache.shape[2], mu.shape[2])
        if cache_size != 0:
            z = z.clone()
            mu = mu.clone()
            z[:, :, :cache_size] = flow_cache[:, :, :cache_size, 0]
            mu[:, :, :cache_size] = flow_cache[:, :, :cache_size, 1]
        z_cache = torch.cat([z[:, :, :prompt_len], z[:, :, -FLOW_CACHE_TAIL:]], dim=2)
        mu_cache = torch.cat([mu[:, :, :prompt_len], mu[:, :, -FLOW_CACHE_TAIL:]], dim=2)
        new_cache = torch.stack([z_cache, mu_cache], dim=-1)
    t_span = cosine_t_span(n_timesteps, mu.dtype)
    return solve_euler(sd, z, t_span, mu, mask, spks, cond, causal=causal), new_cache
