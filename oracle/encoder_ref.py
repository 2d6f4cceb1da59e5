"""CPU restatement of the flow FRONT-END (TEST INFRASTRUCTURE ONLY): everything `flow.inference` does before the CFM.

Functional torch code over a plain state_dict with the reference's own parameter names (`decoder.*` excluded).

Follows (reference file:line, all under /root/reference/cosyvoice):
  flow/flow.py:108-160                      MaskedDiffWithXvec.inference            -> frontend_v1
  flow/flow.py:258-319                      CausalMaskedDiffWithXvec.inference      -> frontend_v2
  transformer/encoder.py:116-165,388-476    BaseEncoder.forward / ConformerEncoder  -> conformer_encoder
  transformer/upsample_encoder.py:33-96     Upsample1D / PreLookaheadLayer          -> _upsample1d / _pre_lookahead
  transformer/upsample_encoder.py:254-319   UpsampleConformerEncoder.forward        -> upsample_conformer_encoder
  transformer/subsampling.py:69-113         LinearNoSubsampling                     -> _embed
  transformer/embedding.py:201-298          EspnetRelPositionalEncoding             -> rel_pos_table
  transformer/encoder_layer.py:109-236      ConformerEncoderLayer (no macaron/cnn)  -> _layer
  transformer/attention.py:200-330          RelPositionMultiHeadedAttention         -> _rel_attention
  transformer/positionwise_feed_forward.py  PositionwiseFeedForward (SiLU)          -> inside _layer
  utils/mask.py:87-200                      subsequent_chunk_mask / add_optional_chunk_mask -> chunk_mask
  flow/length_regulator.py:20-70            InterpolateRegulator.inference          -> length_regulator_inference

Pinned against the real reference modules by oracle/make_golden.py (tests/golden/frontend_v{1,2}.pt) and
tests/test_oracle_golden.py.
"""
import math
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]
D_MODEL, HEADS, D_K = 512, 8, 64


def rel_pos_table(half: int, d_model: int = D_MODEL) -> torch.Tensor:
    """PE(r) for r = half-1 ... -(half-1) (row index half-1-r), the slice `position_encoding(size=half)` returns
    (embedding.py:224-254,272-298): interleaved sin/cos of r * div_term, all in fp32 like the reference."""
    position = torch.arange(0, half, dtype=torch.float32).unsqueeze(1)
    div_term = torch.exp(torch.arange(0, d_model, 2, dtype=torch.float32) * -(math.log(10000.0) / d_model))
    pos = torch.zeros(half, d_model)
    neg = torch.zeros(half, d_model)
    pos[:, 0::2] = torch.sin(position * div_term)
    pos[:, 1::2] = torch.cos(position * div_term)
    neg[:, 0::2] = torch.sin(-1 * position * div_term)
    neg[:, 1::2] = torch.cos(-1 * position * div_term)
    return torch.cat([torch.flip(pos, [0]), neg[1:]], dim=0)            # [2*half-1, d]


def chunk_mask(T: int, chunk: int) -> Optional[torch.Tensor]:
    """utils/mask.py:87-124 with num_left_chunks = -1: query i sees keys [0, min((i//chunk+1)*chunk, T))."""
    if chunk <= 0:
        return None
    i = torch.arange(T).unsqueeze(1)
    j = torch.arange(T).unsqueeze(0)
    return j < torch.clamp((i // chunk + 1) * chunk, max=T)


def _embed(sd: SD, p: str, x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """LinearNoSubsampling: Linear + LayerNorm(eps 1e-5) (+ dropout), then x * sqrt(d) and the rel-pos table."""
    x = F.linear(x, sd[p + "out.0.weight"], sd[p + "out.0.bias"])
    x = F.layer_norm(x, (D_MODEL,), sd[p + "out.1.weight"], sd[p + "out.1.bias"], 1e-5)
    return x * math.sqrt(D_MODEL), rel_pos_table(x.shape[1]).unsqueeze(0).to(x.dtype)


def _rel_attention(sd: SD, p: str, x: torch.Tensor, pos_emb: torch.Tensor, mask: Optional[torch.Tensor]) -> torch.Tensor:
    """attention.py:243-330: scores[i,j] = ((q_i+u).k_j + (q_i+v).P[T-1-(i-j)]) / sqrt(d_k) (rel_shift written as an
    explicit gather), masked_fill(-inf) + softmax + masked_fill(0), .V, linear_out."""
    B, T, _ = x.shape
    q = F.linear(x, sd[p + "linear_q.weight"], sd[p + "linear_q.bias"]).view(B, T, HEADS, D_K)
    k = F.linear(x, sd[p + "linear_k.weight"], sd[p + "linear_k.bias"]).view(B, T, HEADS, D_K).transpose(1, 2)
    v = F.linear(x, sd[p + "linear_v.weight"], sd[p + "linear_v.bias"]).view(B, T, HEADS, D_K).transpose(1, 2)
    pe = F.linear(pos_emb, sd[p + "linear_pos.weight"]).view(1, -1, HEADS, D_K).transpose(1, 2)     # [1,H,2T-1,dk]
    q_u = (q + sd[p + "pos_bias_u"]).transpose(1, 2)
    q_v = (q + sd[p + "pos_bias_v"]).transpose(1, 2)
    ac = torch.matmul(q_u, k.transpose(-2, -1))
    bd_raw = torch.matmul(q_v, pe.transpose(-2, -1))                                                   # [B,H,T,2T-1]
    i = torch.arange(T).unsqueeze(1)
    j = torch.arange(T).unsqueeze(0)
    idx = (T - 1 - (i - j)).expand(B, HEADS, T, T)
    bd = torch.gather(bd_raw, 3, idx)
    scores = (ac + bd) / math.sqrt(D_K)
    if mask is not None:
        m = ~mask.view(1, 1, T, T)
        attn = torch.softmax(scores.masked_fill(m, -float("inf")), dim=-1).masked_fill(m, 0.0)
    else:
        attn = torch.softmax(scores, dim=-1)
    o = torch.matmul(attn, v).transpose(1, 2).contiguous().view(B, T, HEADS * D_K)
    return F.linear(o, sd[p + "linear_out.weight"], sd[p + "linear_out.bias"])


def _layer(sd: SD, p: str, x: torch.Tensor, pos_emb: torch.Tensor, mask: Optional[torch.Tensor]) -> torch.Tensor:
    """encoder_layer.py:191-236 with macaron_style False, use_cnn_module False, normalize_before True (eps 1e-12)."""
    h = F.layer_norm(x, (D_MODEL,), sd[p + "norm_mha.weight"], sd[p + "norm_mha.bias"], 1e-12)
    x = x + _rel_attention(sd, p + "self_attn.", h, pos_emb, mask)
    h = F.layer_norm(x, (D_MODEL,), sd[p + "norm_ff.weight"], sd[p + "norm_ff.bias"], 1e-12)
    h = F.linear(F.silu(F.linear(h, sd[p + "feed_forward.w_1.weight"], sd[p + "feed_forward.w_1.bias"])),
                 sd[p + "feed_forward.w_2.weight"], sd[p + "feed_forward.w_2.bias"])
    return x + h


def conformer_encoder(sd: SD, xs: torch.Tensor, static_chunk_size: int = 0, p: str = "encoder.") -> torch.Tensor:
    """encoder.py:116-165 at B = 1 (pad mask all true): embed -> 6 layers -> after_norm (eps 1e-5)."""
    x, pos = _embed(sd, p + "embed.", xs)
    mask = chunk_mask(x.shape[1], static_chunk_size)
    for i in range(6):
        x = _layer(sd, f"{p}encoders.{i}.", x, pos, mask)
    return F.layer_norm(x, (D_MODEL,), sd[p + "after_norm.weight"], sd[p + "after_norm.bias"], 1e-5)


def _pre_lookahead(sd: SD, p: str, x: torch.Tensor) -> torch.Tensor:
    """upsample_encoder.py:80-96: right-pad 3, conv k4, leaky_relu(0.01), left-pad 2, conv k3, + input."""
    y = F.pad(x.transpose(1, 2), (0, 3))
    y = F.leaky_relu(F.conv1d(y, sd[p + "conv1.weight"], sd[p + "conv1.bias"]))
    y = F.conv1d(F.pad(y, (2, 0)), sd[p + "conv2.weight"], sd[p + "conv2.bias"])
    return y.transpose(1, 2) + x


def _upsample1d(sd: SD, p: str, x: torch.Tensor) -> torch.Tensor:
    """upsample_encoder.py:55-59: nearest x2, left-pad 4, conv k5."""
    y = F.interpolate(x.transpose(1, 2), scale_factor=2.0, mode="nearest")
    y = F.conv1d(F.pad(y, (4, 0)), sd[p + "conv.weight"], sd[p + "conv.bias"])
    return y.transpose(1, 2)


def upsample_conformer_encoder(sd: SD, xs: torch.Tensor, static_chunk_size: int = 0, p: str = "encoder.") -> torch.Tensor:
    """upsample_encoder.py:254-319 at B = 1: embed, pre-lookahead, 6 layers (chunk c), x2 upsample, up_embed, 4 layers
    (chunk 2c), after_norm."""
    x, pos = _embed(sd, p + "embed.", xs)
    mask = chunk_mask(x.shape[1], static_chunk_size)
    x = _pre_lookahead(sd, p + "pre_lookahead_layer.", x)
    for i in range(6):
        x = _layer(sd, f"{p}encoders.{i}.", x, pos, mask)
    x = _upsample1d(sd, p + "up_layer.", x)
    x, pos = _embed(sd, p + "up_embed.", x)
    mask = chunk_mask(x.shape[1], static_chunk_size * 2)
    for i in range(4):
        x = _layer(sd, f"{p}up_encoders.{i}.", x, pos, mask)
    return F.layer_norm(x, (D_MODEL,), sd[p + "after_norm.weight"], sd[p + "after_norm.bias"], 1e-5)


def regulator_segments(n1: int, n2: int, mel_len1: int, mel_len2: int, input_frame_rate: int = 25, sample_rate: int = 22050):
    """The (src_start, src_len, dst_start, dst_len) pieces of InterpolateRegulator.inference (length_regulator.py:52-70):
    prompt tokens -> mel_len1 frames; generated tokens as head(20)/mid/tail(20) when there are more than 40."""
    segs = []
    if n1 != 0:
        segs.append((0, n1, 0, mel_len1))
    if n2 > 40:
        e = int(20 / input_frame_rate * sample_rate / 256)
        segs.append((n1, 20, mel_len1, e))
        segs.append((n1 + 20, n2 - 40, mel_len1 + e, mel_len2 - 2 * e))
        segs.append((n1 + n2 - 20, 20, mel_len1 + mel_len2 - e, e))
    else:
        segs.append((n1, n2, mel_len1, mel_len2))
    return segs


def length_regulator_inference(sd: SD, h: torch.Tensor, n1: int, mel_len1: int, mel_len2: int, input_frame_rate: int = 25,
                               sample_rate: int = 22050, p: str = "length_regulator.") -> torch.Tensor:
    """h [1, n1+n2, 80] -> [1, mel_len1+mel_len2, 80]: piecewise F.interpolate(mode='linear'), then
    4 x [Conv1d k3 pad 1, GroupNorm(1, 80), Mish] + Conv1d 1x1 (length_regulator.py:30-41,52-70)."""
    n2 = h.shape[1] - n1
    parts = [F.interpolate(h[:, s:s + n].transpose(1, 2).contiguous(), size=dn, mode="linear")
             for (s, n, _, dn) in regulator_segments(n1, n2, mel_len1, mel_len2, input_frame_rate, sample_rate)]
    x = torch.cat(parts, dim=2)
    for i in range(4):
        x = F.conv1d(x, sd[f"{p}model.{3 * i}.weight"], sd[f"{p}model.{3 * i}.bias"], padding=1)
        x = F.mish(F.group_norm(x, 1, sd[f"{p}model.{3 * i + 1}.weight"], sd[f"{p}model.{3 * i + 1}.bias"], 1e-5))
    x = F.conv1d(x, sd[f"{p}model.12.weight"], sd[f"{p}model.12.bias"])
    return x.transpose(1, 2).contiguous()


def speaker_projection(sd: SD, embedding: torch.Tensor) -> torch.Tensor:
    """flow.py:127-128: F.normalize(dim=1) + spk_embed_affine_layer."""
    return F.linear(F.normalize(embedding, dim=1), sd["spk_embed_affine_layer.weight"], sd["spk_embed_affine_layer.bias"])


def frontend_v1(sd: SD, token, prompt_token, prompt_feat, embedding, sample_rate: int = 22050, input_frame_rate: int = 25):
    """flow.py:119-147 -> (mu [1,80,T], spks [1,80], cond [1,80,T], mel_len1, mel_len2)."""
    spks = speaker_projection(sd, embedding)
    n1, n2 = prompt_token.shape[1], token.shape[1]
    tok = torch.cat([prompt_token, token], dim=1)
    x = F.embedding(torch.clamp(tok, min=0).long(), sd["input_embedding.weight"])
    h = conformer_encoder(sd, x)
    h = F.linear(h, sd["encoder_proj.weight"], sd["encoder_proj.bias"])
    mel_len1, mel_len2 = prompt_feat.shape[1], int(n2 / input_frame_rate * sample_rate / 256)
    h = length_regulator_inference(sd, h, n1, mel_len1, mel_len2, input_frame_rate, sample_rate)
    cond = torch.zeros(1, mel_len1 + mel_len2, 80)
    cond[:, :mel_len1] = prompt_feat
    return h.transpose(1, 2).contiguous(), spks, cond.transpose(1, 2).contiguous(), mel_len1, mel_len2


def frontend_v2(sd: SD, token, prompt_token, prompt_feat, embedding, static_chunk_size: int = 0, token_mel_ratio: int = 2):
    """flow.py:274-305 (the odd-prompt trim of :279-283 included) -> (mu, spks, cond, mel_len1, mel_len2)."""
    r = prompt_feat.shape[1] % token_mel_ratio
    if r != 0:
        prompt_feat = prompt_feat[:, :prompt_feat.shape[1] - r]
        prompt_token = prompt_token[:, :prompt_feat.shape[1] // token_mel_ratio]
    spks = speaker_projection(sd, embedding)
    tok = torch.cat([prompt_token, token], dim=1)
    x = F.embedding(tok.long(), sd["input_embedding.weight"])
    h = upsample_conformer_encoder(sd, x, static_chunk_size)
    mel_len1, mel_len2 = prompt_feat.shape[1], h.shape[1] - prompt_feat.shape[1]
    h = F.linear(h, sd["encoder_proj.weight"], sd["encoder_proj.bias"])
    cond = torch.zeros(1, mel_len1 + mel_len2, 80)
    cond[:, :mel_len1] = prompt_feat
    return h.transpose(1, 2).contiguous(), spks, cond.transpose(1, 2).contiguous(), mel_len1, mel_len2
