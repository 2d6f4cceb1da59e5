"""Random-init state_dicts with the REFERENCE's parameter names and shapes (TEST INFRASTRUCTURE).

There are no checkpoints on the build or GPU boxes, so tests and benchmarks use random
weights "of the named architecture".  Key names/shapes are those produced by the reference
constructors (checked in the build container by `load_state_dict(strict=True)` inside
oracle/make_golden.py).  Distributions follow the reference initialisers
(ConditionalDecoder.initialize_weights flow/decoder.py:208-220 = kaiming-normal;
HiFT init_weights utils/common.py:102-105 = N(0, 0.01); torch defaults elsewhere) except
that biases / norm affines / snake alphas are made non-trivial so that parity tests
exercise them.
"""
import math
from typing import Dict

import torch

SD = Dict[str, torch.Tensor]


def _g(seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return g


def _kaiming(g, *shape):
    fan_in = 1
    for s in shape[1:]:
        fan_in *= s
    return torch.randn(*shape, generator=g) * math.sqrt(2.0 / fan_in)


def _uniform_default(g, shape, fan_in):
    b = 1.0 / math.sqrt(fan_in)
    return (torch.rand(*shape, generator=g) * 2 - 1) * b


def _bias(g, n, std=0.05):
    return torch.randn(n, generator=g) * std


def _affine(g, sd, p_w, p_b, n):
    sd[p_w] = 1.0 + 0.1 * torch.randn(n, generator=g)
    sd[p_b] = 0.1 * torch.randn(n, generator=g)


# ----------------------------------------------------------------------------- flow estimator
def flow_estimator_sd(causal: bool, seed: int = 0) -> SD:
    """ConditionalDecoder(in=320,out=80, channels [256,256] | causal [256], n_blocks 4, mid 12, heads 8x64)."""
    g = _g(seed)
    sd: SD = {}
    C, TE = 256, 1024
    sd["time_mlp.linear_1.weight"] = _kaiming(g, TE, 320)
    sd["time_mlp.linear_1.bias"] = _bias(g, TE)
    sd["time_mlp.linear_2.weight"] = _kaiming(g, TE, TE)
    sd["time_mlp.linear_2.bias"] = _bias(g, TE)
    norm_idx = 2 if causal else 1

    def block1d(p, cin):
        sd[p + "block.0.weight"] = _kaiming(g, C, cin, 3)
        sd[p + "block.0.bias"] = _bias(g, C)
        _affine(g, sd, p + f"block.{norm_idx}.weight", p + f"block.{norm_idx}.bias", C)

    def resnet(p, cin):
        sd[p + "mlp.1.weight"] = _kaiming(g, C, TE)
        sd[p + "mlp.1.bias"] = _bias(g, C)
        block1d(p + "block1.", cin)
        block1d(p + "block2.", C)
        sd[p + "res_conv.weight"] = _kaiming(g, C, cin, 1)
        sd[p + "res_conv.bias"] = _bias(g, C)

    def tblock(p):
        _affine(g, sd, p + "norm1.weight", p + "norm1.bias", C)
        for n in ("to_q", "to_k", "to_v"):
            sd[p + f"attn1.{n}.weight"] = _kaiming(g, 512, C)
        sd[p + "attn1.to_out.0.weight"] = _kaiming(g, C, 512)
        sd[p + "attn1.to_out.0.bias"] = _bias(g, C)
        _affine(g, sd, p + "norm3.weight", p + "norm3.bias", C)
        sd[p + "ff.net.0.proj.weight"] = _kaiming(g, 1024, C)
        sd[p + "ff.net.0.proj.bias"] = _bias(g, 1024)
        sd[p + "ff.net.2.weight"] = _kaiming(g, C, 1024)
        sd[p + "ff.net.2.bias"] = _bias(g, C)

    def stage(p, cin):
        resnet(p + "0.", cin)
        for i in range(4):
            tblock(p + f"1.{i}.")

    n_down = 1 if causal else 2
    for i in range(n_down):
        p = f"down_blocks.{i}."
        stage(p, 320 if i == 0 else C)
        if i < n_down - 1:      # Downsample1D
            sd[p + "2.conv.weight"] = _kaiming(g, C, C, 3)
            sd[p + "2.conv.bias"] = _bias(g, C)
        else:
            sd[p + "2.weight"] = _kaiming(g, C, C, 3)
            sd[p + "2.bias"] = _bias(g, C)
    for i in range(12):
        stage(f"mid_blocks.{i}.", C)
    for i in range(n_down):
        p = f"up_blocks.{i}."
        stage(p, 2 * C)
        if i < n_down - 1:      # Upsample1D: ConvTranspose1d(C, C, 4, 2, 1), weight [Cin, Cout, k]
            sd[p + "2.conv.weight"] = _uniform_default(g, (C, C, 4), C * 4) * 1.5
            sd[p + "2.conv.bias"] = _bias(g, C)
        else:
            sd[p + "2.weight"] = _kaiming(g, C, C, 3)
            sd[p + "2.bias"] = _bias(g, C)
    block1d("final_block.", C)
    sd["final_proj.weight"] = _kaiming(g, 80, C, 1)
    sd["final_proj.bias"] = _bias(g, 80)
    return sd


# ----------------------------------------------------------------------------- weight-norm helper
def _wn(g, sd, p, shape, std=None, fan_in=None, g_jitter=0.1):
    """old-style weight_norm pair: weight_v [shape], weight_g [shape[0],1,1]."""
    v = torch.randn(*shape, generator=g) * std if std is not None else _uniform_default(g, shape, fan_in)
    n = v.reshape(shape[0], -1).norm(dim=1).reshape(-1, 1, 1)
    sd[p + "weight_v"] = v
    sd[p + "weight_g"] = n * (1.0 + g_jitter * torch.randn(shape[0], 1, 1, generator=g))


def _resblock(g, sd, p, ch, k, std):
    for name in ("convs1", "convs2"):
        for i in range(3):
            _wn(g, sd, f"{p}{name}.{i}.", (ch, ch, k), std=std)
            sd[f"{p}{name}.{i}.bias"] = _bias(g, ch, 0.02)
    for name in ("activations1", "activations2"):
        for i in range(3):
            sd[f"{p}{name}.{i}.alpha"] = 1.0 + 0.3 * torch.randn(ch, generator=g)


# ----------------------------------------------------------------------------- HiFT
def hift_sd(v2: bool, seed: int = 0) -> SD:
    """HiFTGenerator(base 512, nb_harmonics 8, rates [8,8] | [8,5,3]) + ConvRNNF0Predictor(1,80,512)."""
    g = _g(seed)
    sd: SD = {}
    rates, ks, src_k = ([8, 5, 3], [16, 11, 7], [7, 7, 11]) if v2 else ([8, 8], [16, 16], [7, 11])
    sd["m_source.l_linear.weight"] = _uniform_default(g, (1, 9), 9) * 3
    sd["m_source.l_linear.bias"] = _bias(g, 1)
    _wn(g, sd, "conv_pre.", (512, 80, 7), fan_in=80 * 7)
    sd["conv_pre.bias"] = _bias(g, 512)
    for i, (u, k) in enumerate(zip(rates, ks)):
        cin, cout = 512 // 2 ** i, 512 // 2 ** (i + 1)
        _wn(g, sd, f"ups.{i}.", (cin, cout, k), std=0.03)
        sd[f"ups.{i}.bias"] = _bias(g, cout, 0.02)
    down = [1] + rates[::-1][:-1]
    cum = []
    c = 1
    for d in down:
        c *= d
        cum.append(c)
    cum = cum[::-1]
    for i, u in enumerate(cum):
        ch = 512 // 2 ** (i + 1)
        k = 1 if u == 1 else 2 * u
        sd[f"source_downs.{i}.weight"] = _uniform_default(g, (ch, 18, k), 18 * k)
        sd[f"source_downs.{i}.bias"] = _bias(g, ch, 0.02)
        _resblock(g, sd, f"source_resblocks.{i}.", ch, src_k[i], std=0.02)
    for i in range(len(rates)):
        ch = 512 // 2 ** (i + 1)
        for j, k in enumerate((3, 7, 11)):
            _resblock(g, sd, f"resblocks.{i * 3 + j}.", ch, k, std=0.02)
    ch = 512 // 2 ** len(rates)
    _wn(g, sd, "conv_post.", (18, ch, 7), std=0.03)
    sd["conv_post.bias"] = _bias(g, 18, 0.1)
    # f0 predictor: 5 weight-normed convs + classifier; classifier scaled so f0 spans voiced/unvoiced
    cin = 80
    for i in range(5):
        _wn(g, sd, f"f0_predictor.condnet.{2 * i}.", (512, cin, 3), fan_in=cin * 3)
        sd[f"f0_predictor.condnet.{2 * i}.bias"] = _bias(g, 512)
        cin = 512
    sd["f0_predictor.classifier.weight"] = torch.randn(1, 512, generator=g) * 80.0
    sd["f0_predictor.classifier.bias"] = torch.full((1,), 30.0)
    return sd


# ----------------------------------------------------------------------------- BigVGAN trunk
def bigvgan_sd(topology: str = "24k100", seed: int = 0, resblock: str = "1") -> SD:
    """Trunk parameters only (conv_pre, cond layers, ups, AMP blocks, activation_post, conv_post).  resblock "1" = AMPBlock1
    (convs1 / convs2, six activations per block), "2" = AMPBlock2 (convs, three activations; BigVGAN/bigvgan.py:146-232)."""
    g = _g(seed)
    sd: SD = {}
    if topology == "24k100":
        in_ch, rates, ks = 100, [4, 4, 2, 2, 2, 2], [8, 8, 4, 4, 4, 4]
    elif topology == "fork":
        in_ch, rates, ks = 1024, [8, 5, 3, 2], [16, 11, 7, 2]
    else:
        raise ValueError(topology)
    ch0, spk = 1536, 512
    _wn(g, sd, "conv_pre.", (ch0, in_ch, 7), fan_in=in_ch * 7)
    sd["conv_pre.bias"] = _bias(g, ch0)
    sd["cond_layer.weight"] = _uniform_default(g, (ch0, spk, 1), spk)
    sd["cond_layer.bias"] = _bias(g, ch0)
    for i, (u, k) in enumerate(zip(rates, ks)):
        cin, cout = ch0 // 2 ** i, ch0 // 2 ** (i + 1)
        _wn(g, sd, f"ups.{i}.0.", (cin, cout, k), std=0.03)
        sd[f"ups.{i}.0.bias"] = _bias(g, cout, 0.02)
        sd[f"conds.{i}.weight"] = _uniform_default(g, (cout, spk, 1), spk)
        sd[f"conds.{i}.bias"] = _bias(g, cout, 0.02)
        for j, rk in enumerate((3, 7, 11)):
            p = f"resblocks.{i * 3 + j}."
            for name in (("convs1", "convs2") if resblock == "1" else ("convs",)):
                for d in range(3):
                    _wn(g, sd, f"{p}{name}.{d}.", (cout, cout, rk), std=0.02)
                    sd[f"{p}{name}.{d}.bias"] = _bias(g, cout, 0.02)
            for a in range(6 if resblock == "1" else 3):
                sd[f"{p}activations.{a}.act.alpha"] = 0.3 * torch.randn(cout, generator=g)
                sd[f"{p}activations.{a}.act.beta"] = 0.3 * torch.randn(cout, generator=g)
    ch = ch0 // 2 ** len(rates)
    sd["activation_post.act.alpha"] = 0.3 * torch.randn(ch, generator=g)
    sd["activation_post.act.beta"] = 0.3 * torch.randn(ch, generator=g)
    _wn(g, sd, "conv_post.", (1, ch, 7), std=0.01)
    sd["conv_post.bias"] = _bias(g, 1, 0.02)
    return sd


# ----------------------------------------------------------------------------- synthetic inputs
def flow_inputs(n_utt: int, gen_frames, prompt_frames: int = 281, first_index: int = 0):
    """Per-utterance CFM inputs in the reference's layout: mu/cond/z [1,80,T], spks [1,80]; T = prompt + generated.
    Seed 1234+u per utterance index u; cond holds a log-mel-like prompt (N(0,1)*2-4) in its first prompt frames."""
    if isinstance(gen_frames, int):
        gen_frames = [gen_frames] * n_utt
    mus, spks, conds, zs = [], [], [], []
    for i in range(n_utt):
        g = _g(1234 + first_index + i)
        T = prompt_frames + int(gen_frames[i])
        mus.append(torch.randn(1, 80, T, generator=g))
        spks.append(torch.randn(1, 80, generator=g))
        c = torch.zeros(1, 80, T)
        c[:, :, :prompt_frames] = torch.randn(1, 80, prompt_frames, generator=g) * 2 - 4
        conds.append(c)
        zs.append(torch.randn(1, 80, T, generator=_g(4321 + first_index + i)))
    return mus, spks, conds, zs


# ----------------------------------------------------------------------------- flow front-end
def flow_frontend_sd(v2: bool, seed: int = 2) -> SD:
    """Everything of MaskedDiffWithXvec / CausalMaskedDiffWithXvec except `decoder.*`: input_embedding,
    spk_embed_affine_layer, encoder (ConformerEncoder | UpsampleConformerEncoder, d 512, 8 heads, FFN 2048, 6 (+4) blocks,
    rel-pos attention), encoder_proj and, v1 only, the InterpolateRegulator (flow/flow.py:25-62,163-203 with the yaml
    numbers of examples/tts_vc/cosyvoice{,2}/conf/cosyvoice.yaml)."""
    g = _g(seed)
    sd: SD = {}
    D, H, FF, V, SPK = 512, 8, 2048, (6561 if v2 else 4096), (192 if v2 else 512)
    sd["input_embedding.weight"] = torch.randn(V, D, generator=g)
    sd["spk_embed_affine_layer.weight"] = _uniform_default(g, (80, SPK), SPK)
    sd["spk_embed_affine_layer.bias"] = _bias(g, 80)

    def lin(p, n, k, bias=True):
        sd[p + "weight"] = _uniform_default(g, (n, k), k)
        if bias:
            sd[p + "bias"] = _bias(g, n)

    def embed(p):
        lin(p + "out.0.", D, D)
        _affine(g, sd, p + "out.1.weight", p + "out.1.bias", D)

    def layer(p):
        for n in ("q", "k", "v", "out"):
            lin(p + f"self_attn.linear_{n}.", D, D)
        lin(p + "self_attn.linear_pos.", D, D, bias=False)
        b = math.sqrt(6.0 / (H + D // H))          # xavier_uniform on [H, D/H] (transformer/attention.py:219-220)
        sd[p + "self_attn.pos_bias_u"] = (torch.rand(H, D // H, generator=g) * 2 - 1) * b
        sd[p + "self_attn.pos_bias_v"] = (torch.rand(H, D // H, generator=g) * 2 - 1) * b
        lin(p + "feed_forward.w_1.", FF, D)
        lin(p + "feed_forward.w_2.", D, FF)
        _affine(g, sd, p + "norm_ff.weight", p + "norm_ff.bias", D)
        _affine(g, sd, p + "norm_mha.weight", p + "norm_mha.bias", D)

    embed("encoder.embed.")
    if v2:
        sd["encoder.pre_lookahead_layer.conv1.weight"] = _uniform_default(g, (D, D, 4), D * 4)
        sd["encoder.pre_lookahead_layer.conv1.bias"] = _bias(g, D)
        sd["encoder.pre_lookahead_layer.conv2.weight"] = _uniform_default(g, (D, D, 3), D * 3)
        sd["encoder.pre_lookahead_layer.conv2.bias"] = _bias(g, D)
    for i in range(6):
        layer(f"encoder.encoders.{i}.")
    if v2:
        sd["encoder.up_layer.conv.weight"] = _uniform_default(g, (D, D, 5), D * 5)
        sd["encoder.up_layer.conv.bias"] = _bias(g, D)
        embed("encoder.up_embed.")
        for i in range(4):
            layer(f"encoder.up_encoders.{i}.")
    _affine(g, sd, "encoder.after_norm.weight", "encoder.after_norm.bias", D)
    lin("encoder_proj.", 80, D)
    if not v2:
        for i in range(4):
            sd[f"length_regulator.model.{3 * i}.weight"] = _uniform_default(g, (80, 80, 3), 240)
            sd[f"length_regulator.model.{3 * i}.bias"] = _bias(g, 80)
            _affine(g, sd, f"length_regulator.model.{3 * i + 1}.weight", f"length_regulator.model.{3 * i + 1}.bias", 80)
        sd["length_regulator.model.12.weight"] = _uniform_default(g, (80, 80, 1), 80)
        sd["length_regulator.model.12.bias"] = _bias(g, 80)
    return sd


def frontend_inputs(v2: bool, n_tokens: int, n_prompt_tokens: int, seed: int, sample_rate: int = 22050):
    """token / prompt_token int32 [1, n], prompt_feat [1, T_p, 80] time-major log-mel-like, embedding [1, spk].
    T_p = the reference's own prompt ratio: v1 3.75 frames per token at 22.05 kHz/256 -> int(n / 25 * 22050 / 256);
    v2 2 frames per token."""
    g = _g(seed)
    V, SPK = (6561, 192) if v2 else (4096, 512)
    token = torch.randint(0, V, (1, n_tokens), generator=g, dtype=torch.int32)
    prompt_token = torch.randint(0, V, (1, n_prompt_tokens), generator=g, dtype=torch.int32)
    t_p = 2 * n_prompt_tokens if v2 else int(n_prompt_tokens / 25 * sample_rate / 256)
    prompt_feat = torch.randn(1, t_p, 80, generator=g) * 2 - 4
    embedding = torch.randn(1, SPK, generator=g)
    return token, prompt_token, prompt_feat, embedding
