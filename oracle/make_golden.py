"""Generate tests/golden/*.pt from the REAL reference modules (build container only).

    python -m oracle.make_golden

Weights are NOT stored (82 M parameters): each fixture records the seed given to
oracle/init_weights.py; the reference module is loaded from that state_dict with
strict key checking, which also pins the key/shape contract of the weight packer.
Inputs, injected RNG tensors and the reference's outputs are stored (all small).
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_import as R          # noqa: E402
from oracle import init_weights as W        # noqa: E402
from oracle import hift_ref                 # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _save(name, obj):
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, name)
    torch.save(obj, path)
    print(f"wrote {path}  ({os.path.getsize(path) / 1024:.0f} KiB)")


def _gen(seed):
    g = torch.Generator()
    g.manual_seed(seed)
    return g


def estimator(causal: bool):
    est = R.build_estimator(causal).eval()
    est.load_state_dict(W.flow_estimator_sd(causal, seed=0), strict=True)
    g = _gen(100 + causal)
    T = 40 if causal else 45
    x, mu, cond = (torch.randn(2, 80, T, generator=g) for _ in range(3))
    spks = torch.randn(2, 80, generator=g)
    t = torch.tensor([0.1, 0.7])
    mask = torch.ones(2, 1, T)
    with torch.no_grad():
        out = est(x, mask, mu, t, spks, cond)
    _save(f"estimator_{'v2' if causal else 'v1'}.pt",
          dict(weights_seed=0, causal=causal, x=x, mask=mask, mu=mu, t=t, spks=spks, cond=cond, out=out))


def cfm(causal: bool):
    est = R.build_estimator(causal).eval()
    est.load_state_dict(W.flow_estimator_sd(causal, seed=0), strict=True)
    m = R.build_cfm(causal, est)
    g = _gen(200 + causal)
    T, prompt_len = (30, 8) if causal else (61, 11)
    mu = torch.randn(1, 80, T, generator=g)
    spks = torch.randn(1, 80, generator=g)
    cond = torch.zeros(1, 80, T)
    cond[:, :, :prompt_len] = torch.randn(1, 80, prompt_len, generator=g) * 2 - 4
    mask = torch.ones(1, 1, T)
    d = dict(weights_seed=0, causal=causal, mu=mu, mask=mask, spks=spks, cond=cond, prompt_len=prompt_len, n_timesteps=10)
    with torch.no_grad():
        if causal:
            mel, _ = m(mu.clone(), mask, 10, spks=spks, cond=cond)
            d.update(z=m.rand_noise[:, :, :T].clone(), mel=mel, rand_noise_head=m.rand_noise[:, :, :64].clone())
        else:
            torch.manual_seed(4321)
            mel, cache = m(mu.clone(), mask, 10, spks=spks, cond=cond, prompt_len=prompt_len)
            torch.manual_seed(4321)
            z = torch.randn_like(mu)
            # second call re-using the cache (streaming, flow_matching.py:58-65)
            mu2 = torch.randn(1, 80, T, generator=g)
            torch.manual_seed(4322)
            mel2, cache2 = m(mu2.clone(), mask, 10, spks=spks, cond=cond, prompt_len=prompt_len, flow_cache=cache)
            torch.manual_seed(4322)
            z2 = torch.randn_like(mu)
            d.update(z=z, mel=mel, flow_cache=cache, mu2=mu2, z2=z2, mel2=mel2, flow_cache2=cache2)
    _save(f"cfm_{'v2' if causal else 'v1'}.pt", d)


def hift(v2: bool):
    m = R.build_hift(v2).eval()
    sd = W.hift_sd(v2, seed=1)
    m.load_state_dict(sd, strict=True)
    g = _gen(300 + v2)
    T = 11 if v2 else 17
    hop = hift_ref.arch(v2)["hop"]
    while True:     # pick a mel whose predicted f0 covers voiced (>10 Hz) and unvoiced frames
        mel = torch.randn(1, 80, T, generator=g) * 2 - 4
        with torch.no_grad():
            f0 = m.f0_predictor(mel)
        if (f0 > 10).sum() >= 3 and (f0 < 10).sum() >= 2:
            break
    torch.manual_seed(77)
    with torch.no_grad():
        wav, s = m.inference(mel)
    torch.manual_seed(77)
    phase_vec, noise = hift_ref.draw_source_rng(1, T * hop)
    cache_source = s[:, :, :hop * 2].clone() * 0.5
    torch.manual_seed(78)
    with torch.no_grad():
        wav_c, s_c = m.inference(mel, cache_source=cache_source)
        f0 = m.f0_predictor(mel)
    torch.manual_seed(78)
    phase_vec_c, noise_c = hift_ref.draw_source_rng(1, T * hop)
    _save(f"hift_{'v2' if v2 else 'v1'}.pt",
          dict(weights_seed=1, v2=v2, mel=mel, phase_vec=phase_vec, noise=noise, wav=wav, s=s, f0=f0,
               cache_source=cache_source, phase_vec_c=phase_vec_c, noise_c=noise_c, wav_c=wav_c, s_c=s_c))


def act1d():
    R.setup()
    from cosyvoice.BigVGAN.alias_free_activation.torch.act import Activation1d
    from cosyvoice.BigVGAN.nnet.activations import SnakeBeta
    g = _gen(400)
    C, T = 24, 37
    act = SnakeBeta(C, alpha_logscale=True)
    with torch.no_grad():
        act.alpha.copy_(torch.randn(C, generator=g) * 0.4)
        act.beta.copy_(torch.randn(C, generator=g) * 0.4)
    m = Activation1d(activation=act)
    x = torch.randn(2, C, T, generator=g) * 2
    with torch.no_grad():
        y = m(x)
    _save("act1d.pt", dict(x=x, alpha=act.alpha.detach().clone(), beta=act.beta.detach().clone(), y=y,
                           up_filter=m.upsample.filter.reshape(-1).clone(),
                           down_filter=m.downsample.lowpass.filter.reshape(-1).clone()))


def _bigvgan_trunk(m, feat, spk):
    """bigvgan.py:406-436 driven through the REAL module's own layers (the encoders in front of the trunk are out of scope)."""
    with torch.no_grad():
        x = m.conv_pre(feat)
        x = x + m.cond_layer(spk)
        for i in range(m.num_upsamples):
            x = m.ups[i][0](x)
            x = x + m.conds[i](spk)
            xs = None
            for j in range(m.num_kernels):
                r = m.resblocks[i * m.num_kernels + j](x)
                xs = r if xs is None else xs + r
            x = xs / m.num_kernels
        x = m.activation_post(x)
        x = m.conv_post(x)
        return torch.tanh(x).squeeze(1)


def _bigvgan_module(topology, seed):
    m = R.build_bigvgan(topology).eval()
    res = m.load_state_dict(W.bigvgan_sd(topology, seed=seed), strict=False)
    assert not res.unexpected_keys, res.unexpected_keys
    allowed = ("input_embedding.", "encoder_proj.", "mel_proj.")
    bad = [k for k in res.missing_keys if not k.startswith(allowed) and not k.endswith(".filter")]
    assert not bad, bad
    return m


def bigvgan_config4(members=(0, 31, 63), B=64, frames=94):
    """BASELINE config 4 at its own size: the 24 kHz / 100-band trunk on a batch of 64 mels of 94 frames (1 s), inputs
    N(0,1)*2-4 and speaker vectors N(0,1) from torch.Generator seed 5 exactly as tools/bench_configs.py `bigvgan` draws them.
    The REAL reference runs three members one by one (B = 1, as the reference would); only their waveforms are stored, the
    tests regenerate the inputs from the seed."""
    m = _bigvgan_module("24k100", 3)
    g = torch.Generator().manual_seed(5)
    feat = torch.randn(B, 100, frames, generator=g) * 2 - 4
    spk = torch.randn(B, 512, 1, generator=g)
    wavs = {int(i): _bigvgan_trunk(m, feat[i:i + 1], spk[i:i + 1])[0].contiguous() for i in members}
    _save("bigvgan_config4.pt", dict(weights_seed=3, topology="24k100", input_seed=5, batch=B, frames=frames, wavs=wavs))


FORK_POST_GAIN = 0.05     # see bigvgan(): keeps the fork topology's synthetic output out of tanh saturation


def bigvgan(topology="24k100", resblock="1"):
    m = R.build_bigvgan(topology, resblock).eval()
    sd = W.bigvgan_sd(topology, seed=2, resblock=resblock)
    post_gain = 1.0
    if topology == "fork":
        # with the synthetic weights the 1024-channel input drives conv_post far into tanh saturation (rms 0.92): a clipped
        # waveform pins nothing.  Both sides scale conv_post's weight-norm gain; the fixture records the factor.
        post_gain = FORK_POST_GAIN
        sd["conv_post.weight_g"] = sd["conv_post.weight_g"] * post_gain
        sd["conv_post.bias"] = sd["conv_post.bias"] * post_gain
    res = m.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys, res.unexpected_keys
    allowed = ("input_embedding.", "encoder_proj.", "mel_proj.")
    bad = [k for k in res.missing_keys if not k.startswith(allowed) and not k.endswith(".filter")]
    assert not bad, bad
    g = _gen(500)
    B, T = 2, 9
    in_ch = 100 if topology == "24k100" else 1024
    feat = torch.randn(B, in_ch, T, generator=g)
    spk = torch.randn(B, 512, 1, generator=g)
    wav = _bigvgan_trunk(m, feat, spk)
    name = f"bigvgan_{topology}.pt" if resblock == "1" else f"bigvgan_{topology}_amp{resblock}.pt"
    _save(name, dict(weights_seed=2, topology=topology, resblock=resblock, feat=feat, spk=spk, wav=wav, post_gain=post_gain))


def token2wav():
    """The REAL CosyVoiceModel / CosyVoice2Model (cli/model.py) driven over the stub flow / stub vocoder of
    oracle/token2wav_ref.py: pins the restatement of the streaming glue (overlap fade, hift cache, speed)."""
    R.setup()
    from types import SimpleNamespace
    from cosyvoice.cli.model import CosyVoiceModel, CosyVoice2Model
    from oracle import token2wav_ref as T
    out = {}
    m1 = CosyVoiceModel(llm=SimpleNamespace(), flow=T.StubFlow(), hift=T.StubHift(T.SCENARIO_V1["hop"]), fp16=False,
                        sr=T.SCENARIO_V1["sr"])
    m1.device = torch.device("cpu")
    outs, state = T.run_scenario_v1(m1)
    out["v1"] = dict(outs=[o.clone() for o in outs], state={k: (v.clone() if torch.is_tensor(v) else v)
                                                            for k, v in state.items()},
                     flow_calls=list(m1.flow.calls))
    m2 = CosyVoice2Model(llm=SimpleNamespace(), flow=T.StubFlow(input_frame_rate=25), hift=T.StubHift(T.SCENARIO_V2["hop"]),
                         fp16=False)
    m2.device = torch.device("cpu")
    outs, state = T.run_scenario_v2(m2)
    out["v2"] = dict(outs=[o.clone() for o in outs], state={k: v.clone() for k, v in state.items()},
                     flow_calls=list(m2.flow.calls))
    # the reference's fade_in_out alone, on random data (utils/common.py:149-157)
    from cosyvoice.utils.common import fade_in_out
    g = _gen(900)
    a, b = torch.randn(1, 80, 50, generator=g), torch.randn(1, 80, 34, generator=g)
    w = np.hamming(2 * 34)
    out["fade"] = dict(a=a, b=b, n=34, out=fade_in_out(a, b, w))
    x = torch.randn(1, 80, 37, generator=g)
    out["interp"] = dict(x=x, sizes=[28, 37, 46, 111],
                         outs=[torch.nn.functional.interpolate(x, size=n, mode="linear") for n in (28, 37, 46, 111)])
    _save("token2wav.pt", out)


class _CaptureDecoder(torch.nn.Module):
    """Stands in for flow.decoder: records what flow.inference hands to the CFM (flow/flow.py:150-158)."""

    def forward(self, mu, mask, spks, cond, n_timesteps, prompt_len=0, flow_cache=None):
        self.seen = dict(mu=mu.clone(), mask=mask.clone(), spks=spks.clone(), cond=cond.clone(), prompt_len=prompt_len)
        return torch.zeros_like(mu), flow_cache


def frontend(v2: bool):
    """flow.inference of the REAL MaskedDiffWithXvec / CausalMaskedDiffWithXvec: (a) with the decoder replaced by a
    recorder -> the front-end's outputs (mu, spks, cond); (b) with the real CFM + estimator -> the mel."""
    est = R.build_estimator(v2).eval()
    est.load_state_dict(W.flow_estimator_sd(v2, seed=0), strict=True)
    m = R.build_flow(v2, est).eval()
    res = m.load_state_dict(W.flow_frontend_sd(v2, seed=2), strict=False)
    assert not res.unexpected_keys and all(k.startswith("decoder.") for k in res.missing_keys), res
    real_decoder = m.decoder
    cases = []
    if v2:
        specs = [dict(n=23, n_prompt=9, chunk=0, trim=True), dict(n=23, n_prompt=9, chunk=5, trim=False),
                 dict(n=31, n_prompt=0, chunk=0, trim=False)]
    else:
        specs = [dict(n=50, n_prompt=12), dict(n=30, n_prompt=0), dict(n=41, n_prompt=7)]
    for ci, sp in enumerate(specs):
        token, prompt_token, prompt_feat, embedding = W.frontend_inputs(v2, sp["n"], sp["n_prompt"], seed=500 + 10 * v2 + ci)
        if v2 and sp["trim"]:       # odd prompt length -> the trim branch of flow.py:279-283
            g = _gen(77)
            prompt_feat = torch.cat([prompt_feat, torch.randn(1, 1, 80, generator=g)], dim=1)
        d = dict(sp, token=token, prompt_token=prompt_token, prompt_feat=prompt_feat, embedding=embedding)

        def call():
            kw = dict(token=token.clone(), token_len=torch.tensor([token.shape[1]], dtype=torch.int32),
                      prompt_token=prompt_token.clone(),
                      prompt_token_len=torch.tensor([prompt_token.shape[1]], dtype=torch.int32),
                      prompt_feat=prompt_feat.clone(),
                      prompt_feat_len=torch.tensor([prompt_feat.shape[1]], dtype=torch.int32), embedding=embedding.clone())
            if v2:
                m.encoder.static_chunk_size = sp["chunk"]
                return m.inference(**kw, n_timesteps=10)
            return m.inference(**kw, flow_cache=torch.zeros(1, 80, 0, 2), sample_rate=22050, n_timesteps=10)
        with torch.no_grad():
            m.decoder = _CaptureDecoder()
            call()
            seen = m.decoder.seen
            m.decoder = real_decoder
            torch.manual_seed(4400 + ci)
            mel, cache = call()
        T = seen["mu"].shape[2]
        torch.manual_seed(4400 + ci)
        z = real_decoder.rand_noise[:, :, :T].clone() if v2 else torch.randn(1, 80, T)
        d.update(mu=seen["mu"], spks=seen["spks"], cond=seen["cond"], prompt_len=seen["prompt_len"], z=z, mel=mel,
                 flow_cache=cache)
        cases.append(d)
    _save(f"frontend_{'v2' if v2 else 'v1'}.pt", dict(weights_seed=2, estimator_seed=0, v2=v2, cases=cases))


def fullsize():
    """BASELINE-size utterances (75 prompt + 250 tokens, 24 kHz) through the REAL flow.inference of both model versions:
    v1 T = 281 + 937 (noise: torch.randn under manual_seed(4321), what randn_like draws), v2 T = 150 + 500 (rand_noise).
    Inputs are regenerated from their seeds by the tests; only the generated mels are stored."""
    out = {}
    for v2 in (False, True):
        est = R.build_estimator(v2).eval()
        est.load_state_dict(W.flow_estimator_sd(v2, seed=0), strict=True)
        m = R.build_flow(v2, est).eval()
        m.load_state_dict(W.flow_frontend_sd(v2, seed=2), strict=False)
        token, prompt_token, prompt_feat, embedding = W.frontend_inputs(v2, 250, 75, seed=1234, sample_rate=24000)
        kw = dict(token=token, token_len=torch.tensor([250], dtype=torch.int32), prompt_token=prompt_token,
                  prompt_token_len=torch.tensor([75], dtype=torch.int32), prompt_feat=prompt_feat,
                  prompt_feat_len=torch.tensor([prompt_feat.shape[1]], dtype=torch.int32), embedding=embedding)
        with torch.no_grad():
            torch.manual_seed(4321)
            if v2:
                m.encoder.static_chunk_size = 0
                mel, _ = m.inference(**kw, n_timesteps=10)
            else:
                mel, _ = m.inference(**kw, flow_cache=torch.zeros(1, 80, 0, 2), sample_rate=24000, n_timesteps=10)
        out["v2" if v2 else "v1"] = dict(mel=mel.float().contiguous(), input_seed=1234, noise_seed=None if v2 else 4321)
    _save("fullsize.pt", out)


if __name__ == "__main__":
    import warnings
    warnings.filterwarnings("ignore")
    torch.set_num_threads(8)
    estimator(False)
    estimator(True)
    cfm(False)
    cfm(True)
    hift(False)
    hift(True)
    act1d()
    bigvgan("24k100")
    bigvgan("fork")
    bigvgan("24k100", resblock="2")
    bigvgan_config4()
    token2wav()
    frontend(False)
    frontend(True)
    fullsize()
