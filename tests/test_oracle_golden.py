"""Pins the CPU oracle (oracle/*_ref.py) to fixtures produced by the REAL reference
modules (oracle/make_golden.py, run in the build container).  CPU only."""
import os

import pytest
import torch

from oracle import bigvgan_ref, flow_ref, hift_ref, init_weights


def _load(golden_dir, name):
    return torch.load(os.path.join(golden_dir, name), weights_only=True)


def _close(a, b, atol, rtol=0.0):
    err = (a - b).abs().max().item()
    assert a.shape == b.shape
    assert err <= atol + rtol * b.abs().max().item(), f"max-abs err {err}"


@pytest.mark.parametrize("ver", ["v1", "v2"])
def test_estimator_matches_reference(golden_dir, ver):
    d = _load(golden_dir, f"estimator_{ver}.pt")
    sd = init_weights.flow_estimator_sd(d["causal"], seed=d["weights_seed"])
    out = flow_ref.estimator_forward(sd, d["x"], d["mask"], d["mu"], d["t"], d["spks"], d["cond"], causal=d["causal"])
    _close(out, d["out"], atol=5e-5)


def test_cfm_v1_matches_reference_including_flow_cache(golden_dir):
    d = _load(golden_dir, "cfm_v1.pt")
    sd = init_weights.flow_estimator_sd(False, seed=d["weights_seed"])
    mel, cache = flow_ref.cfm_forward(sd, d["mu"].clone(), d["mask"], d["spks"], d["cond"], d["z"],
                                      prompt_len=d["prompt_len"])
    _close(mel, d["mel"], atol=2e-4)
    assert torch.equal(cache, d["flow_cache"])
    mel2, cache2 = flow_ref.cfm_forward(sd, d["mu2"].clone(), d["mask"], d["spks"], d["cond"], d["z2"],
                                        prompt_len=d["prompt_len"], flow_cache=cache)
    _close(mel2, d["mel2"], atol=2e-4)
    assert torch.equal(cache2, d["flow_cache2"])


def test_cfm_v2_matches_reference(golden_dir):
    d = _load(golden_dir, "cfm_v2.pt")
    sd = init_weights.flow_estimator_sd(True, seed=d["weights_seed"])
    mel, cache = flow_ref.cfm_forward(sd, d["mu"].clone(), d["mask"], d["spks"], d["cond"], d["z"], causal=True)
    assert cache is None
    _close(mel, d["mel"], atol=2e-4)


def test_step_times_follow_running_sum():
    ts, dts = flow_ref.step_times(flow_ref.cosine_t_span(10))
    assert len(ts) == 10 and float(ts[0]) == 0.0
    assert abs(float(ts[-1] + dts[-1]) - 1.0) < 1e-6


@pytest.mark.parametrize("ver", ["v1", "v2"])
def test_hift_matches_reference(golden_dir, ver):
    d = _load(golden_dir, f"hift_{ver}.pt")
    sd = init_weights.hift_sd(d["v2"], seed=d["weights_seed"])
    _close(hift_ref.f0_predictor(sd, d["mel"]), d["f0"], atol=1e-3)
    assert (d["f0"] > 10).any() and (d["f0"] < 10).any(), "fixture must cover voiced and unvoiced frames"
    wav, s = hift_ref.hift_inference(sd, d["mel"], d["phase_vec"], d["noise"], v2=d["v2"])
    # the source integrates f0 over thousands of samples (generator.py:148): fp32 rounding of f0
    # (1e-4 Hz) moves the phase by ~1e-4 rad, hence the looser bound on s than on a conv output
    _close(s, d["s"], atol=3e-4)
    _close(wav, d["wav"], atol=1e-3)
    # with the reference's own source injected the decoder is pinned tightly
    _close(hift_ref.decode(sd, d["mel"], d["s"], v2=d["v2"]), d["wav"], atol=2e-5)
    wav, s = hift_ref.hift_inference(sd, d["mel"], d["phase_vec_c"], d["noise_c"],
                                     cache_source=d["cache_source"], v2=d["v2"])
    _close(s, d["s_c"], atol=3e-4)
    _close(wav, d["wav_c"], atol=1e-3)


def test_anti_alias_activation_matches_reference(golden_dir):
    d = _load(golden_dir, "act1d.pt")
    filt = bigvgan_ref.kaiser_sinc_filter()
    _close(filt, d["up_filter"], atol=1e-7)
    _close(filt, d["down_filter"], atol=1e-7)
    _close(bigvgan_ref.anti_alias_act(d["x"], d["alpha"], d["beta"]), d["y"], atol=2e-6)


def test_bigvgan_oracle_on_the_fork_topology_and_ampblock2(golden_dir):
    """oracle/bigvgan_ref.py pinned on the fork's own topology ([8,5,3,2], 1024-channel input; conv_post gain as recorded) and
    on AMPBlock2, against recordings of the REAL reference trunk (oracle/make_golden.py)."""
    d = _load(golden_dir, "bigvgan_fork.pt")
    sd = init_weights.bigvgan_sd("fork", seed=d["weights_seed"])
    sd["conv_post.weight_g"] = sd["conv_post.weight_g"] * d["post_gain"]
    sd["conv_post.bias"] = sd["conv_post.bias"] * d["post_gain"]
    _close(bigvgan_ref.trunk_forward(sd, d["feat"], d["spk"], "fork"), d["wav"], atol=5e-5)
    d = _load(golden_dir, "bigvgan_24k100_amp2.pt")
    sd = init_weights.bigvgan_sd(d["topology"], seed=d["weights_seed"], resblock="2")
    _close(bigvgan_ref.trunk_forward(sd, d["feat"], d["spk"], d["topology"]), d["wav"], atol=2e-5)


def test_bigvgan_trunk_matches_reference(golden_dir):
    d = _load(golden_dir, "bigvgan_24k100.pt")
    sd = init_weights.bigvgan_sd(d["topology"], seed=d["weights_seed"])
    wav = bigvgan_ref.trunk_forward(sd, d["feat"], d["spk"], d["topology"])
    _close(wav, d["wav"], atol=1e-4)


# ---- streaming glue: oracle/token2wav_ref.py against the real CosyVoiceModel / CosyVoice2Model ----
def test_fade_in_out_and_interp_match_reference(golden_dir):
    import numpy as np
    from oracle import token2wav_ref as T
    d = _load(golden_dir, "token2wav.pt")
    f = d["fade"]
    assert torch.equal(T.fade_in_out_ref(f["a"], f["b"], np.hamming(2 * f["n"])), f["out"])      # bit-exact
    for n, ref in zip(d["interp"]["sizes"], d["interp"]["outs"]):
        assert torch.equal(T.interp_linear_ref(d["interp"]["x"], n), ref)


@pytest.mark.parametrize("ver", ["v1", "v2"])
def test_token2wav_oracle_matches_reference_session(golden_dir, ver):
    from oracle import token2wav_ref as T
    d = _load(golden_dir, "token2wav.pt")[ver]
    if ver == "v1":
        m = T.Token2WavRef(T.StubFlow(), T.StubHift(T.SCENARIO_V1["hop"]), sr=T.SCENARIO_V1["sr"])
        outs, state = T.run_scenario_v1(m)
    else:
        m = T.Token2Wav2Ref(T.StubFlow(input_frame_rate=25), T.StubHift(T.SCENARIO_V2["hop"]))
        outs, state = T.run_scenario_v2(m)
    assert [tuple(c) for c in m.flow.calls] == [tuple(c) for c in d["flow_calls"]]
    for i, (o, ref) in enumerate(zip(outs, d["outs"])):
        assert torch.equal(o, ref), f"chunk {i}"                    # fades are double -> fp32, stubs dyadic: bit-exact
    for k, v in d["state"].items():
        assert torch.equal(state[k], v) if torch.is_tensor(v) else tuple(state[k]) == tuple(v), k


def test_sdpa_attention_of_the_port_equals_the_written_out_form():
    """oracle/flow_ref._attention calls F.scaled_dot_product_attention the way diffusers' AttnProcessor2_0 does (so the CPU
    baseline costs what the reference costs); the explicit softmax(QK^T d^-1/2 + M) V form is the independent cross-check
    SURVEY.md 8c asks for.  Includes a mask with padded keys: the {0,1} float mask is ADDED, not applied as a boolean."""
    import torch
    from oracle import flow_ref, init_weights
    sd = init_weights.flow_estimator_sd(False, seed=0)
    p = "mid_blocks.3.1.2.attn1."
    g = torch.Generator().manual_seed(5)
    x = torch.randn(2, 37, 256, generator=g)
    mask = torch.ones(2, 1, 37)
    mask[1, :, 30:] = 0
    am = torch.matmul(mask.transpose(1, 2), mask)
    a = flow_ref._attention(sd, p, x, am)
    b = flow_ref._attention_explicit(sd, p, x, am)
    assert (a - b).abs().max().item() < 2e-5
    assert (flow_ref._attention(sd, p, x, None) - flow_ref._attention_explicit(sd, p, x, None)).abs().max().item() < 2e-5
