"""GPU parity of the streaming glue (cosyvoice_b200/token2wav.py) against the REAL CosyVoiceModel / CosyVoice2Model
sessions recorded in tests/golden/token2wav.pt and against the CPU oracle on fresh inputs.

Tolerance: bit-exact.  The cross-fades are double arithmetic rounded to fp32 once, the speed interpolation repeats
ATen's FMA placement, and the stub flow / stub vocoder only use dyadic constants.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from cosyvoice_b200 import ops
from cosyvoice_b200.hift import B200HiFTGenerator
from cosyvoice_b200.token2wav import Token2Wav, Token2Wav2
from oracle import init_weights
from oracle import token2wav_ref as T

DEV = "cuda"


def _load(golden_dir):
    return torch.load(os.path.join(golden_dir, "token2wav.pt"), weights_only=True)


def test_fade_in_out_kernel_is_bit_exact(golden_dir):
    f = _load(golden_dir)["fade"]
    w = torch.from_numpy(np.hamming(2 * f["n"])).to(DEV)
    a = f["a"].to(DEV).contiguous()
    out = ops.fade_in_out(a.clone(), f["b"].to(DEV).contiguous(), w)
    assert torch.equal(out.cpu(), f["out"])
    # speech-shaped: one row, overlap 5120, against the oracle
    g = torch.Generator().manual_seed(5)
    x, p = torch.randn(1, 9000, generator=g), torch.randn(1, 5120, generator=g)
    w2 = np.hamming(2 * 5120)
    out = ops.fade_in_out(x.to(DEV), p.to(DEV), torch.from_numpy(w2).to(DEV))
    assert torch.equal(out.cpu(), T.fade_in_out_ref(x, p, w2))


def test_fade_in_out_rejects_short_chunks():
    w = torch.from_numpy(np.hamming(68)).to(DEV)
    with pytest.raises(RuntimeError):          # the reference's slice assignment raises on the same inputs
        ops.fade_in_out(torch.zeros(1, 80, 20, device=DEV), torch.zeros(1, 80, 34, device=DEV), w)
    with pytest.raises(RuntimeError):
        ops.fade_in_out(torch.zeros(1, 80, 40, device=DEV), torch.zeros(1, 80, 10, device=DEV), w)
    with pytest.raises(RuntimeError):          # no CPU path
        ops.fade_in_out(torch.zeros(1, 80, 40), torch.zeros(1, 80, 34), w)


def test_interp_linear_kernel_is_bit_exact(golden_dir):
    d = _load(golden_dir)["interp"]
    for n, ref in zip(d["sizes"], d["outs"]):
        assert torch.equal(ops.interp_linear(d["x"].to(DEV), n).cpu(), ref), n


@pytest.mark.parametrize("ver", ["v1", "v2"])
def test_token2wav_session_matches_reference(golden_dir, ver):
    d = _load(golden_dir)[ver]
    if ver == "v1":
        m = Token2Wav(T.StubFlow(), T.StubHift(T.SCENARIO_V1["hop"]), sr=T.SCENARIO_V1["sr"], device=DEV)
        outs, state = T.run_scenario_v1(m, DEV)
    else:
        m = Token2Wav2(T.StubFlow(input_frame_rate=25), T.StubHift(T.SCENARIO_V2["hop"]), device=DEV)
        outs, state = T.run_scenario_v2(m, DEV)
    assert [tuple(c) for c in m.flow.calls] == [tuple(c) for c in d["flow_calls"]]
    for i, (o, ref) in enumerate(zip(outs, d["outs"])):
        assert o.is_cuda and torch.equal(o.cpu(), ref), f"chunk {i}"
    for k, v in d["state"].items():
        if torch.is_tensor(v):
            assert state[k].is_cuda and torch.equal(state[k].cpu(), v), k       # caches stay in HBM
        else:
            assert tuple(state[k]) == tuple(v), k


def test_token2wav_v2_streams_through_the_cuda_vocoder():
    """The same session with the real B200 HiFT behind it: chunk lengths follow the reference's bookkeeping
    (mel_cache_len * 480 samples held back per chunk) and the cached source is re-used (generator.py:408-409)."""
    hift = B200HiFTGenerator(init_weights.hift_sd(True, seed=1), v2=True, dtype=torch.bfloat16, device=DEV)
    m = Token2Wav2(T.StubFlow(input_frame_rate=25), hift, device=DEV)
    m.open_session("u")
    tok = T.scenario_tokens(70, 3, DEV)
    pt, pf, emb = T.scenario_tokens(9, 4, DEV), torch.zeros(1, 18, 80, device=DEV), torch.zeros(1, 192, device=DEV)
    total = 0
    for n, off, fin in T.SCENARIO_V2["steps"]:
        wav = m.token2wav(tok[:, :n], pt, pf, emb, "u", off, finalize=fin)
        assert wav.is_cuda and torch.isfinite(wav).all() and wav.abs().max() <= 0.99 + 1e-6
        new_frames = (n - off) * 2
        cached = 0 if off == 0 else 8
        assert wav.shape[1] == (new_frames + cached) * 480 - (0 if fin else 3840)
        total += wav.shape[1]
        if not fin:
            c = m.hift_cache_dict["u"]
            assert c["mel"].shape == (1, 80, 8) and c["source"].shape == (1, 1, 3840) and c["speech"].shape == (1, 3840)
    assert total == 70 * 2 * 480
    m.close_session("u")
    assert "u" not in m.hift_cache_dict
