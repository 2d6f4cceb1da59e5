"""GPU parity of the flow front-end (cosyvoice_b200/frontend.py + csrc/encoder.cu) through the C ABI, against
oracle/encoder_ref.py and the fixtures recorded from the REAL flow.inference (tests/golden/frontend_v{1,2}.pt).

Tolerances: fp32 engine: mu max-abs < 1e-3 (values of rms ~0.5; fp32 re-association over 6-10 transformer layers),
kernels alone < 2e-5.  bf16 engine: mu rel-rms < 2 %, max-abs < 0.1.  Whole boundary call (front-end + 10-step CFM):
the CFM bars of tests/test_gpu_flow.py (fp32 2e-3; bf16 rel-rms 3 %, max-abs 0.25)."""
import math
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from cosyvoice_b200 import _lib, ops
from cosyvoice_b200.frontend import B200CausalMaskedDiffWithXvec, B200MaskedDiffWithXvec, rel_pos_table
from cosyvoice_b200.layout import SeqLayout
from oracle import encoder_ref as E
from oracle import init_weights

DEV = "cuda"


def _load(golden_dir, name):
    return torch.load(os.path.join(golden_dir, name), weights_only=True)


def _rows(lay, xs, width=None, dtype=torch.float32):
    """list of [len_b, C] -> [rows, width] following the layout"""
    C = xs[0].shape[1]
    buf = torch.zeros(lay.rows, width or C, dtype=dtype)
    for b, x in enumerate(xs):
        buf[lay.row0[b]:lay.row0[b] + x.shape[0], :C] = x
    return buf.to(DEV)


def _unrows(lay, buf, C=None):
    buf = buf.float().cpu()
    return [buf[lay.row0[b]:lay.row0[b] + lay.lens[b], :C or buf.shape[1]] for b in range(lay.n_seq)]


def _attn_ref(q, k, v, P_of_r, u, vb, chunk):
    """Definition of RelPositionMultiHeadedAttention's core for one sequence (double precision)."""
    T = q.shape[0]
    q, k, v = (a.double().view(T, 8, 64) for a in (q, k, v))
    out = torch.zeros(T, 8, 64, dtype=torch.float64)
    for h in range(8):
        s = torch.zeros(T, T, dtype=torch.float64)
        for i in range(T):
            for j in range(T):
                s[i, j] = ((q[i, h] + u[h].double()) @ k[j, h] + (q[i, h] + vb[h].double()) @ P_of_r(i - j)[h].double()) / 8.0
        m = E.chunk_mask(T, chunk)
        if m is not None:
            s = s.masked_fill(~m, -float("inf"))
        out[:, h] = torch.softmax(s, dim=-1) @ v[:, h]
    return out.view(T, 512).float()


@pytest.mark.parametrize("chunk", [0, 5])
def test_relpos_attention_matches_definition(chunk):
    g = torch.Generator().manual_seed(3)
    lens = [37, 70, 1]
    lay = SeqLayout.build(lens, DEV, halo=4)
    half = 128
    P = torch.randn(2 * half - 1, 512, generator=g) * 0.5
    u, vb = torch.randn(8, 64, generator=g) * 0.3, torch.randn(8, 64, generator=g) * 0.3
    qkvs = [torch.randn(n, 1536, generator=g) for n in lens]
    qkv = _rows(lay, qkvs)
    out = torch.full((lay.rows, 512), 7.0, device=DEV)
    ops.relpos_attention_op(qkv, P.to(DEV), half - 1, u.reshape(-1).to(DEV), vb.reshape(-1).to(DEV), out, lay, 8, 0.125, chunk)()
    got = _unrows(lay, out)
    for b, x in enumerate(qkvs):
        want = _attn_ref(x[:, :512], x[:, 512:1024], x[:, 1024:], lambda r: P[half - 1 - r].view(8, 64), u, vb, chunk)
        assert (got[b] - want).abs().max() < 2e-5, b
    full = out.float().cpu()
    for b in range(len(lens)):          # rows past the valid length inside the touched tiles are written as zeros
        assert full[lay.row0[b] + lens[b]:lay.row0[b] + ((lens[b] + 31) // 32) * 32].abs().max() == 0


def test_relpos_attention_rejects_short_table():
    lay = SeqLayout.build([200], DEV)
    qkv = torch.zeros(lay.rows, 1536, device=DEV)
    P = torch.zeros(2 * 64 - 1, 512, device=DEV)
    b = torch.zeros(512, device=DEV)
    with pytest.raises(RuntimeError):
        ops.relpos_attention_op(qkv, P, 63, b, b, torch.zeros(lay.rows, 512, device=DEV), lay, 8, 0.125)()


@pytest.mark.parametrize("C,eps", [(512, 1e-12), (512, 1e-5), (256, 1e-5)])
def test_layernorm_seq(C, eps):
    g = torch.Generator().manual_seed(4)
    lens = [5, 130]
    lay = SeqLayout.build(lens, DEV)
    xs = [torch.randn(n, C, generator=g) * 3 + 1 for n in lens]
    gamma, beta = torch.randn(C, generator=g), torch.randn(C, generator=g)
    for dt, tol in ((torch.float32, 2e-5), (torch.bfloat16, 4e-2)):
        out = torch.full((lay.rows, C), 3.0, dtype=dt, device=DEV)
        ops.layernorm_seq_op(_rows(lay, xs), out, gamma.to(DEV), beta.to(DEV), lay, eps)()
        for got, x in zip(_unrows(lay, out), xs):
            assert (got - torch.nn.functional.layer_norm(x, (C,), gamma, beta, eps)).abs().max() < tol
        assert out.float().cpu()[lens[0]:lay.row0[1]].abs().max() == 0      # pad rows are zeros, not beta


def test_groupnorm1_mish_and_interp_rows_and_embed():
    g = torch.Generator().manual_seed(5)
    F = torch.nn.functional
    lens = [9, 300]
    lay = SeqLayout.build(lens, DEV, align=256)
    xs = [torch.randn(n, 80, generator=g) * 2 + 0.5 for n in lens]
    gamma, beta = torch.randn(80, generator=g), torch.randn(80, generator=g)
    out = torch.full((lay.rows, 128), 5.0, device=DEV)
    ops.groupnorm1_mish_op(_rows(lay, xs, 128), out, gamma.to(DEV), beta.to(DEV), lay)()
    for got, x in zip(_unrows(lay, out, 80), xs):
        want = F.mish(F.group_norm(x.t().unsqueeze(0), 1, gamma, beta, 1e-5)).squeeze(0).t()
        assert (got - want).abs().max() < 2e-5
    assert out.cpu()[lens[0]:lay.row0[1], :80].abs().max() == 0
    # piecewise linear interpolation: bit-exact against F.interpolate (same arithmetic as ATen)
    src_lens, n1 = [62, 30], [12, 0]
    mel = [(int(12 / 25 * 22050 / 256), int(50 / 25 * 22050 / 256)), (0, int(30 / 25 * 22050 / 256))]
    src = SeqLayout.build(src_lens, DEV)
    dst = SeqLayout.build([a + b for a, b in mel], DEV, align=256)
    hs = [torch.randn(n, 80, generator=g) for n in src_lens]
    segs = torch.zeros(2, 4, 4, dtype=torch.int32)
    for b in range(2):
        p = E.regulator_segments(n1[b], src_lens[b] - n1[b], mel[b][0], mel[b][1])
        segs[b, :len(p)] = torch.tensor(p, dtype=torch.int32)
    out = torch.full((dst.rows, 128), 5.0, device=DEV)
    ops.interp_rows_op(_rows(src, hs, 128), src, out, dst, segs.to(DEV), 80)()
    for b, got in enumerate(_unrows(dst, out, 80)):
        parts = [F.interpolate(hs[b][s:s + n].t().unsqueeze(0), size=dn, mode="linear")
                 for (s, n, _, dn) in E.regulator_segments(n1[b], src_lens[b] - n1[b], mel[b][0], mel[b][1])]
        assert torch.equal(got, torch.cat(parts, 2).squeeze(0).t().contiguous()), b
    # embedding gather (negative ids clamp to 0 like torch.clamp(token, min=0))
    table = torch.randn(50, 512, generator=g)
    tl = SeqLayout.build([4, 3], DEV)
    toks = torch.tensor([[3, -2, 49, 7], [1, 0, 5, 0]], dtype=torch.int32)
    out = torch.full((tl.rows, 512), 5.0, device=DEV)
    ops.embed_tokens_op(toks.to(DEV), table.to(DEV), out, tl)()
    got = _unrows(tl, out)
    assert torch.equal(got[0], table[[3, 0, 49, 7]]) and torch.equal(got[1], table[[1, 0, 5]])
    assert out.cpu()[4:tl.row0[1]].abs().max() == 0
    x = torch.randn(3, 192, generator=g)
    y = torch.empty(3, 192, device=DEV)
    ops.l2_normalize(x.to(DEV), y)
    assert (y.cpu() - F.normalize(x, dim=1)).abs().max() < 1e-6


def _flow(v2, dtype, golden):
    sd = dict(init_weights.flow_frontend_sd(v2, seed=golden["weights_seed"]))
    est = init_weights.flow_estimator_sd(v2, seed=golden["estimator_seed"])
    cls = B200CausalMaskedDiffWithXvec if v2 else B200MaskedDiffWithXvec
    return cls(sd, dtype=dtype, device=DEV, estimator_state_dict=est)


def _call(flow, c, v2):
    i32 = lambda n: torch.tensor([n], dtype=torch.int32)
    kw = dict(token=c["token"].to(DEV), token_len=i32(c["token"].shape[1]), prompt_token=c["prompt_token"].to(DEV),
              prompt_token_len=i32(c["prompt_token"].shape[1]), prompt_feat=c["prompt_feat"].to(DEV),
              prompt_feat_len=i32(c["prompt_feat"].shape[1]), embedding=c["embedding"].to(DEV))
    if v2:
        flow.encoder.static_chunk_size = c["chunk"]
        return flow.inference(**kw, n_timesteps=10)
    return flow.inference(**kw, flow_cache=torch.zeros(1, 80, 0, 2), sample_rate=22050, n_timesteps=10)


@pytest.mark.parametrize("ver", ["v1", "v2"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_flow_inference_matches_reference(golden_dir, ver, dtype):
    """The outer boundary call against the REAL reference's recorded outputs: front-end tensors and the final mel."""
    d = _load(golden_dir, f"frontend_{ver}.pt")
    v2 = d["v2"]
    flow = _flow(v2, dtype, d)
    seen = {}
    cfm = flow.decoder

    class Spy:                              # stands where flow.decoder stands: records what flow.inference hands to the CFM
        def __getattr__(self, name):
            return getattr(cfm, name)

        def __call__(self, mu, mask, n_timesteps, temperature=1.0, spks=None, cond=None, **kw):
            seen.update(mu=mu.clone(), spks=spks.clone(), cond=cond.clone())
            return cfm(mu, mask, n_timesteps, temperature, spks=spks, cond=cond, **kw)
    flow.decoder = Spy()
    for i, c in enumerate(d["cases"]):
        cfm._noise = lambda mu, _z=c["z"]: _z.to(mu.device)
        mel, cache = _call(flow, c, v2)
        mu, spks, cond = seen["mu"].cpu(), seen["spks"].cpu(), seen["cond"].cpu()
        assert torch.equal(cond, c["cond"]), i
        assert (spks - c["spks"]).abs().max() < 1e-5, i
        err = (mu - c["mu"]).abs().max().item()
        rel = ((mu - c["mu"]).pow(2).mean().sqrt() / c["mu"].pow(2).mean().sqrt()).item()
        merr = (mel.cpu() - c["mel"]).abs().max().item()
        mrel = ((mel.cpu() - c["mel"]).pow(2).mean().sqrt() / c["mel"].pow(2).mean().sqrt()).item()
        print(f"{ver} {dtype} case {i}: mu max-abs {err:.3e} rel-rms {rel:.3e}; mel max-abs {merr:.3e} rel-rms {mrel:.3e}")
        assert mel.shape == c["mel"].shape and mel.is_cuda
        if dtype == torch.float32:
            assert err < 1e-3 and merr < 2e-3, (i, err, merr)
        else:
            assert rel < 0.02 and err < 0.1, (i, err, rel)
            assert mrel < 0.03 and merr < 0.25, (i, merr, mrel)
        if not v2:
            assert cache.shape == c["flow_cache"].shape
            if dtype == torch.float32:
                assert (cache.cpu() - c["flow_cache"]).abs().max() < 1e-3


@pytest.mark.parametrize("ver", ["v1", "v2"])
def test_inference_batch_equals_per_utterance_calls(golden_dir, ver):
    """Ragged batch through front-end + CFM == the B = 1 boundary calls (fp32 engine, same injected noise)."""
    d = _load(golden_dir, f"frontend_{ver}.pt")
    v2 = d["v2"]
    flow = _flow(v2, torch.float32, d)
    cases = [c for c in d["cases"] if not v2 or c["chunk"] == 0]
    pf = []
    for c in cases:
        p = c["prompt_feat"]
        pf.append(p[:, :p.shape[1] - p.shape[1] % 2] if v2 else p)
    pt = [c["prompt_token"][:, :pf[i].shape[1] // 2] if v2 else c["prompt_token"] for i, c in enumerate(cases)]
    flow.encoder.static_chunk_size = 0
    mels = flow.inference_batch([c["token"] for c in cases], pt, pf, [c["embedding"] for c in cases],
                                zs=[c["z"] for c in cases], sample_rate=22050)
    for i, (m, c) in enumerate(zip(mels, cases)):
        err = (m.cpu() - c["mel"]).abs().max().item()
        assert m.shape == c["mel"].shape and err < 2e-3, (i, err)


def test_frontend_needs_cuda_and_b1():
    d = dict(init_weights.flow_frontend_sd(False, seed=2))
    with pytest.raises(RuntimeError):
        B200MaskedDiffWithXvec(d, device="cpu", estimator_state_dict=init_weights.flow_estimator_sd(False))


def test_cosyvoice2_session_streams_through_the_whole_cuda_back_end(golden_dir):
    """CosyVoice2Model.token2wav's bookkeeping (cli/model.py:334-366) over the REAL B200 flow (front-end + CFM, chunk-masked
    encoder as cli/model.py:314 sets it) and the B200 HiFT: chunk lengths, caches on the device, finite bounded audio;
    the last (finalize) chunk's mel equals a one-shot flow.inference of the same tokens."""
    from cosyvoice_b200.hift import B200HiFTGenerator
    from cosyvoice_b200.token2wav import Token2Wav2
    from oracle import token2wav_ref as T
    d = _load(golden_dir, "frontend_v2.pt")
    flow = _flow(True, torch.bfloat16, d)
    hift = B200HiFTGenerator(init_weights.hift_sd(True, seed=1), v2=True, dtype=torch.bfloat16, device=DEV)
    m = Token2Wav2(flow, hift, device=DEV)
    flow.encoder.static_chunk_size = 2 * flow.input_frame_rate                     # model.py:314
    m.open_session("u")
    g = torch.Generator().manual_seed(11)
    tok = torch.randint(0, 6561, (1, 70), generator=g, dtype=torch.int32)
    pt = torch.randint(0, 6561, (1, 9), generator=g, dtype=torch.int32)
    pf, emb = torch.randn(1, 18, 80, generator=g) * 2 - 4, torch.randn(1, 192, generator=g)
    total = 0
    for n, off, fin in T.SCENARIO_V2["steps"]:
        wav = m.token2wav(tok[:, :n], pt, pf, emb, "u", off, finalize=fin)
        assert wav.is_cuda and torch.isfinite(wav).all() and wav.abs().max() <= 0.99 + 1e-6
        cached = 0 if off == 0 else 8
        assert wav.shape[1] == ((n - off) * 2 + cached) * 480 - (0 if fin else 3840)
        total += wav.shape[1]
    assert total == 70 * 2 * 480
    m.close_session("u")
    i32 = lambda k: torch.tensor([k], dtype=torch.int32)
    mel, _ = flow.inference(token=tok.to(DEV), token_len=i32(70), prompt_token=pt.to(DEV), prompt_token_len=i32(9),
                            prompt_feat=pf.to(DEV), prompt_feat_len=i32(18), embedding=emb.to(DEV))
    assert mel.shape == (1, 80, 140) and torch.isfinite(mel).all()
