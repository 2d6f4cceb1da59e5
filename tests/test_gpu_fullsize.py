"""Parity at BASELINE.json's FULL sizes (10 s utterances: 75 prompt + 250 generated tokens -> T = 281 + 937 frames, L = 239 872 samples).

The small-case tests compare with the oracle element by element; here the oracle is run once on ONE full utterance
(BASELINE config 1, the reference's own CPU-runnable case: a few seconds on the GPU box's host cores), and the batch-32 case
(config 2) is checked through size-independent properties: utterances are independent (a permuted batch gives bit-identical
permuted results; a batch member equals its B = 1 run), the bf16 engine stays within the reference's own bf16-vs-fp32 gap of
the fp32 engine, waveforms are bounded by the clamp, the 16-point STFT kernel's spectrum inverts back to the signal, and the chunk mask
with a chunk longer than the utterance is no mask.

Tolerances: fp32 engine vs oracle: mel max-abs < 2e-3, waveform (injected source) < 1e-3 -- the bars of the small cases;
bf16 vs fp32 engine: mel rel-rms < 3 %, max-abs < 0.25.
"""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from cosyvoice_b200 import ops
from cosyvoice_b200.frontend import B200CausalMaskedDiffWithXvec, B200MaskedDiffWithXvec
from cosyvoice_b200.hift import B200HiFTGenerator
from cosyvoice_b200.layout import SeqLayout
from oracle import encoder_ref, flow_ref, hift_ref, init_weights

DEV = "cuda"
N_PROMPT, N_TOK, SR = 75, 250, 24000
T1, T2 = 281, 937

_FLOW = {}


def _flow(dtype):
    if dtype not in _FLOW:
        _FLOW[dtype] = B200MaskedDiffWithXvec(init_weights.flow_frontend_sd(False, seed=2), dtype=dtype, device=DEV,
                                              estimator_state_dict=init_weights.flow_estimator_sd(False, seed=0))
    return _FLOW[dtype]


def _inputs(n, first=0):
    toks, ptoks, pfeats, embs, zs = [], [], [], [], []
    for u in range(first, first + n):
        t, pt, pf, e = init_weights.frontend_inputs(False, N_TOK, N_PROMPT, seed=1234 + u, sample_rate=SR)
        toks.append(t), ptoks.append(pt), pfeats.append(pf), embs.append(e)
        zs.append(torch.randn(1, 80, T1 + T2, generator=torch.Generator().manual_seed(4321 + u)))
    return toks, ptoks, pfeats, embs, zs


def _golden(ver):
    """mel recorded from the REAL reference's flow.inference at this size (oracle/make_golden.py fullsize())."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fullsize.pt")
    return torch.load(path, weights_only=True)[ver]["mel"]


def _rel_rms(a, b):
    return ((a - b).pow(2).mean().sqrt() / b.pow(2).mean().sqrt()).item()


def test_config1_one_full_utterance_fp32_matches_the_oracle():
    """flow.inference (front-end + 10-step CFM) and HiFT decode of one 10 s utterance against the CPU oracle."""
    toks, ptoks, pfeats, embs, zs = _inputs(1)
    fsd, esd = init_weights.flow_frontend_sd(False, seed=2), init_weights.flow_estimator_sd(False, seed=0)
    torch.set_flush_denormal(True)
    mu, spks, cond, l1, l2 = encoder_ref.frontend_v1(fsd, toks[0], ptoks[0], pfeats[0], embs[0], sample_rate=SR)
    assert (l1, l2) == (T1, T2)
    want = flow_ref.solve_euler(esd, zs[0], flow_ref.cosine_t_span(10), mu, torch.ones(1, 1, T1 + T2), spks, cond)[:, :, T1:]
    got = _flow(torch.float32).inference_batch(toks, ptoks, pfeats, embs, zs=zs, sample_rate=SR)[0]
    err = (got.cpu() - want).abs().max().item()
    assert (want - _golden("v1")).abs().max().item() < 1e-4         # the oracle itself against the reference's recording
    gerr = (got.cpu() - _golden("v1")).abs().max().item()
    print(f"full-size flow fp32: max-abs {err:.3e} vs the oracle, {gerr:.3e} vs the reference's recorded mel")
    assert got.shape == (1, 80, T2) and err < 2e-3 and gerr < 2e-3, (err, gerr)
    hsd = init_weights.hift_sd(False, seed=1)
    src = torch.randn(1, 1, T2 * 256, generator=torch.Generator().manual_seed(8)) * 0.1
    want_w = hift_ref.decode(hsd, want, src, v2=False)
    hift = B200HiFTGenerator(hsd, v2=False, dtype=torch.float32, device=DEV)
    wav = hift.inference_batch([want], source=src)[0][0].cpu()
    werr = (wav - want_w).abs().max().item()
    print(f"full-size hift fp32: max-abs {werr:.3e}")
    assert wav.shape == (1, T2 * 256) and werr < 1e-3, werr


def test_config3_size_one_full_v2_utterance_fp32_matches_the_oracle():
    """CosyVoice2 at the size of BASELINE's streaming config once finalised (75 prompt + 250 tokens -> T = 150 + 500, L = 240 000):
    CausalMaskedDiffWithXvec.inference (UpsampleConformer front-end + causal 10-step CFM) and the v2 HiFT decode against the oracle."""
    tok, ptok, pfeat, emb = init_weights.frontend_inputs(True, N_TOK, N_PROMPT, seed=1234, sample_rate=SR)
    fsd, esd = init_weights.flow_frontend_sd(True, seed=2), init_weights.flow_estimator_sd(True, seed=0)
    torch.set_flush_denormal(True)
    mu, spks, cond, l1, l2 = encoder_ref.frontend_v2(fsd, tok, ptok, pfeat, emb, static_chunk_size=0)
    assert (l1, l2) == (150, 500)
    flow = B200CausalMaskedDiffWithXvec(fsd, dtype=torch.float32, device=DEV, estimator_state_dict=esd)
    # the causal CFM's noise is a slice of its fixed seed-0 buffer (flow_matching.py:212-213, :235): same values as the reference's
    z = torch.randn([1, 80, 50 * 300], generator=torch.Generator().manual_seed(0))[:, :, :l1 + l2]
    assert torch.equal(z, flow.decoder.rand_noise[:, :, :l1 + l2])
    want = flow_ref.solve_euler(esd, z, flow_ref.cosine_t_span(10), mu, torch.ones(1, 1, l1 + l2), spks, cond, causal=True)[:, :, l1:]
    i32 = lambda n: torch.tensor([n], dtype=torch.int32)
    got, _ = flow.inference(token=tok.to(DEV), token_len=i32(tok.shape[1]), prompt_token=ptok.to(DEV),
                            prompt_token_len=i32(ptok.shape[1]), prompt_feat=pfeat.to(DEV), prompt_feat_len=i32(pfeat.shape[1]),
                            embedding=emb.to(DEV), n_timesteps=10, finalize=True)
    err = (got.cpu() - want).abs().max().item()
    assert (want - _golden("v2")).abs().max().item() < 1e-4         # the oracle itself against the reference's recording
    gerr = (got.cpu() - _golden("v2")).abs().max().item()
    print(f"full-size v2 flow fp32: max-abs {err:.3e} vs the oracle, {gerr:.3e} vs the reference's recorded mel")
    assert got.shape == (1, 80, 500) and err < 2e-3 and gerr < 2e-3, (err, gerr)
    hsd = init_weights.hift_sd(True, seed=1)
    src = torch.randn(1, 1, 500 * 480, generator=torch.Generator().manual_seed(8)) * 0.1
    want_w = hift_ref.decode(hsd, want, src, v2=True)
    hift = B200HiFTGenerator(hsd, v2=True, dtype=torch.float32, device=DEV)
    wav = hift.inference_batch([want], source=src)[0][0].cpu()
    werr = (wav - want_w).abs().max().item()
    print(f"full-size v2 hift fp32: max-abs {werr:.3e}")
    assert wav.shape == (1, 500 * 480) and werr < 1e-3, werr


def test_config2_batch_members_are_independent_and_bf16_is_within_the_gap():
    B = 32
    toks, ptoks, pfeats, embs, zs = _inputs(B)
    flow = _flow(torch.bfloat16)
    mels = [m.clone() for m in flow.inference_batch(toks, ptoks, pfeats, embs, zs=zs, sample_rate=SR)]
    assert all(m.shape == (1, 80, T2) and torch.isfinite(m).all() for m in mels)
    # permutation: utterances never mix (flow/flow.py:125 asserts B = 1: the batch must behave as 32 separate calls)
    perm = torch.randperm(B, generator=torch.Generator().manual_seed(0)).tolist()
    pm = flow.inference_batch([toks[i] for i in perm], [ptoks[i] for i in perm], [pfeats[i] for i in perm],
                              [embs[i] for i in perm], zs=[zs[i] for i in perm], sample_rate=SR)
    for k, i in enumerate(perm):
        assert torch.equal(pm[k], mels[i]), (k, i)
    # a member of the batch equals its own B = 1 run bit for bit (same tiles, same kernels)
    one = flow.inference_batch(toks[5:6], ptoks[5:6], pfeats[5:6], embs[5:6], zs=zs[5:6], sample_rate=SR)[0]
    assert torch.equal(one, mels[5])
    # bf16 engine against the fp32 engine at full size
    ref32 = _flow(torch.float32).inference_batch(toks[:2], ptoks[:2], pfeats[:2], embs[:2], zs=zs[:2], sample_rate=SR)
    for b in range(2):
        rel, mx = _rel_rms(mels[b], ref32[b]), (mels[b] - ref32[b]).abs().max().item()
        print(f"full-size bf16 vs fp32 engine, utterance {b}: rel-rms {rel:.3e} max-abs {mx:.3e}")
        assert rel < 0.03 and mx < 0.25, (b, rel, mx)
    # vocoder on the whole batch: bounded by the clamp (generator.py:381), right length, independent of batch position
    hift = B200HiFTGenerator(init_weights.hift_sd(False, seed=1), v2=False, dtype=torch.bfloat16, device=DEV)
    g = torch.Generator().manual_seed(3)
    pv = torch.rand(B, 9, 1, generator=g) * 6.28 - 3.14
    pv[:, 0] = 0
    noise = torch.randn(B, 9, T2 * 256, generator=g)
    wavs, _ = hift.inference_batch(mels, phase_vec=pv, noise=noise)
    assert all(w.shape == (1, T2 * 256) and torch.isfinite(w).all() and w.abs().max() <= 0.99 + 1e-6 for w in wavs)
    w5, _ = hift.inference_batch(mels[5:6], phase_vec=pv[5:6], noise=noise[5:6])
    assert torch.equal(w5[0], wavs[5])


def test_stft_istft_identity_and_unmasked_chunk_at_full_length():
    # STFT(16, hop 4, hann, center/reflect) of generator.py:333-339 at L = 239 872
    L = T2 * 256
    F = L // 4 + 1
    lay = SeqLayout.build([F, F], DEV)
    s = torch.randn(2, L, generator=torch.Generator().manual_seed(1)).to(DEV) * 0.3
    spec = torch.zeros(lay.rows, 64, device=DEV)
    ops.stft16(s, spec, lay)
    re, im = spec[:, 0:9], spec[:, 9:18]
    # the spectrum our kernel produced inverts back to the signal through torch's own iSTFT
    spec_c = torch.complex(re, im)
    for b in range(2):
        r0 = lay.row0[b]
        back = torch.istft(spec_c[r0:r0 + F].t().contiguous()[None], 16, 4, 16, window=torch.hann_window(16, device=DEV), center=True)
        assert (back[0] - s[b]).abs().max().item() < 1e-4
    # rel-pos attention: a chunk mask at least as long as the sequence is no mask (utils/mask.py:122)
    g = torch.Generator().manual_seed(2)
    T = N_PROMPT + N_TOK
    la = SeqLayout.build([T], DEV, halo=8)
    qkv = torch.zeros(la.rows, 1536, device=DEV)
    qkv[:T] = torch.randn(T, 1536, generator=g).to(DEV)
    P = (torch.randn(2 * 512 - 1, 512, generator=g) * 0.3).to(DEV)
    u = (torch.randn(512, generator=g) * 0.2).to(DEV)
    outs = []
    for chunk in (0, T, 4096):
        o = torch.zeros(la.rows, 512, device=DEV)
        ops.relpos_attention_op(qkv, P, 511, u, u, o, la, 8, 0.125, chunk)()
        outs.append(o)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
