"""Parity at BASELINE.json's FULL sizes (10 s utterances: 75 prompt + 250 generated tokens -> T = 281 + 937 frames, L = 239 872 samples).

The small-case tests compare with the oracle element by element; here the oracle is run once on ONE full utterance
(BASELINE config 1, the reference's own CPU-runnable case: a few seconds on the GPU box's host cores), and the batch-32 case
(config 2) is checked through size-independent properties: utterances are independent (a permuted batch gives bit-identical
permuted results; a batch member equals the same utterance in a smaller batch), the bf16 engine stays within the reference's own bf16-vs-fp32 gap of
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
    # a member of the batch equals the same utterance in a smaller batch bit for bit (same tiles, same kernels: the fused
    # FFN kernel is selected from 48 row tiles up, which an 8-utterance batch reaches at both resolutions) ...
    sub = flow.inference_batch(toks[4:12], ptoks[4:12], pfeats[4:12], embs[4:12], zs=zs[4:12], sample_rate=SR)
    assert torch.equal(sub[1], mels[5])
    # ... and its own B = 1 run, which takes the four-launch FFN of small layouts (flow.FUSED_FFN_MIN_TILES), to bf16
    # rounding of two algebraically equal evaluation orders
    one = flow.inference_batch(toks[5:6], ptoks[5:6], pfeats[5:6], embs[5:6], zs=zs[5:6], sample_rate=SR)[0]
    rel1 = _rel_rms(one, mels[5])
    print(f"B = 1 (unfused FFN) vs batch member (fused FFN): rel-rms {rel1:.3e}")
    assert rel1 < 0.02, rel1          # the size of the bf16-vs-fp32 gap itself (0.9 %), amplified over 10 Euler steps
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


# ----------------------------------------------------------------------------------------------- config 4 (BigVGAN) at its size
def _snr_db(x, ref):
    return (10 * torch.log10(ref.pow(2).mean() / (x - ref).pow(2).mean().clamp_min(1e-30))).item()


def test_config4_bigvgan_batch64_of_94_frames_matches_the_reference_recordings():
    """BASELINE config 4 at its own size: the 24 kHz / 100-band BigVGAN trunk on 64 mels of 94 frames (1 s each), inputs
    drawn exactly as tools/bench_configs.py `bigvgan` draws them (torch.Generator seed 5).  Three members were run one by
    one through the REAL reference trunk (oracle/make_golden.py bigvgan_config4 -> tests/golden/bigvgan_config4.pt):
    fp32 engine max-abs < 2e-4 against them (the small-case bar); a permuted batch returns bit-identical permuted waveforms;
    the bf16 (tcgen05) trunk keeps an SNR > 30 dB (and max-abs < 0.1 on the +-1 tanh output) against the reference's fp32
    waveform.  The reference never runs this vocoder in half precision, so the bound is ours, stated as a signal-to-noise
    ratio: bf16 operands (2^-9 relative rounding, fp32 accumulation) through the trunk's ~75 convolutions measure 36 dB."""
    from cosyvoice_b200.bigvgan import B200BigVGAN
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bigvgan_config4.pt")
    d = torch.load(path, weights_only=True)
    B, frames = d["batch"], d["frames"]
    g = torch.Generator().manual_seed(d["input_seed"])
    feat = torch.randn(B, 100, frames, generator=g) * 2 - 4
    spk = torch.randn(B, 512, 1, generator=g)
    sd = init_weights.bigvgan_sd(d["topology"], seed=d["weights_seed"])
    g32 = B200BigVGAN(sd, d["topology"], dtype=torch.float32, device=DEV)
    wav32 = g32.inference(feat, spk).cpu()
    assert wav32.shape == (B, frames * 256)
    for i, want in d["wavs"].items():
        err = (wav32[i] - want).abs().max().item()
        print(f"config 4 fp32 member {i}: max-abs {err:.3e}")
        assert err < 2e-4, (i, err)
    del g32
    gbf = B200BigVGAN(sd, d["topology"], dtype=torch.bfloat16, device=DEV)
    wavb = gbf.inference(feat, spk).clone()
    perm = torch.randperm(B, generator=torch.Generator().manual_seed(1)).tolist()
    wavp = gbf.inference(feat[perm], spk[perm])
    for k, i in enumerate(perm):
        assert torch.equal(wavp[k], wavb[i]), (k, i)
    for i, want in d["wavs"].items():
        err, snr = (wavb[i].cpu() - want).abs().max().item(), _snr_db(wavb[i].cpu(), want)
        print(f"config 4 bf16 member {i}: max-abs {err:.3e}, SNR {snr:.1f} dB, peak {want.abs().max().item():.3f}")
        assert err < 0.1 and snr > 30.0, (i, err, snr)


# ----------------------------------------------------------------------------------------------- config 5 (ragged sweep) subset
def test_config5_ragged_subset_members_equal_their_b1_runs_and_the_oracle(monkeypatch):
    """BASELINE config 5 on a 16-utterance subset drawn with the sweep's own seed (tools/bench_configs.py `sweep`: generated
    duration ~ U[2, 30] s after a 3 s prompt, tokens = round(25 d)), batched exactly as the sweep batches them
    (shard.make_batches: one stride bucket per batch) through AcousticBackend.synthesize_tokens_batch on the bf16 engine:
      * the subset spans at least three stride buckets and the programs are re-bound across batches;
      * every member's waveform equals its own B = 1 run bit for bit (the reference asserts B = 1, flow/flow.py:125) with the
        FFN kernel choice pinned (CVB200_FUSED_FFN=1; by default small layouts take the four-launch FFN and the two agree
        to bf16 rounding, checked on the first batch: mel rel-rms < 1 %);
    and on the fp32 engine the shortest member (10 Euler steps) and the longest member (T ~ 3000 frames; 2 Euler steps to keep the
    CPU oracle to seconds -- every step is the same computation) match the CPU oracle: mel max-abs < 2e-3, waveform with the
    oracle's mel and an injected source < 1e-3."""
    from cosyvoice_b200 import shard
    from cosyvoice_b200.pipeline import AcousticBackend
    HOP, PROMPT, N = 256, 281, 16
    g = torch.Generator().manual_seed(2048)
    dur = 2.0 + 28.0 * torch.rand(2048, generator=g)
    n_tok = [int(round(25.0 * float(x))) for x in dur[:N]]
    gen_frames = [int(n / 25 * SR / 256) for n in n_tok]
    tot = [PROMPT + f for f in gen_frames]
    batches = shard.make_batches(list(range(N)), tot, row_budget=48 * 1024, align=256, max_batch=64)
    strides = {(tot[b[0]] + 8 + 255) // 256 * 256 for b in batches}
    assert len(strides) >= 3 and sum(len(b) for b in batches) == N, (strides, batches)

    def inputs(u):
        t, pt, pf, e = init_weights.frontend_inputs(False, n_tok[u], 75, seed=1234 + u, sample_rate=SR)
        return t, pt, pf, e, torch.randn(1, 80, tot[u], generator=torch.Generator().manual_seed(4321 + u))
    fsd, esd, hsd = (init_weights.flow_frontend_sd(False, seed=2), init_weights.flow_estimator_sd(False, seed=0),
                     init_weights.hift_sd(False, seed=1))
    be_auto = AcousticBackend(esd, hsd, v2=False, dtype=torch.bfloat16, device=DEV, frontend_sd=fsd)
    cols = list(zip(*[inputs(u) for u in batches[0]]))
    auto_b = [m.clone() for m in be_auto.flow.inference_batch(*[list(c) for c in cols], 10, SR)]
    auto_1 = be_auto.flow.inference_batch(*[[c[0]] for c in cols], 10, SR)[0]
    rel = _rel_rms(auto_1, auto_b[0])
    print(f"config 5 member {batches[0][0]}: B = 1 vs batch member under size-based kernel selection, mel rel-rms {rel:.3e}")
    assert rel < 0.01, rel
    del be_auto
    monkeypatch.setenv("CVB200_FUSED_FFN", "1")
    be = AcousticBackend(esd, hsd, v2=False, dtype=torch.bfloat16, device=DEV, frontend_sd=fsd)
    wavs = {}
    for b in batches:
        cols = list(zip(*[inputs(u) for u in b]))
        torch.manual_seed(99)                       # HiFT's own source RNG (phase / noise draws): same seed for the B = 1 runs
        out = be.synthesize_tokens_batch(*[list(c) for c in cols], sample_rate=SR)
        for u, w in zip(b, out):
            assert w.shape == (1, gen_frames[u] * HOP) and torch.isfinite(w).all() and w.abs().max().item() <= 0.99 + 1e-6
            wavs[u] = w.clone()
    # B = 1: the flow is deterministic given z; HiFT draws its source noise from the global generator, whose stream
    # depends on the batch composition, so the bit-for-bit comparison is made on the MEL and on the vocoder with an injected source
    for b in batches:
        cols = list(zip(*[inputs(u) for u in b]))
        mels = [m.clone() for m in be.flow.inference_batch(*[list(c) for c in cols], 10, SR)]
        L_max = max(gen_frames[u] for u in b) * HOP
        src = torch.zeros(len(b), 1, L_max)
        for k, u in enumerate(b):
            src[k, :, :gen_frames[u] * HOP] = torch.randn(1, gen_frames[u] * HOP, generator=torch.Generator().manual_seed(7000 + u)) * 0.1
        wb = [w.clone() for w in be.hift.inference_batch(mels, source=src, uniform_stride=True)[0]]
        for k, u in enumerate(b):
            one = be.flow.inference_batch(*[[c[k]] for c in cols], 10, SR)[0]
            assert torch.equal(one, mels[k]), ("mel", u)
            w1 = be.hift.inference_batch([one], source=src[k:k + 1, :, :gen_frames[u] * HOP], uniform_stride=True)[0][0]
            assert torch.equal(w1, wb[k]), ("wav", u)
    del be
    monkeypatch.delenv("CVB200_FUSED_FFN")
    # ---- fp32 engine against the CPU oracle: shortest (10 steps) and longest (2 steps) member
    torch.set_flush_denormal(True)
    be32 = AcousticBackend(esd, hsd, v2=False, dtype=torch.float32, device=DEV, frontend_sd=fsd)
    order = sorted(range(N), key=lambda u: tot[u])
    for u, steps in ((order[0], 10), (order[-1], 2)):
        t, pt, pf, e, z = inputs(u)
        mu, spks, cond, l1, l2 = encoder_ref.frontend_v1(fsd, t, pt, pf, e, sample_rate=SR)
        assert (l1, l2) == (PROMPT, gen_frames[u])
        want = flow_ref.solve_euler(esd, z, flow_ref.cosine_t_span(steps), mu, torch.ones(1, 1, l1 + l2), spks, cond)[:, :, l1:]
        got = be32.flow.inference_batch([t], [pt], [pf], [e], [z], steps, SR)[0].cpu()
        err = (got - want).abs().max().item()
        src = torch.randn(1, 1, l2 * HOP, generator=torch.Generator().manual_seed(8)) * 0.1
        want_w = hift_ref.decode(hsd, want, src, v2=False)
        wav = be32.hift.inference_batch([want], source=src)[0][0].cpu()
        werr = (wav - want_w).abs().max().item()
        print(f"config 5 member {u} (T = {tot[u]}, {steps} Euler steps) fp32 vs oracle: mel {err:.3e}, waveform {werr:.3e}")
        assert err < 2e-3 and werr < 1e-3, (u, err, werr)


def test_sweep_to_host_equals_batch_by_batch_calls():
    """AcousticBackend.synthesize_tokens_batches_to_host (pinned results, device->host copies on a side stream under the next
    batch) returns, batch by batch, exactly what synthesize_tokens_batch returns on the device -- also when consecutive
    batches reuse one HiFT program (same stride bucket), i.e. the copy of batch i is not overtaken by batch i + 1."""
    from cosyvoice_b200.pipeline import AcousticBackend
    esd, hsd, fsd = (init_weights.flow_estimator_sd(False, seed=0), init_weights.hift_sd(False, seed=1),
                     init_weights.flow_frontend_sd(False, seed=2))
    be = AcousticBackend(esd, hsd, v2=False, dtype=torch.bfloat16, device=DEV, frontend_sd=fsd)
    n_tok = [[60, 75], [70, 64], [150], [66, 72]]              # batches 0, 1 and 3 share a stride bucket

    def batch(k):
        cols = [[], [], [], [], []]
        for i, n in enumerate(n_tok[k]):
            t, pt, pf, e = init_weights.frontend_inputs(False, n, 25, seed=100 * k + i, sample_rate=SR)
            T = pf.shape[1] + int(n / 25 * SR / 256)
            for c, v in zip(cols, (t, pt, pf, e, torch.randn(1, 80, T, generator=torch.Generator().manual_seed(k * 10 + i)))):
                c.append(v)
        return tuple(cols)
    batches = [batch(k) for k in range(len(n_tok))]
    torch.manual_seed(5)
    want = [[w.cpu() for w in be.synthesize_tokens_batch(*b, sample_rate=SR)] for b in batches]
    torch.manual_seed(5)
    got = be.synthesize_tokens_batches_to_host(iter(batches), sample_rate=SR)
    assert be.last_sweep["batches"] == len(batches)
    for k, (g, w) in enumerate(zip(got, want)):
        assert len(g) == len(w)
        for a, b in zip(g, w):
            assert a.is_pinned() and a.shape == b.shape and torch.equal(a, b), k
