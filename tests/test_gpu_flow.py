"""GPU parity of the flow estimator / CFM solver (through the C ABI) against the committed golden
vectors of the REAL reference and against the CPU oracle on the same seeded inputs.

Tolerances: FP32 path -- the reference's own estimator-replacement bar
rtol 1e-2 / atol 1e-5 per call (cosyvoice/bin/export_onnx.py:100), mel after 10 steps max-abs <= 2e-3.
BF16 path -- no worse than the reference's own bf16-vs-fp32 gap: rel-rms <= 3 %, max-abs <= 0.25.
"""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from cosyvoice_b200 import _lib
from cosyvoice_b200.flow import B200CausalConditionalCFM, B200ConditionalCFM, B200Estimator
from oracle import flow_ref, init_weights

DEV = "cuda"
BF16_REL_RMS, BF16_MAX_ABS = 0.03, 0.25


def _load(golden_dir, name):
    return torch.load(os.path.join(golden_dir, name), weights_only=True)


def _rel_rms(a, b):
    return ((a - b).pow(2).mean().sqrt() / b.pow(2).mean().sqrt()).item()


_EST = {}


def _estimator(causal, dtype, engine=None):
    key = (causal, dtype, engine)
    if key not in _EST:
        sd = init_weights.flow_estimator_sd(causal, seed=0)
        _EST[key] = B200Estimator(sd, causal=causal, dtype=dtype, device=DEV, engine=engine)
    return _EST[key]


@pytest.mark.parametrize("ver", ["v1", "v2"])
def test_estimator_fp32_matches_reference_golden(golden_dir, ver):
    d = _load(golden_dir, f"estimator_{ver}.pt")
    est = _estimator(d["causal"], torch.float32)
    out = est(d["x"].to(DEV), d["mask"].to(DEV), d["mu"].to(DEV), d["t"].to(DEV), d["spks"].to(DEV), d["cond"].to(DEV)).cpu()
    ref = d["out"]
    err = (out - ref).abs()
    assert (err <= 1e-5 + 1e-2 * ref.abs()).float().mean().item() > 0.999, err.max().item()
    assert err.max().item() < 2e-3, err.max().item()


@pytest.mark.parametrize("ver", ["v1", "v2"])
def test_estimator_stage_taps_fp32(golden_dir, ver):
    """Intermediate activations against the oracle's taps: localises a wrong layer."""
    d = _load(golden_dir, f"estimator_{ver}.pt")
    causal = d["causal"]
    sd = init_weights.flow_estimator_sd(causal, seed=0)
    taps_ref = {}
    flow_ref.estimator_forward(sd, d["x"], d["mask"], d["mu"], d["t"], d["spks"], d["cond"], causal=causal, taps=taps_ref)
    est = _estimator(causal, torch.float32)
    taps = {}
    est(d["x"].to(DEV), d["mask"].to(DEV), d["mu"].to(DEV), d["t"].to(DEV), d["spks"].to(DEV), d["cond"].to(DEV), taps=taps)
    # only buffers that are not overwritten later in the forward can be compared after the fact
    for name in [n for n in taps if n.startswith("up")]:
        buf, L = taps[name]
        ref = taps_ref[name]
        for b in range(L.n_seq):
            got = buf[L.row0[b]:L.row0[b] + L.lens[b]].float().cpu().t()
            want = ref[b][:, :L.lens[b]]
            assert (got - want).abs().max().item() < 2e-3 * max(1.0, want.abs().max().item()), name


@pytest.mark.parametrize("ver", ["v1", "v2"])
@pytest.mark.parametrize("engine", [_lib.ENGINE_SIMT, _lib.ENGINE_TCGEN05])
def test_estimator_bf16_within_reference_bf16_gap(golden_dir, ver, engine):
    d = _load(golden_dir, f"estimator_{ver}.pt")
    est = _estimator(d["causal"], torch.bfloat16, engine)
    out = est(d["x"].to(DEV), d["mask"].to(DEV), d["mu"].to(DEV), d["t"].to(DEV), d["spks"].to(DEV), d["cond"].to(DEV)).cpu()
    ref = d["out"]
    assert _rel_rms(out, ref) < BF16_REL_RMS, _rel_rms(out, ref)
    assert (out - ref).abs().max().item() < BF16_MAX_ABS


def test_cfm_v1_fp32_golden_with_flow_cache(golden_dir):
    d = _load(golden_dir, "cfm_v1.pt")
    cfm = B200ConditionalCFM(_estimator(False, torch.float32))
    cfm._noise = lambda mu: d["z"].to(mu.device)
    mel, cache = cfm(d["mu"].clone().to(DEV), d["mask"].to(DEV), 10, spks=d["spks"].to(DEV), cond=d["cond"].to(DEV),
                     prompt_len=d["prompt_len"])
    assert (mel.cpu() - d["mel"]).abs().max().item() < 2e-3
    assert torch.equal(cache.cpu(), d["flow_cache"])
    cfm._noise = lambda mu: d["z2"].to(mu.device)
    mel2, cache2 = cfm(d["mu2"].clone().to(DEV), d["mask"].to(DEV), 10, spks=d["spks"].to(DEV), cond=d["cond"].to(DEV),
                       prompt_len=d["prompt_len"], flow_cache=cache)
    assert (mel2.cpu() - d["mel2"]).abs().max().item() < 2e-3
    assert torch.equal(cache2.cpu(), d["flow_cache2"])


def test_cfm_v2_fp32_golden_and_fixed_noise(golden_dir):
    d = _load(golden_dir, "cfm_v2.pt")
    cfm = B200CausalConditionalCFM(_estimator(True, torch.float32))
    assert torch.equal(cfm.rand_noise[:, :, :64], d["rand_noise_head"])      # seed-0 buffer, flow_matching.py:212-213
    mel, cache = cfm(d["mu"].clone().to(DEV), d["mask"].to(DEV), 10, spks=d["spks"].to(DEV), cond=d["cond"].to(DEV))
    assert cache is None
    assert (mel.cpu() - d["mel"]).abs().max().item() < 2e-3


@pytest.mark.parametrize("ffn", ["by_size", "fused"])
@pytest.mark.parametrize("ver", ["v1", "v2"])
def test_cfm_bf16_tcgen05_within_bound(golden_dir, ver, ffn, monkeypatch):
    """Both FFN paths of the bf16 engine against the real reference's mel: the four launches a layout this small
    selects, and block_ffn (folded LayerNorm, hidden tensor in TMEM) forced onto the same small layout."""
    d = _load(golden_dir, f"cfm_{ver}.pt")
    causal = d["causal"]
    if ffn == "fused":
        monkeypatch.setenv("CVB200_FUSED_FFN", "1")
        est = B200Estimator(init_weights.flow_estimator_sd(causal, seed=0), causal=causal, dtype=torch.bfloat16, device=DEV,
                            engine=_lib.ENGINE_TCGEN05)
        with torch.inference_mode():
            names = [getattr(op, "kind", "") for op in est.program(est.layout_for([d["mu"].shape[-1]] * 2)).ops]
        assert "block_ffn" in names, names
    else:
        est = _estimator(causal, torch.bfloat16, _lib.ENGINE_TCGEN05)
        with torch.inference_mode():
            names = [getattr(op, "kind", "") for op in est.program(est.layout_for([d["mu"].shape[-1]] * 2)).ops]
        assert "block_ffn" not in names, names
    cfm = (B200CausalConditionalCFM if causal else B200ConditionalCFM)(est)
    mel = cfm.solve_batch([d["mu"]], [d["spks"]], [d["cond"]], [d["z"]], 10)[0].cpu()
    assert _rel_rms(mel, d["mel"]) < BF16_REL_RMS, _rel_rms(mel, d["mel"])
    assert (mel - d["mel"]).abs().max().item() < BF16_MAX_ABS


@pytest.mark.parametrize("graph", [False, True])
def test_varlen_batch_equals_per_utterance_oracle(graph):
    """Ragged batch (3 utterances, different lengths): each result equals the oracle's B=1 solve."""
    g = torch.Generator().manual_seed(11)
    lens = [29, 64, 47]
    sd = init_weights.flow_estimator_sd(False, seed=0)
    mus = [torch.randn(1, 80, l, generator=g) for l in lens]
    zs = [torch.randn(1, 80, l, generator=g) for l in lens]
    spks = [torch.randn(1, 80, generator=g) for _ in lens]
    conds = []
    for l in lens:
        c = torch.zeros(1, 80, l)
        c[:, :, :7] = torch.randn(1, 80, 7, generator=g) * 2 - 4
        conds.append(c)
    cfm = B200ConditionalCFM(_estimator(False, torch.float32))
    cfm.use_cuda_graph = graph
    mels = cfm.solve_batch(mus, spks, conds, zs, 4)
    if graph:
        mels = cfm.solve_batch(mus, spks, conds, zs, 4)       # second call replays the captured graph
    for b, l in enumerate(lens):
        t_span = flow_ref.cosine_t_span(4)
        want = flow_ref.solve_euler(sd, zs[b], t_span, mus[b], torch.ones(1, 1, l), spks[b], conds[b])
        assert (mels[b].cpu() - want).abs().max().item() < 2e-3, (b, l)


@pytest.mark.parametrize("causal", [False, True])
def test_program_rebinds_lengths_under_one_stride_bucket(causal):
    """Streaming reuse (BASELINE config 3): a program / CUDA graph captured at one length serves other lengths
    that fit the same 128-row stride bucket (longer, then shorter: stale rows of the longer run must not leak).
    Each result equals the oracle's own B=1 solve at that length."""
    g = torch.Generator().manual_seed(21)
    sd = init_weights.flow_estimator_sd(causal, seed=0)
    cfm = (B200CausalConditionalCFM if causal else B200ConditionalCFM)(_estimator(causal, torch.float32))
    cfm.use_cuda_graph = True
    n_prog0 = None
    for l in (60, 110, 37):
        mu, z = torch.randn(1, 80, l, generator=g), torch.randn(1, 80, l, generator=g)
        spk = torch.randn(1, 80, generator=g)
        cond = torch.zeros(1, 80, l)
        cond[:, :, :9] = torch.randn(1, 80, 9, generator=g) * 2 - 4
        mel = cfm.solve_batch([mu], [spk], [cond], [z], 3)[0].cpu()
        want = flow_ref.solve_euler(sd, z, flow_ref.cosine_t_span(3), mu, torch.ones(1, 1, l), spk, cond, causal=causal)
        assert (mel - want).abs().max().item() < 2e-3, l
        if n_prog0 is None:
            n_prog0 = len(cfm._graphs)
    assert len(cfm._graphs) == n_prog0        # one graph served all three lengths


def test_evicted_estimator_programs_are_reclaimed_by_reference_counting():
    """Programs dropped from the LRU cache (32 entries) go away by reference counting alone -- no cycle through the
    launch list -- so a ragged sweep does not pile dead programs up in HBM between Python gc passes."""
    import gc
    import weakref
    est = B200Estimator(init_weights.flow_estimator_sd(False, seed=0), causal=False, dtype=torch.bfloat16, device=DEV)
    gc.collect()
    gc.disable()
    try:
        refs = [weakref.ref(est.program(est.layout_for([100 + 256 * k]))) for k in range(34)]
        assert refs[0]() is None and refs[1]() is None
        assert all(r() is not None for r in refs[2:])
    finally:
        gc.enable()


def test_caches_trimmed_by_memory_give_the_same_results(monkeypatch):
    """ops.trim_cache: with the memory bound forced to zero every new stride bucket evicts all older programs and graphs
    (the situation of a ragged sweep on a nearly full GPU); results are bit-identical to the unbounded run and the caches
    stay at two entries."""
    from cosyvoice_b200 import ops as ops_mod
    sd = init_weights.flow_estimator_sd(False, seed=0)
    g = torch.Generator().manual_seed(33)
    cases = []
    for l in (60, 330, 700, 60):          # stride buckets of 256, 512, 768 rows (v1 aligns to 256), then the first again
        cases.append((torch.randn(1, 80, l, generator=g), torch.randn(1, 80, l, generator=g), torch.randn(1, 80, generator=g),
                      torch.zeros(1, 80, l)))

    def run():
        est = B200Estimator(sd, causal=False, dtype=torch.bfloat16, device=DEV)
        cfm = B200ConditionalCFM(est)
        cfm.use_cuda_graph = True
        outs = [cfm.solve_batch([mu], [spk], [cond], [z], 3)[0].clone() for mu, z, spk, cond in cases]
        return outs, est, cfm
    want, est0, _ = run()
    assert len(est0._programs) == 3
    monkeypatch.setattr(ops_mod, "CACHE_MEMORY_FRACTION", 0.0)
    got, est1, cfm1 = run()
    assert len(est1._programs) <= 2 and len(cfm1._graphs) <= 2      # the newest entry plus the one kept while it was built
    for a, b in zip(got, want):
        assert torch.equal(a, b)
