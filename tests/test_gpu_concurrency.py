"""Re-entrancy of the boundary objects (SURVEY.md 8b "Threading"): the reference keeps several `tts()` sessions alive at once
and runs the LLM on another thread + CUDA stream (cosyvoice/cli/model.py:62-69,119,183-189), and its flow / hift allocate per
call.  Here programs (buffers + launch lists) are cached, so calls are serialised per object on the host (lock) and ordered
across streams on the device (event), without any device-wide synchronisation.  Four threads, each on its own CUDA stream,
hammer ONE flow object and ONE vocoder with different utterances of the same stride bucket (= the same cached programs):
every result must equal the single-threaded one bit for bit."""
import threading

import pytest
import torch

pytestmark = pytest.mark.gpu

from cosyvoice_b200.frontend import B200CausalMaskedDiffWithXvec
from cosyvoice_b200.hift import B200HiFTGenerator
from oracle import init_weights

DEV = "cuda"


def test_sessions_on_four_threads_and_streams_share_one_flow_and_one_vocoder():
    flow = B200CausalMaskedDiffWithXvec(init_weights.flow_frontend_sd(True, seed=2), dtype=torch.bfloat16, device=DEV,
                                        estimator_state_dict=init_weights.flow_estimator_sd(True, seed=0))
    hift = B200HiFTGenerator(init_weights.hift_sd(True, seed=1), v2=True, dtype=torch.bfloat16, device=DEV)
    n_utt = 8
    cases = []
    for u in range(n_utt):
        tok, ptok, pfeat, emb = init_weights.frontend_inputs(True, 20 + u, 6, seed=700 + u)      # 26..33 tokens: one stride bucket
        g = torch.Generator().manual_seed(800 + u)
        T = 2 * (20 + u)
        pv = torch.rand(1, 9, 1, generator=g) * 6.28 - 3.14
        pv[:, 0] = 0
        cases.append(dict(tok=tok.to(DEV), ptok=ptok.to(DEV), pfeat=pfeat.to(DEV), emb=emb.to(DEV), pv=pv,
                          noise=torch.randn(1, 9, T * 480, generator=g)))
    i32 = lambda n: torch.tensor([n], dtype=torch.int32)

    def synth(c):
        mel, _ = flow.inference(token=c["tok"], token_len=i32(c["tok"].shape[1]), prompt_token=c["ptok"],
                                prompt_token_len=i32(c["ptok"].shape[1]), prompt_feat=c["pfeat"],
                                prompt_feat_len=i32(c["pfeat"].shape[1]), embedding=c["emb"])
        wavs, _ = hift.inference_batch([mel], phase_vec=c["pv"], noise=c["noise"])
        return mel, wavs[0]

    want = [synth(c) for c in cases]            # single thread, default stream (also builds programs / captures graphs)
    torch.cuda.synchronize()
    errors, results = [], {}

    def worker(k):
        try:
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                for rep in range(4):
                    for u in range(k, n_utt, 4):
                        mel, wav = synth(cases[u])
                        results[(k, rep, u)] = (mel, wav)
                s.synchronize()
        except Exception as e:          # noqa: BLE001
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    assert len(results) == 4 * n_utt
    for (k, rep, u), (mel, wav) in results.items():
        assert torch.equal(mel, want[u][0]), (k, rep, u, "mel")
        assert torch.equal(wav, want[u][1]), (k, rep, u, "wav")


def test_cold_graph_capture_under_contention_with_different_stride_buckets():
    """Four threads start with COLD program and CUDA-graph caches, each on utterances of a different stride bucket, so every
    thread builds its programs and captures its solver graph while the others allocate, copy and launch (round-1 advice:
    a capture in torch's default "global" error mode is invalidated by another thread's cudaMalloc / H2D copy; the solver
    captures thread-locally).  Results must equal those of a fresh single-threaded object bit for bit."""
    def build():
        flow = B200CausalMaskedDiffWithXvec(init_weights.flow_frontend_sd(True, seed=2), dtype=torch.bfloat16, device=DEV,
                                            estimator_state_dict=init_weights.flow_estimator_sd(True, seed=0))
        hift = B200HiFTGenerator(init_weights.hift_sd(True, seed=1), v2=True, dtype=torch.bfloat16, device=DEV)
        return flow, hift
    n_tok = [20, 70, 130, 190]                   # mel lengths 2*(n+6): 52 / 152 / 272 / 392 frames = four 128-row buckets
    cases = []
    for u, n in enumerate(n_tok):
        tok, ptok, pfeat, emb = init_weights.frontend_inputs(True, n, 6, seed=900 + u)
        g = torch.Generator().manual_seed(950 + u)
        pv = torch.rand(1, 9, 1, generator=g) * 6.28 - 3.14
        pv[:, 0] = 0
        cases.append(dict(tok=tok, ptok=ptok, pfeat=pfeat, emb=emb, pv=pv, noise=torch.randn(1, 9, 2 * n * 480, generator=g)))
    i32 = lambda n: torch.tensor([n], dtype=torch.int32)

    def synth(flow, hift, c):
        # host tensors on purpose: the H2D copies and allocations of one thread run while another thread captures
        mel, _ = flow.inference(token=c["tok"].to(DEV), token_len=i32(c["tok"].shape[1]), prompt_token=c["ptok"].to(DEV),
                                prompt_token_len=i32(c["ptok"].shape[1]), prompt_feat=c["pfeat"].to(DEV),
                                prompt_feat_len=i32(c["pfeat"].shape[1]), embedding=c["emb"].to(DEV))
        wavs, _ = hift.inference_batch([mel], phase_vec=c["pv"], noise=c["noise"])
        return mel, wavs[0]

    flow0, hift0 = build()
    want = [synth(flow0, hift0, c) for c in cases]
    torch.cuda.synchronize()
    flow, hift = build()                         # cold caches
    errors, results = [], {}
    start = threading.Barrier(4)

    def worker(k):
        try:
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                start.wait()
                for rep in range(2):
                    results[(k, rep)] = synth(flow, hift, cases[k])
                s.synchronize()
        except Exception as e:          # noqa: BLE001
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for (k, rep), (mel, wav) in results.items():
        assert torch.equal(mel, want[k][0]), (k, rep, "mel")
        assert torch.equal(wav, want[k][1]), (k, rep, "wav")
