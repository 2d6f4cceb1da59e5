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
