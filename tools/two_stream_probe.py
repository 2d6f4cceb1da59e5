"""Does running two half-batches on two CUDA streams beat one full batch on one stream?  (Tails of persistent kernels --
wave quantisation, ~28 % of block_ffn at 320 tiles -- could be filled by the other half's kernels; HBM-bound launches could
overlap tensor-bound ones.)  Builder probe on the bench workload:
    python tools/two_stream_probe.py [--batch 32] [--reps 5]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from cosyvoice_b200 import synthetic
from cosyvoice_b200.pipeline import AcousticBackend
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--ways", type=int, default=2)
a = ap.parse_args()


def backend():
    return AcousticBackend(synthetic.flow_estimator_sd(False, seed=0), synthetic.hift_sd(False, seed=1), v2=False,
                           dtype=torch.bfloat16, device="cuda", frontend_sd=synthetic.flow_frontend_sd(False, seed=2))


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


toks, ptoks, pfeats, embs, zs = bench.utterance_inputs(a.batch)
be = backend()
res = be.prepare_resident_tokens(toks, ptoks, pfeats, embs, zs, sample_rate=bench.SR)
one = timed(lambda: be.step_resident(res), a.reps)
print(f"one stream, batch {a.batch}: {one:.2f} ms/step")
del be, res
torch.cuda.empty_cache()

W = a.ways
h = a.batch // W
bes = [backend() for _ in range(W)]
streams = [torch.cuda.Stream() for _ in range(W)]
ress = []
for k in range(W):
    sl = slice(k * h, (k + 1) * h)
    with torch.cuda.stream(streams[k]):
        ress.append(bes[k].prepare_resident_tokens(toks[sl], ptoks[sl], pfeats[sl], embs[sl], zs[sl], sample_rate=bench.SR))
torch.cuda.synchronize()


def both():
    main = torch.cuda.current_stream()
    for k in range(W):
        streams[k].wait_stream(main)
        with torch.cuda.stream(streams[k]):
            bes[k].step_resident(ress[k])
    for k in range(W):
        main.wait_stream(streams[k])


two = timed(both, a.reps)
print(f"{W} streams x batch {h}: {two:.2f} ms/step  ({one / two:.3f}x)")
