"""One eager (no CUDA graph) pass of the bench workload, for `ncu` launch lists:
    python tools/profile_step.py [--batch 32] [--timesteps 10]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from cosyvoice_b200 import synthetic
from cosyvoice_b200.pipeline import AcousticBackend

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--timesteps", type=int, default=10)
a = ap.parse_args()
import bench  # noqa: E402  (utterance_inputs: the bench's synthetic tokens / prompt / noise)

be = AcousticBackend(synthetic.flow_estimator_sd(False, seed=0), synthetic.hift_sd(False, seed=1), v2=False,
                     dtype=torch.bfloat16, device="cuda", frontend_sd=synthetic.flow_frontend_sd(False, seed=2))
be.cfm.use_cuda_graph = False
toks, ptoks, pfeats, embs, zs = bench.utterance_inputs(a.batch)
res = be.prepare_resident_tokens(toks, ptoks, pfeats, embs, zs, sample_rate=bench.SR)     # runs the front-end once (binding)
be.step_resident(res, a.timesteps)                                                         # front-end + CFM + HiFT
torch.cuda.synchronize()
print("ok", float(res["hprog"].wav.abs().max()))
