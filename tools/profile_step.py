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
be = AcousticBackend(synthetic.flow_estimator_sd(False, seed=0), synthetic.hift_sd(False, seed=1), v2=False,
                     dtype=torch.bfloat16, device="cuda")
be.cfm.use_cuda_graph = False
mus, spks, conds, zs = synthetic.flow_inputs(a.batch, 937, 281)
res = be.prepare_resident(mus, spks, conds, zs, [281] * a.batch)
be.step_resident(res, a.timesteps)
torch.cuda.synchronize()
print("ok", float(res["hprog"].wav.abs().max()))
