"""Flow front-end micro-benchmark (run on the GPU box): tokens -> mu for a batch, CUDA events, per-kernel shares.
    python tools/frontend_bench.py [--batch 32] [--v2] [--tokens 250] [--prompt 75] [--dtype bf16]
Also the command profiled for profiles/r01_ncu_relpos.md (ncu -k regex:relpos_attention -c 1 --set full ...)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cosyvoice_b200 import _lib, ops, synthetic
from cosyvoice_b200.frontend import B200CausalMaskedDiffWithXvec, B200MaskedDiffWithXvec

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--tokens", type=int, default=250)
ap.add_argument("--prompt", type=int, default=75)
ap.add_argument("--v2", action="store_true")
ap.add_argument("--chunk", type=int, default=0)
ap.add_argument("--dtype", default="bf16")
ap.add_argument("--reps", type=int, default=20)
a = ap.parse_args()
dt = torch.bfloat16 if a.dtype == "bf16" else torch.float32
cls = B200CausalMaskedDiffWithXvec if a.v2 else B200MaskedDiffWithXvec
flow = cls(synthetic.flow_frontend_sd(a.v2, seed=2), dtype=dt, device="cuda",
           estimator_state_dict=synthetic.flow_estimator_sd(a.v2, seed=0))
flow.encoder.static_chunk_size = a.chunk
sr = 24000
toks, n_prompt, mel_lens, embs = [], [], [], []
for u in range(a.batch):
    t, pt, pf, e = synthetic.frontend_inputs(a.v2, a.tokens, a.prompt, seed=1234 + u, sample_rate=sr)
    toks.append(torch.cat([pt, t], 1).reshape(-1)); n_prompt.append(a.prompt); embs.append(e)
    mel_lens.append(flow._mel_lens(a.prompt, a.tokens, pf.shape[1], sr))
prog = flow.encode_batch(toks, n_prompt, mel_lens, embs, sr)
for _ in range(3):
    prog.run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n0 = _lib.launch_count()
e0.record()
for _ in range(a.reps):
    prog.run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.reps
launches = (_lib.launch_count() - n0) // a.reps
ops.PROFILE = []
prog.run()
torch.cuda.synchronize()
rec, ops.PROFILE = ops.PROFILE, None
kinds = {}
for kind, fl, x, y in rec:
    d = kinds.setdefault(kind, dict(launches=0, ms=0.0, flops=0))
    d["launches"] += 1; d["ms"] += x.elapsed_time(y); d["flops"] += fl
for d in kinds.values():
    d["tflops"] = d["flops"] / (d["ms"] * 1e-3) / 1e12 if d["ms"] > 0 else None
print(json.dumps({"workload": f"flow front-end ({'v2 UpsampleConformer' if a.v2 else 'v1 Conformer + length regulator'}), batch {a.batch} x "
                              f"({a.prompt}+{a.tokens}) tokens, chunk {a.chunk}", "dtype": a.dtype, "ms": ms, "launches": launches,
                  "algorithmic_gflop": prog.flops / 1e9, "tflops": prog.flops / (ms * 1e-3) / 1e12, "kinds": kinds}))
