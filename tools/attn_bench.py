"""Attention micro-benchmark (run on the GPU box): python tools/attn_bench.py [--T 1218] [--nseq 64]"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cosyvoice_b200 import _lib, ops
from cosyvoice_b200.layout import SeqLayout

ap = argparse.ArgumentParser()
ap.add_argument("--T", type=int, default=1218)
ap.add_argument("--nseq", type=int, default=64)
a = ap.parse_args()
lay = SeqLayout.build([a.T] * a.nseq, "cuda", align=256)
qkv = (torch.randn(lay.rows, 1536, device="cuda") * 1.5).to(torch.bfloat16)
fl = a.nseq * 4 * 512 * a.T * a.T
outs = {}
for name, impl in (("mma.sync", _lib.ATTN_MMA_SYNC), ("tcgen05", _lib.ATTN_TCGEN05), ("tcgen05 v2", _lib.ATTN_TCGEN05_V2)):
    out = torch.zeros(lay.rows, 512, device="cuda", dtype=torch.bfloat16)
    op = ops.attention_op(qkv, out, lay, 8, 0.125, impl=impl)
    for _ in range(3):
        op()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        op()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    outs[name] = out
    print(f"{name:10s} T={a.T} nseq={a.nseq}: {ms * 1e3:8.1f} us  {fl / ms / 1e9:7.1f} TFLOP/s")
print("max |tcgen05 - mma.sync| =", (outs["tcgen05"].float() - outs["mma.sync"].float()).abs().max().item())
print("max |tcgen05 v2 - mma.sync| =", (outs["tcgen05 v2"].float() - outs["mma.sync"].float()).abs().max().item())
