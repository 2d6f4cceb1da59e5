"""Micro-benchmark of the transformer block's second half (run on the GPU box): the fused cvb200_block_ffn kernel against the
separate launches it replaces (LayerNorm, FFN-in GEMM + GELU, FFN-out GEMM + residual, next LayerNorm) on the bf16 stream.
    python tools/ffn_bench.py [--T 609] [--nseq 64]
CUDA events, 3 warm-up + 10 timed launches, L2 flushed between launches."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from cosyvoice_b200 import _lib, ops
from cosyvoice_b200.layout import SeqLayout
from cosyvoice_b200.packing import linear_plan

DEV, bf = "cuda", torch.bfloat16


def timed(fns, reps=10):
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)
    for _ in range(3):
        for f in fns:
            f()
    torch.cuda.synchronize()
    ms = 0.0
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for f in fns:
            f()
        e1.record()
        torch.cuda.synchronize()
        ms += e0.elapsed_time(e1)
    return ms / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--T", type=int, default=609)
    ap.add_argument("--nseq", type=int, default=64)
    a = ap.parse_args()
    lay = SeqLayout.build([a.T] * a.nseq, DEV, align=128)
    R = lay.rows
    torch.manual_seed(0)
    x = (torch.randn(R, 256, device=DEV)).to(bf)
    n, f = torch.zeros(R, 256, device=DEV, dtype=bf), torch.zeros(R, 1024, device=DEV, dtype=bf)
    w1, b1 = torch.randn(1024, 256) * .06, torch.randn(1024) * .1
    w2, b2 = torch.randn(256, 1024) * .03, torch.randn(256) * .1
    g, be = torch.ones(256, device=DEV), torch.zeros(256, device=DEV)
    p1, p2 = linear_plan(w1, b1, bf, DEV), linear_plan(w2, b2, bf, DEV)
    fl = 4 * 256 * 1024 * a.T * a.nseq
    unfused = [ops.layernorm_op(x, n, g, be, R), ops.gemm_op(p1, n, lay, f, engine=_lib.ENGINE_TCGEN05, act=_lib.ACT_GELU),
               ops.gemm_op(p2, f, lay, x, engine=_lib.ENGINE_TCGEN05, res=x), ops.layernorm_op(x, n, g, be, R)]
    w1f, u1, b1f = ops.fold_layernorm(w1.to(DEV), b1.to(DEV), g, be, bf)
    w2d, b2d = w2.to(DEV, bf).contiguous(), b2.to(DEV)
    stats = torch.zeros(R, 2, device=DEV)
    fused = [ops.block_ffn_op(x, w1f, u1, b1f, w2d, b2d, lay, stats_out=stats)]
    fused_plain = [ops.block_ffn_op(x, w1f, u1, b1f, w2d, b2d, lay)]
    for name, fns in (("LN + FFN-in + FFN-out + LN (4 launches)", unfused), ("block_ffn (+ row stats)", fused),
                      ("block_ffn (no stats)", fused_plain), ("LN alone", unfused[:1]), ("FFN-in alone", unfused[1:2]),
                      ("FFN-out alone", unfused[2:3])):
        ms = timed(fns)
        print(f"{name:42s} rows={R:7d} tiles={R // 128:4d}  {ms * 1e3:8.1f} us  {fl / ms / 1e9:7.1f} TFLOP/s (FFN flops)")
        x.copy_(torch.randn(R, 256, device=DEV).to(bf))      # keep the in-place stream bounded


if __name__ == "__main__":
    main()
