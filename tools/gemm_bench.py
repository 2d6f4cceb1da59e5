"""Micro-benchmark of the tap GEMM on the shapes of the hot path (run on the GPU box):
    python tools/gemm_bench.py [--rows 81920]
Prints per-shape time and TFLOP/s (CUDA events, 3 warm-up + 10 timed launches, L2 flushed between launches)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from cosyvoice_b200 import _lib, ops
from cosyvoice_b200.layout import SeqLayout
from cosyvoice_b200.packing import conv1d_plan, conv_transpose1d_plan, linear_plan

DEV = "cuda"
bf = torch.bfloat16


BN_OVERRIDE = 0


def bench(name, plan, A, lay, out0, reps=10, **kw):
    if BN_OVERRIDE and plan.n_pad % BN_OVERRIDE == 0:
        plan.bn = BN_OVERRIDE      # experiment: N-tile width (tail balance / resident-W mode)
        name = f"{name} [bn {BN_OVERRIDE}]"
    op = ops.gemm_op(plan, A, lay, out0, engine=_lib.ENGINE_TCGEN05, **kw)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)
    for _ in range(3):
        op()
    torch.cuda.synchronize()
    ms = 0.0
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        op()
        e1.record()
        torch.cuda.synchronize()
        ms += e0.elapsed_time(e1)
    ms /= reps
    fl = plan.flops_per_row * sum(lay.lens)
    print(f"{name:38s} M={lay.rows:8d} N={plan.N:5d} K={plan.n_seg * plan.seg_len:5d} phases={plan.n_phase}  "
          f"{ms * 1e3:9.1f} us  {fl / ms / 1e9:8.1f} TFLOP/s")
    return ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=64 * 1280)
    ap.add_argument("--only", default="")
    ap.add_argument("--bn", type=int, default=0, help="override the plan's N tile (0 = packing._choose_bn)")
    a = ap.parse_args()
    global BN_OVERRIDE
    BN_OVERRIDE = a.bn
    nseq = 64
    T = a.rows // nseq - 62
    lay = SeqLayout.build([T] * nseq, DEV, align=256)
    R = lay.rows
    z = lambda c, dt=bf: torch.randn(R, c, device=DEV).to(dt)
    X = torch.randn(R, 256, device=DEV)
    cases = {
        "qkv 256->1536": lambda: bench("qkv 256->1536", linear_plan(torch.randn(1536, 256) * .06, None, bf, DEV), z(256), lay, z(1536), bias=False),
        "ffn1 256->1024 gelu": lambda: bench("ffn1 256->1024 gelu", linear_plan(torch.randn(1024, 256) * .06, torch.randn(1024), bf, DEV), z(256), lay, z(1024), act=_lib.ACT_GELU),
        "ffn2 1024->256 +res f32": lambda: bench("ffn2 1024->256 +res f32", linear_plan(torch.randn(256, 1024) * .03, torch.randn(256), bf, DEV), z(1024), lay, X, res=X),
        "out 512->256 +res f32": lambda: bench("out 512->256 +res f32", linear_plan(torch.randn(256, 512) * .04, torch.randn(256), bf, DEV), z(512), lay, X, res=X),
        "conv k3 256->256 +gn": lambda: bench("conv k3 256->256 +gn", conv1d_plan(torch.randn(256, 256, 3) * .03, torch.randn(256), bf, DEV), z(256), lay, z(256), gn_stats=torch.zeros(nseq, 8, 2, dtype=torch.float64, device=DEV)),
        "conv k3 512->256 +gn": lambda: bench("conv k3 512->256 +gn", conv1d_plan(torch.randn(256, 512, 3) * .03, torch.randn(256), bf, DEV), z(512), lay, z(256), gn_stats=torch.zeros(nseq, 8, 2, dtype=torch.float64, device=DEV)),
        "proj 256->80 f32": lambda: bench("proj 256->80 f32", linear_plan(torch.randn(80, 256) * .06, torch.randn(80), bf, DEV), z(256), lay, torch.zeros(R, 80, device=DEV)),
    }
    # HiFT last stage: 128 channels at 64x rate (batch 32 x 1024-row stride x 64)
    def hift_cases():
        lay2 = SeqLayout.build([937 * 64 + 1] * 32, DEV, align=128 * 64)
        R2 = lay2.rows
        a_ = torch.randn(R2, 128, device=DEV).to(bf)
        b_ = torch.zeros(R2, 128, device=DEV, dtype=bf)
        xr = torch.randn(R2, 128, device=DEV)
        al = (1 + .3 * torch.randn(128)).to(DEV)
        for k, d in ((11, 5), (7, 1), (3, 1)):
            plan = conv1d_plan(torch.randn(128, 128, k) * .03, torch.randn(128), bf, DEV, dilation=d)
            bench(f"hift conv k{k} d{d} 128 snake->bf16", plan, a_, lay2, b_, act=_lib.ACT_SNAKE, act_alpha=al)
            bench(f"hift conv k{k} 128 +res f32, out1 snake", plan, a_, lay2, xr, res=xr, out1=b_, out1_mode=_lib.OUT1_SNAKE, out1_alpha=al)
        lay1 = SeqLayout.build([937 * 8] * 32, DEV, align=128 * 8)
        x1 = torch.randn(lay1.rows, 256, device=DEV).to(bf)
        plan = conv_transpose1d_plan(torch.randn(256, 128, 16) * .03, torch.randn(128), bf, DEV, stride=8, padding=4)
        bench("hift convT k16 s8 256->128 f32", plan, x1, lay1, xr)
    def ffn_fused_case():
        w1 = (torch.randn(1024, 256, device=DEV) * .06).to(bf)
        w2 = (torch.randn(256, 1024, device=DEV) * .03).to(bf)
        b1, b2 = torch.randn(1024, device=DEV), torch.randn(256, device=DEV)
        op = ops.ffn_fused_op(z(256), w1, b1, w2, b2, X, lay)
        flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)
        for _ in range(3):
            op()
        torch.cuda.synchronize()
        ms = 0.0
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            op()
            e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        ms /= 10
        fl = 2 * 2 * 256 * 1024 * sum(lay.lens)
        print(f"{'ffn fused 256->1024->256 +res':38s} M={lay.rows:8d} N=  256 K= 1024 phases=1  {ms * 1e3:9.1f} us  {fl / ms / 1e9:8.1f} TFLOP/s")
    cases["ffn fused"] = ffn_fused_case
    for name, fn in cases.items():
        if a.only in name:
            fn()
    if a.only in "hift":
        hift_cases()


if __name__ == "__main__":
    main()
