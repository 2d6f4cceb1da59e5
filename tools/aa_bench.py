"""Stand-alone timing of the anti-aliased SnakeBeta kernels against the HBM roofline.
    python tools/aa_bench.py            rows layout (trunk: fp32 in -> bf16 out) and [B,C,T] layout (fp32 / bf16)"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from cosyvoice_b200 import _lib, ops
from cosyvoice_b200.bigvgan import kaiser_sinc_filter12
from cosyvoice_b200.layout import SeqLayout

dev = torch.device("cuda:0")
PEAK = 6546.6
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
filt = kaiser_sinc_filter12(dev)
REF = None
if "--ref" in sys.argv:       # the reference's own fwd_cuda built for sm_100a (oracle/build_ref.py): the kernel to beat
    from oracle import build_ref
    REF = build_ref.load()
    assert REF is not None, "oracle/_ref not built"


def timed(fn, reps=5):
    fn()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


B, frames = 64, 938
P = 1
for u, c in zip([4, 4, 2, 2, 2, 2], [384, 192, 96, 48, 24, 12]):
    P *= u
    lay = SeqLayout.build([frames * P] * B, dev, align=128 * P)
    ld = (c + 63) // 64 * 64
    for xdt in (torch.float32, torch.bfloat16):
        x = torch.randn(lay.rows, ld, device=dev).to(xdt)
        out = torch.empty(lay.rows, ld, device=dev, dtype=torch.bfloat16)
        al, be = torch.randn(c, device=dev) * 0.3, torch.randn(c, device=dev) * 0.3
        ms = timed(lambda: ops.anti_alias_act_rows(x, out, filt, al, be, c, lay))
        n = B * frames * P * c
        gb = n * (x.element_size() + 2) / 1e9
        print(json.dumps({"kernel": "anti_alias_act_rows", "C": c, "ld": ld, "rows": lay.rows, "x": str(xdt)[6:], "ms": round(ms, 4),
                          "algorithmic_GBps": round(gb / ms * 1e3, 1), "frac": round(gb / ms * 1e3 / PEAK, 3)}))
        del x, out
for dt in (torch.float32, torch.bfloat16):
    for c, T in ((384, 15008), (96, 60032), (24, 240128)):
        x = torch.randn(B, c, T, device=dev).to(dt)
        al, be = torch.randn(c, device=dev) * 0.3, torch.randn(c, device=dev) * 0.3
        y = torch.empty_like(x)
        fn = lambda: _lib.check(_lib.load().cvb200_anti_alias_act_nct(
            x.data_ptr(), y.data_ptr(), ops.dt_of(x), B, c, T, filt.data_ptr(), filt.data_ptr(), al.data_ptr(), be.data_ptr(),
            torch.cuda.current_stream().cuda_stream))
        ms = timed(fn)
        gb = 2 * x.numel() * x.element_size() / 1e9
        print(json.dumps({"kernel": "anti_alias_act_nct", "shape": [B, c, T], "dtype": str(dt)[6:], "ms": round(ms, 4),
                          "algorithmic_GBps": round(gb / ms * 1e3, 1), "frac": round(gb / ms * 1e3 / PEAK, 3)}))
        if REF is not None:
            f3 = filt.reshape(1, 1, -1)
            ms_r = timed(lambda: REF.forward(x, f3, f3, al, be))
            print(json.dumps({"kernel": "reference fwd_cuda (sm_100a build, --use_fast_math)", "shape": [B, c, T],
                              "dtype": str(dt)[6:], "ms": round(ms_r, 4), "algorithmic_GBps": round(gb / ms_r * 1e3, 1),
                              "frac": round(gb / ms_r * 1e3 / PEAK, 3), "speedup_ours": round(ms_r / ms, 2)}))
        del x, y
