import sys, torch
sys.path.insert(0, "/root/repo")
from cosyvoice_b200 import ops
from cosyvoice_b200.layout import SeqLayout
for T, nseq in ((609, 64), (1218, 64)):
    lay = SeqLayout.build([T] * nseq, "cuda", align=128)
    R = lay.rows
    h = torch.randn(R, 256, device="cuda").to(torch.bfloat16)
    o = torch.zeros_like(h)
    g, b = torch.ones(256, device="cuda"), torch.zeros(256, device="cuda")
    st = torch.zeros(nseq, 8, 2, dtype=torch.float64, device="cuda")
    st[:, :, 0] = 0.0; st[:, :, 1] = T * 32.0
    op = ops.norm_mish_op(h, o, g, b, lay, gn_stats=st)
    for _ in range(5): op()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): op()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 50 * 1e3
    print(f"norm_mish rows={R}: {us:.1f} us, {R * 256 * 4 / us / 1e6:.2f} TB/s")
