"""cvb200_sine_source_noalloc (include/cvb200.h): the NSF sine source with its fp64 frame-prefix scratch inside the
library's device image instead of a caller workspace.  Called through the C ABI twice in a row (different f0 / lengths,
so stale scratch from the first call would show) and compared with cvb200_sine_source on the same inputs.  Both entry
points run the same two kernels, so the outputs are required to be identical; the test also runs while torch's caching
allocator is watched: the new entry point must not allocate.

Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

from cosyvoice_b200 import _lib
from cosyvoice_b200.layout import SeqLayout

pytestmark = pytest.mark.gpu

HOP, SR = 256, 24000.0


def _inputs(lens, seed, dev):
    g = torch.Generator().manual_seed(seed)
    lay = SeqLayout.build(lens, dev)
    f0 = torch.zeros(lay.rows)
    for b, l in enumerate(lens):
        v = torch.rand(l, generator=g) * 300 + 60
        v[torch.rand(l, generator=g) < 0.2] = 0.0          # unvoiced frames (uv gate, generator.py:159-163)
        f0[lay.row0[b]:lay.row0[b] + l] = v
    l_max = max(lens) * HOP
    noise = torch.randn(len(lens), 9, l_max, generator=g)
    phase = (torch.rand(len(lens), 9, generator=g) * 2 - 1) * 3.14159
    phase[:, 0] = 0
    lin_w, lin_b = torch.randn(9, generator=g) * 0.3, torch.randn(1, generator=g) * 0.1
    return lay, [t.to(dev).contiguous() for t in (f0, noise, phase, lin_w, lin_b)], l_max


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def test_noalloc_sine_source_matches_the_workspace_entry_point_twice():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    lib = _lib.load()
    for call, (lens, seed) in enumerate((([37, 12, 50], 11), ([9, 61], 12))):
        lay, (f0, noise, phase, lin_w, lin_b), l_max = _inputs(lens, seed, dev)
        ws = torch.zeros(lay.rows, dtype=torch.float64, device=dev)
        want = torch.zeros(len(lens), l_max, device=dev)
        got = torch.zeros(len(lens), l_max, device=dev)
        _lib.check(lib.cvb200_sine_source(f0.data_ptr(), C.byref(lay.c), HOP, SR, noise.data_ptr(), phase.data_ptr(),
                                          lin_w.data_ptr(), lin_b.data_ptr(), ws.data_ptr(), want.data_ptr(), l_max,
                                          _stream()), "sine_source")
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()["allocation.all.allocated"]
        _lib.check(lib.cvb200_sine_source_noalloc(f0.data_ptr(), C.byref(lay.c), HOP, SR, noise.data_ptr(),
                                                  phase.data_ptr(), lin_w.data_ptr(), lin_b.data_ptr(), got.data_ptr(),
                                                  l_max, _stream()), "sine_source_noalloc")
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == before
        assert want.abs().max().item() > 0.01
        assert torch.equal(got, want), f"call {call}: max-abs diff {(got - want).abs().max().item():.3e}"


def test_noalloc_sine_source_rejects_layouts_beyond_its_scratch():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    lib = _lib.load()
    lay = SeqLayout.build([70000], dev)     # 70 016 frame rows > CVB200_NOALLOC_MAX_ROWS = 65 536
    one = torch.zeros(16, device=dev)
    st = lib.cvb200_sine_source_noalloc(one.data_ptr(), C.byref(lay.c), HOP, SR, one.data_ptr(), one.data_ptr(),
                                        one.data_ptr(), one.data_ptr(), one.data_ptr(), 70000 * HOP, _stream())
    assert st == 1
    assert b"CVB200_NOALLOC_MAX_ROWS" in lib.cvb200_last_error()
