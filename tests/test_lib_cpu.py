"""No-GPU checks of the C-ABI library: it loads, exports every symbol include/cvb200.h declares, and argument
validation fails loudly (no compute is launched here)."""
import os
import re

import pytest
import torch

from cosyvoice_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_lib.LIB_PATH):
        from cosyvoice_b200 import build
        build.build()
    return _lib.load()


def test_header_symbols_are_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "cvb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(cvb200_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/cvb200.h but not exported"
    assert names == set(_lib.exported_symbols())


def test_null_arguments_are_rejected_with_message(lib):
    st = lib.cvb200_layernorm(None, 256, None, 256, 0, None, None, 8, 256, 1e-5, None)
    assert st == 1
    assert b"null" in lib.cvb200_last_error()
    with pytest.raises(RuntimeError, match="null"):
        _lib.check(st, "layernorm")


def test_ops_refuse_cpu_tensors(lib):
    from cosyvoice_b200 import ops
    x = torch.zeros(128, 256)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.layernorm_op(x, x.clone(), torch.ones(256), torch.zeros(256), 128)


def test_struct_sizes_match_c_layout():
    import ctypes as C
    # 8 pointers + 16 4-byte fields + 2 pointers (folded-LayerNorm row statistics)
    assert C.sizeof(_lib.Epilogue) == 8 * 8 + 16 * 4 + 2 * 8 + 2 * 8 + 2 * 4
    assert C.sizeof(_lib.Layout) == 4 * 8 + 3 * 4 + 4
    assert C.sizeof(_lib.Gemm) == 6 * 8 + 11 * 4 + 2 * 2 * 32 + 4 + C.sizeof(_lib.Epilogue)


def test_product_package_never_reaches_the_oracle():
    """oracle/ is test infrastructure: nothing under cosyvoice_b200/ may import or execute it (static scan of every
    module of the package), and a missing shared library is a loud error, not a fallback."""
    import subprocess
    import sys
    pkg = os.path.join(ROOT, "cosyvoice_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), os.path.join(dirpath, f)
                assert not re.search(r"import_module\(\s*['\"]oracle|__import__\(\s*['\"]oracle", src), os.path.join(dirpath, f)
    code = ("import os, sys; os.environ['CVB200_LIB'] = '/nonexistent/libcvb200.so'; sys.path.insert(0, %r)\n"
            "from cosyvoice_b200 import _lib\n"
            "try:\n    _lib.load()\nexcept RuntimeError as e:\n    print('LOUD', e)\n" % ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert "LOUD" in out.stdout and "not found" in out.stdout, out.stdout + out.stderr


def test_cache_trim_and_async_staging_on_cpu():
    """Host-side helpers that have a CPU meaning: ops.trim_cache bounds an LRU dict by count (oldest first, never below the
    newest `keep` entries; the memory bound only exists on CUDA), and layout.to_device_async / copy_async_ are plain copies
    when no CUDA device is involved."""
    import torch
    from cosyvoice_b200 import ops
    from cosyvoice_b200.layout import SeqLayout, copy_async_, to_device_async
    cache = {k: object() for k in range(5)}
    ops.trim_cache(cache, "cpu", max_entries=4)
    assert list(cache) == [2, 3, 4]                      # room for one more below the bound of 4
    ops.trim_cache(cache, "cpu", max_entries=32)
    assert list(cache) == [2, 3, 4]
    ops.trim_cache(cache, "cpu", max_entries=1)
    assert cache == {}
    src = torch.arange(6, dtype=torch.int64)
    out = to_device_async(src, "cpu", torch.int32)
    assert out.dtype == torch.int32 and out.tolist() == list(range(6))
    dst = torch.zeros(6, dtype=torch.float32)
    copy_async_(dst, src)
    assert dst.tolist() == [0.0, 1.0, 2.0, 3.0, 4.0, 5.0]
    lay = SeqLayout.build([5, 130], "cpu")
    lay.set_lens([7, 100])
    assert lay.seq_len.tolist() == [7, 100] and lay.max_len == 100
