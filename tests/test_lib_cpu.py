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
