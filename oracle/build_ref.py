"""Compile the reference's OWN native kernel for this path into oracle/_ref/ -- TEST INFRASTRUCTURE.

The reference ships one native op: `fwd_cuda` (cosyvoice/BigVGAN/alias_free_activation/cuda/
anti_alias_activation_cuda.cu:214-256, bound as `forward` in anti_alias_activation.cpp:21-23), which it JIT-builds at
import time with torch.utils.cpp_extension (cuda/load.py:63-98).  This recipe does the same build from the sources
WHERE THEY LIE under /root/reference (nothing is copied into the repo), with the reference's own flags
(-O3 --use_fast_math, load.py:70-78) except the arch list, which becomes sm_100a: load.py asks for sm_70/sm_80 SASS
only, which cannot run on a B200 at all.

Output: oracle/_ref/anti_alias_activation_cuda.so (git-ignored, travels to the GPU box).  It is used only by
tests/test_gpu_ref_kernel.py (parity of cvb200_anti_alias_act_nct against the real kernel) and by
tools/aa_bench.py --ref (the "kernel to beat on the same box").

    python -m oracle.build_ref
"""
import glob
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REFERENCE_ROOT = os.environ.get("CVB_REFERENCE_ROOT", "/root/reference")
SRC = os.path.join(REFERENCE_ROOT, "cosyvoice", "BigVGAN", "alias_free_activation", "cuda")
NAME = "anti_alias_activation_cuda"


def built_path():
    hits = sorted(glob.glob(os.path.join(OUT, NAME + "*.so")))
    return hits[0] if hits else None


def build(force: bool = False, verbose: bool = False):
    """Returns the path of the built module, or None when /root/reference is absent (GPU box: prebuilt file only)."""
    have = built_path()
    if have and not force:
        return have
    if not os.path.isdir(SRC):
        return have
    os.makedirs(OUT, exist_ok=True)
    os.environ["TORCH_CUDA_ARCH_LIST"] = ""          # as load.py:14 does; the arch comes from the explicit -gencode
    from torch.utils import cpp_extension
    cpp_extension.load(
        name=NAME,
        sources=[os.path.join(SRC, "anti_alias_activation.cpp"), os.path.join(SRC, "anti_alias_activation_cuda.cu")],
        build_directory=OUT,
        extra_cflags=["-O3"],
        extra_cuda_cflags=["-O3", "--use_fast_math", "-gencode", "arch=compute_100a,code=sm_100a",
                           "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__",
                           "--expt-relaxed-constexpr", "--expt-extended-lambda"],
        is_python_module=False,                      # loading it would need nothing more, but keep build != use
        verbose=verbose,
    )
    return built_path()


def load():
    """Import the prebuilt module (tests / aa_bench only).  None when it was never built."""
    path = built_path()
    if path is None:
        return None
    import torch  # noqa: F401  (the module links against libtorch)
    spec = importlib.util.spec_from_file_location(NAME, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose=True)
    print("reference kernel:", p or "not built (no /root/reference)")
