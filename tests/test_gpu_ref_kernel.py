"""cvb200_anti_alias_act_nct against the reference's OWN native kernel (`fwd_cuda`,
cosyvoice/BigVGAN/alias_free_activation/cuda/anti_alias_activation_cuda.cu:214-256), compiled for sm_100a from the
reference tree by oracle/build_ref.py into oracle/_ref/ (prebuilt in the build container; it travels to the GPU box).

What the comparison shows (and the tolerances).  The reference has TWO implementations of this activation: the torch
module Activation1d (alias_free_activation/torch/act.py:26-31), which is the default (`use_cuda_kernel: False`,
examples/tts_vc/cosyvoice2/conf/cosyvoice_bigvgan_tts.yaml:42) and the one the oracle restates, and the native kernel.
They do not agree with each other: the native kernel differs from the torch module by up to 5e-2 on the first and
last 3 samples of every sequence (its replicate padding of the 2x intermediate), and in bf16 it rounds the 2x
intermediate to bf16 (error up to 8e-2 anywhere).  cvb200_anti_alias_act_nct follows the torch module everywhere:
  fp32: |ours - fp64 oracle| <= 2e-5 everywhere; |ours - native| <= 2e-5 on the interior samples [3, T-3);
  bf16: ours within half a bf16 ulp of the fp64 oracle (only the final rounding) and never further from it than the
        native kernel is.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

from cosyvoice_b200 import ops
from oracle import bigvgan_ref, build_ref

DEV = "cuda"
_CACHE = []


def _ref():
    if not _CACHE:
        _CACHE.append(build_ref.load())
    if _CACHE[0] is None:
        pytest.skip("oracle/_ref/anti_alias_activation_cuda.so not built (python -m oracle.build_ref in the build container)")
    return _CACHE[0]


@pytest.mark.parametrize("shape", [(2, 5, 700), (1, 24, 4096 + 37), (3, 96, 1000), (1, 1, 1), (2, 7, 11)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_matches_reference_native_kernel(shape, dtype):
    B, Cc, T = shape
    g = torch.Generator().manual_seed(B * 1000 + Cc * 10 + T)
    x = (torch.randn(B, Cc, T, generator=g) * 2).to(dtype)
    al, be = torch.randn(Cc, generator=g) * 0.3, torch.randn(Cc, generator=g) * 0.3
    filt = bigvgan_ref.kaiser_sinc_filter()
    xd, fd, ad, bd = x.to(DEV), filt.to(DEV), al.to(DEV), be.to(DEV)
    ref = _ref().forward(xd, fd, fd, ad, bd)
    ours = ops.anti_alias_act_nct(xd, fd, fd, ad, bd)
    torch.cuda.synchronize()
    assert ours.dtype == ref.dtype and ours.shape == ref.shape
    truth = bigvgan_ref.anti_alias_act(x.double(), al.double(), be.double(), filt.double())
    e_ref = (ref.double().cpu() - truth).abs().max().item()
    e_ours = (ours.double().cpu() - truth).abs().max().item()
    diff = (ours.float() - ref.float()).abs().max().item()
    if dtype == torch.float32:
        assert e_ours <= 2e-5, e_ours
        if T > 6:
            interior = (ours - ref)[..., 3:T - 3].abs().max().item()
            assert interior <= 2e-5, interior
    else:
        half_ulp = truth.abs().max().item() * 2.0 ** -8 * 1.05 + 1e-6
        assert e_ours <= half_ulp and e_ours <= e_ref + 1e-6, (e_ours, e_ref, half_ulp)
    del diff


def test_empty_sequence_is_a_no_op_in_both():
    filt = bigvgan_ref.kaiser_sinc_filter().to(DEV)
    a = torch.zeros(4, device=DEV)
    x = torch.zeros(1, 4, 0, device=DEV)
    assert _ref().forward(x, filt, filt, a, a).shape == ops.anti_alias_act_nct(x, filt, filt, a, a).shape == (1, 4, 0)
