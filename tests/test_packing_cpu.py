"""Host logic (no GPU): weight packing + varlen layout reproduce torch's conv semantics through the
tap-GEMM addressing contract (include/cvb200.h)."""
import pytest
import torch
import torch.nn.functional as F

from cosyvoice_b200.layout import SeqLayout
from cosyvoice_b200.packing import conv1d_plan, conv_transpose1d_plan, linear_plan
from tapgemm_emul import emulate, from_rows, to_rows

torch.manual_seed(0)
DT, DEV = torch.float32, "cpu"


def _cmp(got, want, tol=1e-4):
    for g, w in zip(got, want):
        assert g.shape == w.shape, (g.shape, w.shape)
        assert (g - w).abs().max().item() < tol, (g - w).abs().max().item()


@pytest.mark.parametrize("k,d,causal", [(3, 1, False), (3, 1, True), (7, 1, False), (11, 5, False), (7, 3, False)])
def test_conv1d_same_and_causal(k, d, causal):
    cin, cout, lens = 96, 40, [37, 150, 5]
    w, b = torch.randn(cout, cin, k), torch.randn(cout)
    lay = SeqLayout.build(lens, DEV, halo=32)
    xs = [torch.randn(cin, l) for l in lens]
    plan = conv1d_plan(w, b, DT, DEV, dilation=d, pad_left=(k - 1) * d if causal else None)
    D = emulate(plan, to_rows(xs, lay, 128), lay)
    if causal:
        want = [F.conv1d(F.pad(x[None], ((k - 1) * d, 0)), w, b, dilation=d)[0] for x in xs]
    else:
        want = [F.conv1d(x[None], w, b, dilation=d, padding=(k - 1) * d // 2)[0] for x in xs]
    _cmp(from_rows(D[:, 0], lay, cout), want)


def test_linear_and_column_slice_view():
    lens = [20, 3]
    lay = SeqLayout.build(lens, DEV)
    w, b = torch.randn(24, 64), torch.randn(24)
    xs = [torch.randn(64, l) for l in lens]
    buf = torch.zeros(lay.rows, 128)
    buf[:, 64:] = to_rows(xs, lay, 64)
    plan = linear_plan(w, b, DT, DEV)
    D = emulate(plan, buf[:, 64:], lay)
    _cmp(from_rows(D[:, 0], lay, 24), [F.linear(x.t(), w, b).t() for x in xs])


@pytest.mark.parametrize("u,k,pad,cin", [(2, 3, 1, 64), (8, 16, 4, 18), (15, 30, 7, 18), (3, 6, 1, 18)])
def test_strided_conv_through_grouped_view(u, k, pad, cin):
    """Downsample1D (components/decoder.py:65) and HiFT source_downs (generator.py:290-300)."""
    cout = 32
    lens_out = [9, 21]
    lens_in = [l * u + (1 if u != 2 else 0) for l in lens_out]        # HiFT: F = T*P + 1 frames
    if u == 2:
        lens_in = [2 * l - 1 for l in lens_out]                       # odd T: ceil(T/2) outputs
    lay_out = SeqLayout.build(lens_out, DEV, align=128)
    lay_in = lay_out.derive(u, 1, lens_in)
    w, b = torch.randn(cout, cin, k), torch.randn(cout)
    xs = [torch.randn(cin, l) for l in lens_in]
    plan = conv1d_plan(w, b, DT, DEV, stride=u, pad_left=pad, c_pad=64)
    D = emulate(plan, to_rows(xs, lay_in, 64), lay_out)
    want = [F.conv1d(x[None], w, b, stride=u, padding=pad)[0] for x in xs]
    assert [x.shape[1] for x in want] == lens_out
    _cmp(from_rows(D[:, 0], lay_out, cout), want)


def test_stride2_conv_on_column_slice_with_pitch():
    """v1 Downsample1D reads the skip half of a [h | skip] buffer (col_pitch = 2*C)."""
    Cc, lens = 64, [11, 30]
    lay = SeqLayout.build(lens, DEV, align=256)
    lay2 = lay.derive(1, 2, [(l + 1) // 2 for l in lens])
    w, b = torch.randn(48, Cc, 3), torch.randn(48)
    xs = [torch.randn(Cc, l) for l in lens]
    cat = torch.randn(lay.rows, 2 * Cc)
    cat[:, Cc:] = to_rows(xs, lay, Cc)
    plan = conv1d_plan(w, b, DT, DEV, stride=2, pad_left=1, col_pitch=2 * Cc)
    D = emulate(plan, cat[:, Cc:], lay2, a_cols=3 * Cc)
    _cmp(from_rows(D[:, 0], lay2, 48), [F.conv1d(x[None], w, b, stride=2, padding=1)[0] for x in xs])


@pytest.mark.parametrize("s,k,shift", [(2, 4, 0), (8, 16, 0), (8, 16, 1), (5, 11, 0), (3, 7, 1), (4, 8, 0), (2, 2, 0)])
def test_conv_transpose_phases(s, k, shift):
    """Upsample1D (components/decoder.py:142), HiFT ups (+ ReflectionPad1d((1,0)) via shift), BigVGAN ups."""
    cin, cout, lens = 64, 24, [7, 19]
    pad = (k - s) // 2
    lay_in = SeqLayout.build(lens, DEV)
    out_lens = [l * s + shift for l in lens]
    lay_out = lay_in.derive(s, 1, out_lens)
    w, b = torch.randn(cin, cout, k), torch.randn(cout)
    xs = [torch.randn(cin, l) for l in lens]
    plan = conv_transpose1d_plan(w, b, DT, DEV, stride=s, padding=pad, shift=shift)
    lay_m = lay_in.with_lens([l + (1 if shift else 0) for l in lens])
    D = emulate(plan, to_rows(xs, lay_in, 64), lay_m, out_limit=out_lens)
    out = D.reshape(lay_in.rows * s, cout)                      # [rows, n_phase, N] -> interleaved output rows
    want = [F.conv_transpose1d(x[None], w, b, stride=s, padding=pad)[0] for x in xs]
    got = from_rows(out, lay_out, cout)
    if shift:
        got = [g[:, 1:] for g in got]                           # row 0 is fixed up by reflect_row0
    _cmp(got, want)


def test_upsample_crop_to_skip_length():
    """flow/decoder.py:309: ConvTranspose1d(4,2,1) output cropped to the (odd) skip length."""
    lens_full = [11, 30]
    lay = SeqLayout.build(lens_full, DEV, align=256)
    lay2 = lay.derive(1, 2, [(l + 1) // 2 for l in lens_full])
    w, b = torch.randn(64, 64, 4), torch.randn(64)
    xs = [torch.randn(64, l) for l in lay2.lens]
    plan = conv_transpose1d_plan(w, b, DT, DEV, stride=2, padding=1)
    D = emulate(plan, to_rows(xs, lay2, 64), lay2, out_limit=lens_full)
    got = from_rows(D.reshape(lay.rows, 64), lay, 64)
    want = [F.conv_transpose1d(x[None], w, b, stride=2, padding=1)[0][:, :l] for x, l in zip(xs, lens_full)]
    _cmp(got, want)
    # rows past the crop stay zero
    full = D.reshape(lay.rows, 64)
    assert full[lay.row0[0] + 11].abs().max() == 0
