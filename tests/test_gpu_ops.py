"""GPU parity of the individual kernels (through the C ABI) against torch fp32 references."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from cosyvoice_b200 import _lib, ops
from cosyvoice_b200.layout import SeqLayout
from cosyvoice_b200.packing import conv1d_plan, conv_transpose1d_plan, linear_plan
from tapgemm_emul import emulate, from_rows, to_rows

DEV = "cuda"


def _sync():
    torch.cuda.synchronize()


def _rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def _run_gemm(plan, A, lay, N_out_cols, engine, out_dtype=torch.float32, **kw):
    out = torch.zeros(lay.rows * plan.n_phase, N_out_cols, dtype=out_dtype, device=DEV)
    ops.gemm_op(plan, A, lay, out, engine=engine, **kw)()
    _sync()
    return out


@pytest.mark.parametrize("case", ["linear", "conv_k3", "conv_k11_d5", "causal", "stride8", "convT_s8", "convT_s2"])
def test_tapgemm_simt_fp32_matches_emulation(case):
    torch.manual_seed(1)
    lens = [37, 150, 5]
    cin, cout = 128, 80
    lay = SeqLayout.build(lens, DEV, halo=32)
    layc = SeqLayout.build(lens, "cpu", halo=32)
    if case == "linear":
        plan_c = linear_plan(torch.randn(cout, cin), torch.randn(cout), torch.float32, "cpu")
    elif case == "conv_k3":
        plan_c = conv1d_plan(torch.randn(cout, cin, 3), torch.randn(cout), torch.float32, "cpu")
    elif case == "conv_k11_d5":
        plan_c = conv1d_plan(torch.randn(cout, cin, 11), torch.randn(cout), torch.float32, "cpu", dilation=5)
    elif case == "causal":
        plan_c = conv1d_plan(torch.randn(cout, cin, 3), torch.randn(cout), torch.float32, "cpu", pad_left=2)
    elif case == "stride8":
        plan_c = conv1d_plan(torch.randn(cout, 18, 16), torch.randn(cout), torch.float32, "cpu", stride=8, pad_left=4, c_pad=64)
    elif case == "convT_s8":
        plan_c = conv_transpose1d_plan(torch.randn(cin, cout, 16), torch.randn(cout), torch.float32, "cpu", stride=8, padding=4)
    else:
        plan_c = conv_transpose1d_plan(torch.randn(cin, cout, 4), torch.randn(cout), torch.float32, "cpu", stride=2, padding=1)
    import copy
    plan_g = copy.copy(plan_c)
    plan_g.W, plan_g.bias = plan_c.W.to(DEV), plan_c.bias.to(DEV)
    if case == "stride8":
        lay_in_c = layc.derive(8, 1, [l * 8 + 1 for l in lens])
        A = torch.randn(lay_in_c.rows, 64)
        for b in range(3):   # zero the padding rows like every producer does
            A[lay_in_c.row0[b] + lay_in_c.lens[b]:lay_in_c.row0[b] + lay_in_c.strides[b]] = 0
    else:
        A = to_rows([torch.randn(cin, l) for l in lens], layc, cin)
    want = emulate(plan_c, A, layc)
    got = _run_gemm(plan_g, A.to(DEV), lay, cout, _lib.ENGINE_SIMT)
    got = got.reshape(lay.rows, plan_c.n_phase, cout).cpu()
    assert (got - want).abs().max().item() < 2e-3 * max(1.0, want.abs().max().item())


@pytest.mark.parametrize("case", ["linear_256_1536", "linear_1024_256", "conv_k3_320", "conv_k11_d5_128", "convT_s8",
                                  "final_proj_80", "conv_post_18", "stride8"])
def test_tapgemm_tcgen05_matches_simt(case):
    """Same bf16 operands through both engines: the tensor-core path must agree with the FFMA path to fp32
    accumulation-order noise."""
    torch.manual_seed(2)
    lens = [300, 129, 5, 640]
    bf = torch.bfloat16
    lay = SeqLayout.build(lens, DEV, halo=32)
    kw = {}
    cin = 256
    if case == "linear_256_1536":
        plan = linear_plan(torch.randn(1536, 256) * 0.06, None, bf, DEV)
    elif case == "linear_1024_256":
        cin = 1024
        plan = linear_plan(torch.randn(256, 1024) * 0.03, torch.randn(256), bf, DEV)
        kw = dict(act=_lib.ACT_GELU)
    elif case == "conv_k3_320":
        cin = 320
        plan = conv1d_plan(torch.randn(256, 320, 3) * 0.03, torch.randn(256), bf, DEV)
    elif case == "conv_k11_d5_128":
        cin = 128
        plan = conv1d_plan(torch.randn(128, 128, 11) * 0.03, torch.randn(128), bf, DEV, dilation=5)
    elif case == "convT_s8":
        plan = conv_transpose1d_plan(torch.randn(256, 128, 16) * 0.03, torch.randn(128), bf, DEV, stride=8, padding=4)
    elif case == "final_proj_80":
        plan = linear_plan(torch.randn(80, 256) * 0.06, torch.randn(80), bf, DEV)
    elif case == "conv_post_18":
        cin = 128
        plan = conv1d_plan(torch.randn(18, 128, 7) * 0.03, torch.randn(18), bf, DEV)
    else:
        cin = 64
        plan = conv1d_plan(torch.randn(256, 18, 16) * 0.1, torch.randn(256), bf, DEV, stride=8, pad_left=4, c_pad=64)
    if case == "stride8":
        lay_in = lay.derive(8, 1, [l * 8 + 1 for l in lens])
        A = torch.randn(lay_in.rows, 64, device=DEV)
        for b in range(len(lens)):
            A[lay_in.row0[b] + lay_in.lens[b]:lay_in.row0[b] + lay_in.strides[b]] = 0
        A[:, 18:] = 0
    else:
        A = to_rows([torch.randn(cin, l) for l in lens], SeqLayout.build(lens, "cpu", halo=32), cin).to(DEV)
    A = A.to(bf)
    ncol = 32 if plan.N == 18 else plan.N
    ref = _run_gemm(plan, A, lay, ncol, _lib.ENGINE_SIMT, **kw)
    got = _run_gemm(plan, A, lay, ncol, _lib.ENGINE_TCGEN05, **kw)
    err = (got - ref).abs().max().item()
    assert err < 2e-3 * max(1.0, ref.abs().max().item()), err
    assert ref.abs().max().item() > 0.1


@pytest.mark.parametrize("case", ["linear_256_1536", "linear_1024_256", "conv_k3_320", "conv_k11_d5_128", "convT_s8",
                                  "final_proj_80", "conv_post_18", "stride8"])
def test_tapgemm_tcgen05_matches_torch_fp32(case):
    """The eight tcgen05 shapes against torch itself (F.linear / F.conv1d / F.conv_transpose1d in fp32 on the bf16-rounded
    weights and inputs, one call per sequence on the un-packed [C, T] tensors): the reference never sees packing.py's
    tap / phase / segment layout, so a packing bug shared by both engines cannot hide.
    Tolerance: fp32 accumulation order over K <= 3520 products of bf16 values, 2e-3 of the output scale."""
    torch.manual_seed(2)
    lens = [300, 129, 5, 640]
    bf = torch.bfloat16
    lay = SeqLayout.build(lens, DEV, halo=32)
    r = lambda *sh, s=1.0: (torch.randn(*sh) * s).to(bf).float()          # bf16-rounded fp32 master copies
    kw, cin, n_phase = {}, 256, 1
    if case == "linear_256_1536":
        W = r(1536, 256, s=0.06)
        plan, ref = linear_plan(W, None, bf, DEV), (lambda x: F.linear(x.t(), W).t())
    elif case == "linear_1024_256":
        cin, W, b = 1024, r(256, 1024, s=0.03), torch.randn(256)
        plan, ref = linear_plan(W, b, bf, DEV), (lambda x: F.linear(x.t(), W, b).t())
    elif case == "conv_k3_320":
        cin, W, b = 320, r(256, 320, 3, s=0.03), torch.randn(256)
        plan, ref = conv1d_plan(W, b, bf, DEV), (lambda x: F.conv1d(x[None], W, b, padding=1)[0])
    elif case == "conv_k11_d5_128":
        cin, W, b = 128, r(128, 128, 11, s=0.03), torch.randn(128)
        plan, ref = conv1d_plan(W, b, bf, DEV, dilation=5), (lambda x: F.conv1d(x[None], W, b, padding=25, dilation=5)[0])
    elif case == "convT_s8":
        W, b, n_phase = r(256, 128, 16, s=0.03), torch.randn(128), 8
        plan = conv_transpose1d_plan(W, b, bf, DEV, stride=8, padding=4)
        ref = lambda x: F.conv_transpose1d(x[None], W, b, stride=8, padding=4)[0]
    elif case == "final_proj_80":
        W, b = r(80, 256, s=0.06), torch.randn(80)
        plan, ref = linear_plan(W, b, bf, DEV), (lambda x: F.linear(x.t(), W, b).t())
    elif case == "conv_post_18":
        cin, W, b = 128, r(18, 128, 7, s=0.03), torch.randn(18)
        plan, ref = conv1d_plan(W, b, bf, DEV), (lambda x: F.conv1d(x[None], W, b, padding=3)[0])
    else:       # strided analysis conv (source_downs: kernel 16, stride 8, 4 zeros on the left), 18 channels padded to 64
        cin, W, b = 18, r(256, 18, 16, s=0.1), torch.randn(256)
        plan = conv1d_plan(W, b, bf, DEV, stride=8, pad_left=4, c_pad=64)
        ref = lambda x: F.conv1d(F.pad(x, (4, 16))[None], W, b, stride=8)[0]
    layc = SeqLayout.build(lens, "cpu", halo=32)
    if case == "stride8":
        lay_in = layc.derive(8, 1, [l * 8 + 1 for l in lens])
        xs = [r(cin, l * 8 + 1) for l in lens]
        A = to_rows(xs, lay_in, 64)
    else:
        xs = [r(cin, l) for l in lens]
        A = to_rows(xs, layc, cin)
    ncol = 32 if plan.N == 18 else plan.N
    got = _run_gemm(plan, A.to(DEV).to(bf), lay, ncol, _lib.ENGINE_TCGEN05, **kw).cpu()
    got = got.reshape(lay.rows, n_phase, ncol)
    for bi, l in enumerate(lens):
        want = ref(xs[bi])[:, :l * n_phase]                                     # [N, l * n_phase]
        have = got[lay.row0[bi]:lay.row0[bi] + l, :, :plan.N].reshape(l * n_phase, plan.N).t()
        err = (have - want).abs().max().item()
        assert err < 2e-3 * max(1.0, want.abs().max().item()), (case, bi, err)
        assert want.abs().max().item() > 0.1


def test_tapgemm_epilogue_features_tcgen05():
    """bias + residual + scale + accumulate + dual output (snake) + GroupNorm statistics, bf16 engine vs torch."""
    torch.manual_seed(3)
    bf = torch.bfloat16
    lens = [200, 77]
    lay = SeqLayout.build(lens, DEV)
    layc = SeqLayout.build(lens, "cpu")
    w, b = torch.randn(256, 256, 3) * 0.04, torch.randn(256)
    plan = conv1d_plan(w, b, bf, DEV)
    xs = [torch.randn(256, l) for l in lens]
    A = to_rows(xs, layc, 256).to(DEV).to(bf)
    res = torch.randn(lay.rows, 256, device=DEV)
    out0 = torch.randn(lay.rows, 256, device=DEV)
    prev = out0.clone()
    out1 = torch.zeros(lay.rows, 256, device=DEV, dtype=bf)
    alpha = (1 + 0.3 * torch.randn(256)).to(DEV)
    stats = torch.zeros(2, 8, 2, dtype=torch.float64, device=DEV)
    ops.gemm_op(plan, A, lay, out0, engine=_lib.ENGINE_TCGEN05, res=res, out_scale=0.5, accumulate=True, out1=out1,
                out1_mode=_lib.OUT1_SNAKE, out1_alpha=alpha, gn_stats=stats)()
    _sync()
    for bi, l in enumerate(lens):
        r0 = lay.row0[bi]
        xa = A[r0:r0 + l].float().t()[None]
        conv = F.conv1d(xa, plan.W[0].float().reshape(256, 3, 256).permute(0, 2, 1).contiguous(), b.to(DEV), padding=1)[0].t()
        want = (conv + res[r0:r0 + l]) * 0.5 + prev[r0:r0 + l]
        assert (out0[r0:r0 + l] - want).abs().max().item() < 5e-3
        sn = want + torch.sin(want * alpha) ** 2 / (alpha + 1e-9)
        assert (out1[r0:r0 + l].float() - sn).abs().max().item() < 0.05
        assert out0[r0 + l:r0 + lay.strides[bi]].abs().max().item() == 0
        g = want.reshape(l, 8, 32)
        s_want = g.sum(dim=(0, 2)).double()
        q_want = (g * g).sum(dim=(0, 2)).double()
        assert (stats[bi, :, 0] - s_want).abs().max().item() < 1e-2 * max(1.0, s_want.abs().max().item())
        assert (stats[bi, :, 1] - q_want).abs().max().item() < 1e-3 * q_want.abs().max().item()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_layernorm_and_norm_mish(dtype):
    torch.manual_seed(4)
    lens = [50, 131]
    lay = SeqLayout.build(lens, DEV)
    X = torch.randn(lay.rows, 256, device=DEV) * 2 + 0.5
    g, b = (1 + 0.1 * torch.randn(256)).to(DEV), (0.1 * torch.randn(256)).to(DEV)
    out = torch.zeros(lay.rows, 256, device=DEV, dtype=dtype)
    ops.layernorm_op(X, out, g, b, lay.rows)()
    _sync()
    want = F.layer_norm(X, (256,), g, b)
    tol = 1e-4 if dtype == torch.float32 else 3e-2
    assert (out.float() - want).abs().max().item() < tol
    if dtype == torch.bfloat16:
        # cvb200_layernorm_act: the same norm over a residual stream stored in the operand dtype (exact on the same bf16 values)
        Xb = X.to(dtype)
        out32 = torch.zeros(lay.rows, 256, device=DEV)
        ops.layernorm_op(Xb, out32, g, b, lay.rows)()
        _sync()
        assert (out32 - F.layer_norm(Xb.float(), (256,), g, b)).abs().max().item() < 1e-4
    # per-row LayerNorm + Mish + time bias (causal Block1D)
    h = X.to(dtype)
    tb = torch.randn(2, 512, device=DEV)[:, 256:]
    ops.norm_mish_op(h, out, g, b, lay, seq_bias=tb)()
    _sync()
    for bi, l in enumerate(lens):
        r0 = lay.row0[bi]
        w_ = F.mish(F.layer_norm(h[r0:r0 + l].float(), (256,), g, b)) + tb[bi]
        assert (out[r0:r0 + l].float() - w_).abs().max().item() < tol
        assert out[r0 + l:r0 + lay.strides[bi]].abs().max().item() == 0
    # GroupNorm(8) variant with statistics supplied
    stats = torch.zeros(2, 8, 2, dtype=torch.float64, device=DEV)
    for bi, l in enumerate(lens):
        r0 = lay.row0[bi]
        gq = h[r0:r0 + l].double().reshape(l, 8, 32)
        stats[bi, :, 0] = gq.sum(dim=(0, 2))
        stats[bi, :, 1] = (gq * gq).sum(dim=(0, 2))
    ops.norm_mish_op(h, out, g, b, lay, gn_stats=stats)()
    _sync()
    for bi, l in enumerate(lens):
        r0 = lay.row0[bi]
        w_ = F.mish(F.group_norm(h[r0:r0 + l].float().t()[None], 8, g, b))[0].t()
        assert (out[r0:r0 + l].float() - w_).abs().max().item() < tol


@pytest.mark.parametrize("dtype,lens,impl", [(torch.float32, [70, 200], 0),
                                             (torch.bfloat16, [70, 200, 1, 64, 128, 650], _lib.ATTN_MMA_SYNC),
                                             (torch.bfloat16, [70, 200, 1, 64, 128, 650, 129, 1218], _lib.ATTN_TCGEN05),
                                             (torch.bfloat16, [70, 200, 1, 64, 128, 650, 129, 1218], _lib.ATTN_TCGEN05_V2),
                                             (torch.bfloat16, [609] * 64, _lib.ATTN_TCGEN05_V2),
                                             (torch.bfloat16, [300], _lib.ATTN_TCGEN05_V2)])
def test_attention(dtype, lens, impl):
    torch.manual_seed(5)
    lay = SeqLayout.build(lens, DEV)
    qkv = (torch.randn(lay.rows, 1536, device=DEV) * 1.5).to(dtype)
    out = torch.zeros(lay.rows, 512, device=DEV, dtype=dtype)
    ops.attention_op(qkv, out, lay, 8, 0.125, impl=impl)()
    _sync()
    for bi, l in enumerate(lens):
        r0 = lay.row0[bi]
        x = qkv[r0:r0 + l].float()
        q, k, v = (x[:, i * 512:(i + 1) * 512].reshape(l, 8, 64).transpose(0, 1) for i in range(3))
        want = torch.softmax(q @ k.transpose(1, 2) * 0.125, -1) @ v
        want = want.transpose(0, 1).reshape(l, 512)
        err = (out[r0:r0 + l].float() - want).abs().max().item()
        assert err < (1e-4 if dtype == torch.float32 else 3e-2), (bi, l, err)


def test_time_embedding_path():
    torch.manual_seed(6)
    t = torch.tensor([0.0, 0.1, 0.7, 1.0], device=DEV)
    out = torch.zeros(4, 320, device=DEV)
    ops.time_sinusoid(t, out)
    half = 160
    k = math.log(10000) / (half - 1)
    e = 1000.0 * t[:, None] * torch.exp(torch.arange(half, device=DEV).float() * -k)[None]
    want = torch.cat([e.sin(), e.cos()], -1)
    _sync()
    assert (out - want).abs().max().item() < 2e-3        # sin/cos of arguments up to 1000 rad
    W, b = torch.randn(1024, 320, device=DEV) * 0.05, torch.randn(1024, device=DEV)
    y = torch.zeros(4, 1024, device=DEV)
    ops.small_linear(want.contiguous(), W, b, y, post_act=_lib.ACT_SILU)
    _sync()
    assert (y - F.silu(F.linear(want, W, b))).abs().max().item() < 1e-4
    W2 = torch.randn(512, 1024, device=DEV) * 0.03
    y2 = torch.zeros(4, 512, device=DEV)
    ops.small_linear(y, W2, None, y2, pre_act=_lib.ACT_MISH)
    _sync()
    assert (y2 - F.linear(F.mish(y), W2)).abs().max().item() < 1e-4


def test_pack_cfg_and_euler():
    torch.manual_seed(7)
    lens = [33, 140]
    lay = SeqLayout.build(lens, DEV)
    R = lay.rows
    x, mu, cond = (torch.randn(R, 80, device=DEV) for _ in range(3))
    spks = torch.randn(2, 80, device=DEV)
    out = torch.full((2 * R, 320), 7.0, device=DEV)
    ops.pack_cfg_input(x, mu, cond, spks, out, lay)
    _sync()
    for bi, l in enumerate(lens):
        r0 = lay.row0[bi]
        want_c = torch.cat([x[r0:r0 + l], mu[r0:r0 + l], spks[bi][None].expand(l, -1), cond[r0:r0 + l]], 1)
        assert torch.equal(out[r0:r0 + l], want_c)
        want_u = torch.cat([x[r0:r0 + l], torch.zeros(l, 240, device=DEV)], 1)
        assert torch.equal(out[R + r0:R + r0 + l], want_u)
        assert out[r0 + l:r0 + lay.strides[bi]].abs().max().item() == 0
    v = torch.randn(2 * R, 80, device=DEV)
    x0 = x.clone()
    ops.euler_cfg(x, v, 0.05, 0.7)
    _sync()
    want = x0 + 0.05 * (1.7 * v[:R] - 0.7 * v[R:])
    assert (x - want).abs().max().item() < 1e-5


def test_layout_transposes_roundtrip():
    torch.manual_seed(8)
    lens = [45, 300, 7]
    lay = SeqLayout.build(lens, DEV)
    src = torch.randn(3, 80, 300, device=DEV)
    rows = torch.zeros(lay.rows, 80, device=DEV)
    ops.nct_to_rows(src, rows, lay)
    back = torch.zeros_like(src)
    ops.rows_to_nct(rows, back, lay)
    _sync()
    for bi, l in enumerate(lens):
        assert torch.equal(back[bi, :, :l], src[bi, :, :l])
        assert torch.equal(rows[lay.row0[bi]:lay.row0[bi] + l], src[bi, :, :l].t())
        assert rows[lay.row0[bi] + l:lay.row0[bi] + lay.strides[bi]].abs().max().item() == 0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_anti_alias_activation_nct_matches_reference_golden(golden_dir, dtype):
    """Drop-in for the reference's fwd_cuda: checked against the output of the reference's torch Activation1d."""
    import os
    d = torch.load(os.path.join(golden_dir, "act1d.pt"), weights_only=True)
    y = ops.anti_alias_act_nct(d["x"].to(DEV).to(dtype), d["up_filter"].to(DEV), d["down_filter"].to(DEV),
                               d["alpha"].to(DEV), d["beta"].to(DEV))
    _sync()
    tol = 5e-6 if dtype == torch.float32 else 5e-2
    assert (y.float().cpu() - d["y"]).abs().max().item() < tol
    # longer, strip-crossing rows against the oracle
    from oracle import bigvgan_ref
    g = torch.Generator().manual_seed(9)
    x = torch.randn(2, 5, 700, generator=g) * 2
    al, be = torch.randn(5, generator=g) * 0.3, torch.randn(5, generator=g) * 0.3
    filt = bigvgan_ref.kaiser_sinc_filter()
    want = bigvgan_ref.anti_alias_act(x, al, be, filt)
    y = ops.anti_alias_act_nct(x.to(DEV).to(dtype), filt.to(DEV), filt.to(DEV), al.to(DEV), be.to(DEV))
    assert (y.float().cpu() - want).abs().max().item() < (1e-5 if dtype == torch.float32 else 6e-2)
    # empty input is a silent no-op like the reference (.cu:195-198)
    e = ops.anti_alias_act_nct(torch.zeros(1, 4, 0, device=DEV), filt.to(DEV), filt.to(DEV), al[:4].to(DEV), be[:4].to(DEV))
    assert e.shape == (1, 4, 0)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("T", [4096, 4097, 8191, 12289, 5])
def test_anti_alias_nct_vector_staging_any_alignment(dtype, T):
    """The [B,C,T] kernel stages with 16-byte accesses aligned to the global address: odd T (rows start at odd element
    offsets), chunk boundaries at 4096 and a base pointer that is itself only element-aligned."""
    from oracle import bigvgan_ref
    g = torch.Generator().manual_seed(T)
    B, Cc = 2, 3
    x = (torch.randn(B, Cc, T, generator=g) * 2).to(dtype)
    al, be = torch.randn(Cc, generator=g) * 0.3, torch.randn(Cc, generator=g) * 0.3
    filt = bigvgan_ref.kaiser_sinc_filter()
    want = bigvgan_ref.anti_alias_act(x.double(), al.double(), be.double(), filt.double())
    tol = 2e-5 if dtype == torch.float32 else want.abs().max().item() * 2.0 ** -8 * 1.05 + 1e-6
    for shift in (0, 1, 3):
        flat = torch.zeros(B * Cc * T + shift + 16, dtype=dtype, device=DEV)
        xv = flat[shift:shift + B * Cc * T].view(B, Cc, T)
        xv.copy_(x.to(DEV))
        assert xv.is_contiguous() and xv.data_ptr() == flat.data_ptr() + shift * flat.element_size()
        y = ops.anti_alias_act_nct(xv, filt.to(DEV), filt.to(DEV), al.to(DEV), be.to(DEV))
        _sync()
        assert (y.double().cpu() - want).abs().max().item() <= tol, (shift, T)
        assert flat[shift + B * Cc * T:].abs().max().item() == 0 and (shift == 0 or flat[:shift].abs().max().item() == 0)


def test_anti_alias_activation_rows_varlen():
    from oracle import bigvgan_ref
    g = torch.Generator().manual_seed(10)
    lens, Cc = [150, 64, 3], 96
    lay = SeqLayout.build(lens, DEV)
    layc = SeqLayout.build(lens, "cpu")
    xs = [torch.randn(Cc, l, generator=g) * 2 for l in lens]
    al, be = torch.randn(Cc, generator=g) * 0.3, torch.randn(Cc, generator=g) * 0.3
    filt = bigvgan_ref.kaiser_sinc_filter()
    A = to_rows(xs, layc, Cc).to(DEV)
    out = torch.full((lay.rows, Cc), 5.0, device=DEV)
    ops.anti_alias_act_rows(A, out, filt.to(DEV), al.to(DEV), be.to(DEV), Cc, lay)
    _sync()
    got = from_rows(out.cpu(), layc, Cc)
    for b, x in enumerate(xs):
        want = bigvgan_ref.anti_alias_act(x[None], al, be, filt)[0]
        assert (got[b] - want).abs().max().item() < 1e-5, b
        assert out[lay.row0[b] + lens[b]:lay.row0[b] + lay.strides[b]].abs().max().item() == 0


def test_snake_and_reflect():
    torch.manual_seed(11)
    lens = [40, 7]
    lay = SeqLayout.build(lens, DEV)
    x = torch.randn(lay.rows, 64, device=DEV)
    al = (1 + 0.3 * torch.randn(64)).to(DEV)
    out = torch.zeros(lay.rows, 64, device=DEV, dtype=torch.bfloat16)
    ops.snake(x, out, al, 64, lay)
    _sync()
    for b, l in enumerate(lens):
        r0 = lay.row0[b]
        want = x[r0:r0 + l] + torch.sin(x[r0:r0 + l] * al) ** 2 / (al + 1e-9)
        assert (out[r0:r0 + l].float() - want).abs().max().item() < 3e-2
        assert out[r0 + l:r0 + lay.strides[b]].abs().max().item() == 0
    ops.reflect_row0(x, 64, lay)
    _sync()
    for b in range(2):
        assert torch.equal(x[lay.row0[b]], x[lay.row0[b] + 2])


@pytest.mark.parametrize("lens", [[200, 77], [128], [1000, 3, 640, 129]])
def test_ffn_fused_matches_torch(lens):
    """cvb200_ffn_fused (hidden activations kept in TMEM) against torch on the same bf16-rounded operands:
    x += W2 . bf16(gelu(W1 . a + b1)) + b2.  Tolerance = bf16 operand rounding of a K=1024 reduction."""
    torch.manual_seed(31)
    lay = SeqLayout.build(lens, DEV)
    R = lay.rows
    a = (torch.randn(R, 256, device=DEV)).to(torch.bfloat16)
    w1 = (torch.randn(1024, 256, device=DEV) * (2.0 / 256) ** 0.5).to(torch.bfloat16)
    w2 = (torch.randn(256, 1024, device=DEV) * (2.0 / 1024) ** 0.5).to(torch.bfloat16)
    b1 = torch.randn(1024, device=DEV) * 0.1
    b2 = torch.randn(256, device=DEV) * 0.1
    x0 = torch.randn(R, 256, device=DEV)
    x = x0.clone()
    out1 = torch.full((R, 256), 7.0, device=DEV, dtype=torch.bfloat16)
    ops.ffn_fused_op(a, w1, b1, w2, b2, x, lay, out1=out1)()
    _sync()
    h = torch.nn.functional.gelu(a.float() @ w1.float().t() + b1).to(torch.bfloat16).float()
    want = x0 + h @ w2.float().t() + b2
    for b, l in enumerate(lens):
        r0 = lay.row0[b]
        got = x[r0:r0 + l]
        err = (got - want[r0:r0 + l]).abs().max().item()
        assert err < 3e-2, (b, l, err)
        assert (out1[r0:r0 + l].float() - got).abs().max().item() <= 2e-2 * max(1.0, got.abs().max().item())
        assert x[r0 + l:r0 + lay.strides[b]].abs().max().item() == 0
        assert out1[r0 + l:r0 + lay.strides[b]].float().abs().max().item() == 0
    # second launch on the same buffers (barrier phases / TMEM reuse across launches)
    x2 = x0.clone()
    ops.ffn_fused_op(a, w1, b1, w2, b2, x2, lay)()
    _sync()
    assert torch.equal(x2, x)


@pytest.mark.parametrize("lens", [[200, 77], [128], [1000, 3, 640, 129], [609] * 40])
def test_block_ffn_matches_torch(lens):
    """cvb200_block_ffn (norm3 + FeedForward + residual in one tcgen05 kernel on the bf16 stream, LayerNorm folded into the
    first GEMM, row statistics of the new stream as a side output) against torch:
        x' = bf16(x + W2 . bf16(gelu(W1 . LN(x) + b1)) + b2),  stats = (rstd, -mean rstd) of x'.
    Tolerances = bf16 rounding of the operands (K = 1024 reduction) and of the output.  The rows carry a common offset of
    up to 3 sigma (the folded form must cancel the mean term).  The last case has more tiles than SMs (several tiles per CTA:
    barrier phases, X-tile prefetch, TMEM reuse across tiles)."""
    torch.manual_seed(41)
    lay = SeqLayout.build(lens, DEV)
    R = lay.rows
    x0 = torch.zeros(R, 256, device=DEV)
    for b, l in enumerate(lens):
        x0[lay.row0[b]:lay.row0[b] + l] = torch.randn(l, 256, device=DEV) * 1.5 + 3 * torch.randn(l, 1, device=DEV)
    x0 = x0.to(torch.bfloat16)
    w1 = torch.randn(1024, 256, device=DEV) * (2.0 / 256) ** 0.5
    w2 = (torch.randn(256, 1024, device=DEV) * (2.0 / 1024) ** 0.5).to(torch.bfloat16)
    b1, b2 = torch.randn(1024, device=DEV) * 0.1, torch.randn(256, device=DEV) * 0.1
    g3, be3 = 1 + 0.2 * torch.randn(256, device=DEV), 0.1 * torch.randn(256, device=DEV)
    F = torch.nn.functional
    w1f, u1, b1f = ops.fold_layernorm(w1, b1, g3, be3, torch.bfloat16)
    # the folded operands reproduce the unfolded Linear(LayerNorm(x)) up to the bf16 rounding of W1 * gamma
    a = F.layer_norm(x0.float(), (256,))
    pre = a @ w1f.float().t() + b1f
    pre_unfolded = F.layer_norm(x0.float(), (256,), g3, be3) @ w1.t() + b1
    assert (pre - pre_unfolded).abs().max().item() < 0.05
    h = F.gelu(pre).to(torch.bfloat16).float()
    want_x = x0.float() + h @ w2.float().t() + b2
    x = x0.clone()
    out1 = torch.full((R, 256), 7.0, device=DEV, dtype=torch.bfloat16)
    stats = torch.full((R, 2), 7.0, device=DEV)
    ops.block_ffn_op(x, w1f, u1, b1f, w2, b2, lay, out1=out1, stats_out=stats)()
    _sync()
    mean = x.float().mean(dim=1)
    rstd = torch.rsqrt(x.float().var(dim=1, unbiased=False) + 1e-5)
    for b, l in enumerate(lens):
        r0 = lay.row0[b]
        got = x[r0:r0 + l].float()
        scale = max(1.0, want_x[r0:r0 + l].abs().max().item())
        err = (got - want_x[r0:r0 + l]).abs().max().item()
        assert err < 2e-2 * scale, (b, l, err, scale)
        assert torch.equal(out1[r0:r0 + l], x[r0:r0 + l])
        assert torch.allclose(stats[r0:r0 + l, 0], rstd[r0:r0 + l], rtol=2e-4, atol=1e-6), (b, l)
        assert torch.allclose(stats[r0:r0 + l, 1], -mean[r0:r0 + l] * rstd[r0:r0 + l], rtol=2e-4, atol=2e-4), (b, l)
        for buf in (x, out1, stats):
            assert buf[r0 + l:r0 + lay.strides[b]].float().abs().max().item() == 0
    # without the optional outputs, and a second launch on the same buffers: identical stream
    x2 = x0.clone()
    ops.block_ffn_op(x2, w1f, u1, b1f, w2, b2, lay)()
    _sync()
    assert torch.equal(x2, x)


@pytest.mark.parametrize("lens", [[200, 77], [128], [1000, 3, 640, 129], [609] * 40])
def test_block_out_ffn_equals_out_projection_then_block_ffn(lens):
    """cvb200_block_out_ffn (attention out-projection + residual + norm3 + FeedForward + residual in ONE kernel) against the two
    launches it replaces -- cvb200_tapgemm (out-projection, bias, bf16 residual) then cvb200_block_ffn: BIT-identical stream,
    copy and statistics (same K order in the accumulator, same epilogue arithmetic) -- and against torch on the same
    bf16-rounded operands.  Rows past the valid lengths come out as zeros."""
    torch.manual_seed(43)
    lay = SeqLayout.build(lens, DEV)
    R = lay.rows
    x0 = torch.zeros(R, 256, device=DEV)
    o = torch.zeros(R, 512, device=DEV)
    for b, l in enumerate(lens):
        x0[lay.row0[b]:lay.row0[b] + l] = torch.randn(l, 256, device=DEV) * 1.5 + 3 * torch.randn(l, 1, device=DEV)
        o[lay.row0[b]:lay.row0[b] + l] = torch.randn(l, 512, device=DEV)
    o[lay.row0[0] + lens[0]:lay.row0[0] + lay.strides[0]] = 5.0          # garbage in pad rows of the attention output must not leak
    x0, o = x0.to(torch.bfloat16), o.to(torch.bfloat16)
    wo = (torch.randn(256, 512, device=DEV) * (1.0 / 512) ** 0.5).to(torch.bfloat16)
    bo = torch.randn(256, device=DEV) * 0.1
    w1 = torch.randn(1024, 256, device=DEV) * (2.0 / 256) ** 0.5
    w2 = (torch.randn(256, 1024, device=DEV) * (2.0 / 1024) ** 0.5).to(torch.bfloat16)
    b1, b2 = torch.randn(1024, device=DEV) * 0.1, torch.randn(256, device=DEV) * 0.1
    g3, be3 = 1 + 0.2 * torch.randn(256, device=DEV), 0.1 * torch.randn(256, device=DEV)
    w1f, u1, b1f = ops.fold_layernorm(w1, b1, g3, be3, torch.bfloat16)
    # the two-launch form
    plan = linear_plan(wo.float(), bo, torch.bfloat16, DEV)
    xa = x0.clone()
    ops.gemm_op(plan, o, lay, xa, engine=_lib.ENGINE_TCGEN05, res=xa)()
    x_mid = xa.clone()
    out1a, sta = torch.full((R, 256), 7.0, device=DEV, dtype=torch.bfloat16), torch.full((R, 2), 7.0, device=DEV)
    ops.block_ffn_op(xa, w1f, u1, b1f, w2, b2, lay, out1=out1a, stats_out=sta)()
    # the fused form
    xb = x0.clone()
    out1b, stb = torch.full((R, 256), 7.0, device=DEV, dtype=torch.bfloat16), torch.full((R, 2), 7.0, device=DEV)
    ops.block_ffn_op(xb, w1f, u1, b1f, w2, b2, lay, out1=out1b, stats_out=stb, attn_out=o, w_out=wo, b_out=bo)()
    _sync()
    assert torch.equal(xb, xa) and torch.equal(out1b, out1a) and torch.equal(stb, sta)
    # torch on the same operands
    F = torch.nn.functional
    mid = (x0.float() + o.float() @ wo.float().t() + bo).to(torch.bfloat16).float()
    for b, l in enumerate(lens):
        r0 = lay.row0[b]
        assert (x_mid[r0:r0 + l].float() - mid[r0:r0 + l]).abs().max().item() < 2e-2 * max(1.0, mid[r0:r0 + l].abs().max().item())
    h = F.gelu(F.layer_norm(mid, (256,)) @ w1f.float().t() + b1f).to(torch.bfloat16).float()
    want = mid + h @ w2.float().t() + b2
    for b, l in enumerate(lens):
        r0 = lay.row0[b]
        err = (xb[r0:r0 + l].float() - want[r0:r0 + l]).abs().max().item()
        assert err < 3e-2 * max(1.0, want[r0:r0 + l].abs().max().item()), (b, l, err)
        for buf in (xb, out1b, stb):
            assert buf[r0 + l:r0 + lay.strides[b]].float().abs().max().item() == 0
    # second launch without the optional outputs on the same buffers (barrier phases, TMEM reuse)
    xc = x0.clone()
    ops.block_ffn_op(xc, w1f, u1, b1f, w2, b2, lay, attn_out=o, w_out=wo, b_out=bo)()
    _sync()
    assert torch.equal(xc, xb)


def test_folded_layernorm_projection_matches_layernorm_then_projection():
    """epilogue.row_stats: the QKV projection reading the RAW bf16 stream with norm1 folded into its weights (statistics
    from cvb200_block_ffn's side output) against LayerNorm -> bf16 -> projection on the same engine, and against torch."""
    from cosyvoice_b200.packing import folded_linear_plan, linear_plan
    torch.manual_seed(43)
    lens = [300, 77, 640]
    lay = SeqLayout.build(lens, DEV)
    R = lay.rows
    x = torch.zeros(R, 256, device=DEV)
    for b, l in enumerate(lens):
        x[lay.row0[b]:lay.row0[b] + l] = torch.randn(l, 256, device=DEV) * 1.3 + 2 * torch.randn(l, 1, device=DEV)
    x = x.to(torch.bfloat16)
    w = torch.randn(1536, 256) * 0.06
    g, be = 1 + 0.2 * torch.randn(256), 0.1 * torch.randn(256)
    pf = folded_linear_plan(w, None, g, be, torch.bfloat16, DEV)
    xf = x.float()
    mean, rstd = xf.mean(dim=1), torch.rsqrt(xf.var(dim=1, unbiased=False) + 1e-5)
    stats = torch.stack([rstd, -mean * rstd], dim=1).contiguous()
    out = torch.zeros(R, 1536, device=DEV, dtype=torch.bfloat16)
    ops.gemm_op(pf, x, lay, out, engine=_lib.ENGINE_TCGEN05, row_stats=stats, row_u=pf.extra["row_u"])()
    _sync()
    want = torch.nn.functional.layer_norm(xf, (256,), g.to(DEV), be.to(DEV)) @ w.to(DEV).t()
    for b, l in enumerate(lens):
        r0 = lay.row0[b]
        err = (out[r0:r0 + l].float() - want[r0:r0 + l]).abs().max().item()
        assert err < 3e-2 * max(1.0, want[r0:r0 + l].abs().max().item()), (b, err)
        assert out[r0 + l:r0 + lay.strides[b]].float().abs().max().item() == 0


@pytest.mark.parametrize("lens", [[300, 77, 640], [50], [609] * 12])
def test_row_partials_fold_layernorm_between_two_gemms(lens):
    """epilogue.row_parts_out / row_parts: a 256-wide GEMM (out-projection + bf16 residual) hands (sum, sum of squares) per
    32-column chunk of the rows IT STORES to the next GEMM, which folds LayerNorm from them -- no LayerNorm launch in between.
    Checked: the partial sums against torch on the stored stream; the consumer against the row_stats path fed with torch's
    statistics of the same stream (identical operands, so only the variance formula differs) and against torch's
    LayerNorm -> Linear.  The small case runs in the GEMM's latency mode (64-wide N tiles: two chunks per tile)."""
    from cosyvoice_b200.packing import folded_linear_plan
    torch.manual_seed(47)
    lay = SeqLayout.build(lens, DEV)
    R = lay.rows
    x0 = torch.zeros(R, 256, device=DEV)
    o = torch.zeros(R, 512, device=DEV)
    for b, l in enumerate(lens):
        x0[lay.row0[b]:lay.row0[b] + l] = torch.randn(l, 256, device=DEV) * 1.3 + 2 * torch.randn(l, 1, device=DEV)
        o[lay.row0[b]:lay.row0[b] + l] = torch.randn(l, 512, device=DEV)
    x0, o = x0.to(torch.bfloat16), o.to(torch.bfloat16)
    p_out = linear_plan(torch.randn(256, 512) * (1.0 / 512) ** 0.5, torch.randn(256) * 0.1, torch.bfloat16, DEV)
    w = torch.randn(1536, 256) * 0.06
    g, be = 1 + 0.2 * torch.randn(256), 0.1 * torch.randn(256)
    pf = folded_linear_plan(w, None, g, be, torch.bfloat16, DEV)
    x = x0.clone()
    parts = torch.full((R, 8, 2), 7.0, device=DEV)
    ops.gemm_op(p_out, o, lay, x, engine=_lib.ENGINE_TCGEN05, res=x, row_parts_out=parts)()
    # the producer's other outputs are unchanged by the extra side output
    x_ref = x0.clone()
    ops.gemm_op(p_out, o, lay, x_ref, engine=_lib.ENGINE_TCGEN05, res=x_ref)()
    _sync()
    assert torch.equal(x, x_ref)
    xf = x.float().reshape(R, 8, 32)
    assert torch.allclose(parts[..., 0], xf.sum(-1), rtol=1e-5, atol=1e-3)
    assert torch.allclose(parts[..., 1], (xf * xf).sum(-1), rtol=1e-5, atol=1e-3)
    # consumer: partial sums vs finished statistics
    q1 = torch.zeros(R, 1536, device=DEV, dtype=torch.bfloat16)
    ops.gemm_op(pf, x, lay, q1, engine=_lib.ENGINE_TCGEN05, row_parts=parts, row_u=pf.extra["row_u"])()
    xr = x.float()
    mean, rstd = xr.mean(dim=1), torch.rsqrt(xr.var(dim=1, unbiased=False) + 1e-5)
    q2 = torch.zeros(R, 1536, device=DEV, dtype=torch.bfloat16)
    ops.gemm_op(pf, x, lay, q2, engine=_lib.ENGINE_TCGEN05, row_stats=torch.stack([rstd, -mean * rstd], 1).contiguous(),
                row_u=pf.extra["row_u"])()
    _sync()
    want = torch.nn.functional.layer_norm(xr, (256,), g.to(DEV), be.to(DEV)) @ w.to(DEV).t()
    for b, l in enumerate(lens):
        r0 = lay.row0[b]
        scale = max(1.0, want[r0:r0 + l].abs().max().item())
        assert (q1[r0:r0 + l].float() - q2[r0:r0 + l].float()).abs().max().item() < 2e-2 * scale, b      # a bf16 ulp or two
        assert (q1[r0:r0 + l].float() - want[r0:r0 + l]).abs().max().item() < 3e-2 * scale, b
        assert q1[r0 + l:r0 + lay.strides[b]].float().abs().max().item() == 0
        assert parts[r0 + l:r0 + lay.strides[b]].abs().max().item() == 0
