"""Host logic of the flow front-end (no GPU): the weight packing of cosyvoice_b200/frontend.py reproduces the reference
layers through the tap-GEMM addressing contract (CPU emulation in tests/tapgemm_emul.py), and the host-side tables equal
the oracle's.  Reference: transformer/upsample_encoder.py:33-96 (Upsample1D, PreLookaheadLayer), transformer/attention.py:66-76
(linear_q/k/v), flow/length_regulator.py:30-41,52-70, transformer/embedding.py:224-254."""
import math

import pytest
import torch
import torch.nn.functional as F

from cosyvoice_b200 import frontend
from cosyvoice_b200.layout import SeqLayout
from oracle import encoder_ref as E
from oracle import init_weights
from tapgemm_emul import emulate, from_rows, to_rows

DEV = "cpu"


@pytest.fixture(scope="module")
def w2():
    sd = init_weights.flow_frontend_sd(True, seed=2)
    return sd, frontend._Weights(sd, True, torch.float32, DEV)


@pytest.fixture(scope="module")
def w1():
    sd = init_weights.flow_frontend_sd(False, seed=2)
    return sd, frontend._Weights(sd, False, torch.float32, DEV)


def _cmp(got, want, tol=2e-4):
    for g, w in zip(got, want):
        assert g.shape == w.shape, (g.shape, w.shape)
        assert (g - w).abs().max().item() < tol, (g - w).abs().max().item()


def test_tables_equal_the_oracles():
    for half in (1, 7, 384):
        assert torch.equal(frontend.rel_pos_table(half), E.rel_pos_table(half))
    for n1, n2 in ((12, 50), (0, 30), (7, 41), (75, 250), (0, 40)):
        for sr in (22050, 24000):
            l1, l2 = int(n1 / 25 * sr / 256), int(n2 / 25 * sr / 256)
            assert frontend.regulator_segments(n1, n2, l1, l2, 25, sr) == E.regulator_segments(n1, n2, l1, l2, 25, sr)


def test_upsample1d_as_two_phase_plan(w2):
    """nearest x2 + left-pad 4 + conv k5 == two output phases over x[m-2..m] with pre-summed taps (ragged batch: the rows
    before a sequence are the previous utterance's zero padding)."""
    sd, w = w2
    lens = [9, 1, 40]
    lay = SeqLayout.build(lens, DEV, halo=8)
    xs = [torch.randn(512, l) for l in lens]
    D = emulate(w.up, to_rows(xs, lay, 512), lay, out_limit=[2 * l for l in lens])          # [rows, 2, 512]
    out = D.reshape(lay.rows * 2, 512)
    up = lay.derive(2, 1, [2 * l for l in lens])
    want = [E._upsample1d(sd, "encoder.up_layer.", x.t()[None])[0].t() for x in xs]
    _cmp(from_rows(out, up, 512), want)


def test_pre_lookahead_plans(w2):
    sd, w = w2
    lens = [5, 33]
    lay = SeqLayout.build(lens, DEV, halo=8)
    xs = [torch.randn(512, l) for l in lens]
    A = to_rows(xs, lay, 512)
    h = F.leaky_relu(emulate(w.pl1, A, lay)[:, 0])
    y = emulate(w.pl2, h, lay)[:, 0] + A
    for b, l in enumerate(lens):            # pad rows stay zero (they are the next conv's padding)
        y[lay.row0[b] + l:lay.row0[b] + lay.strides[b]] = 0
    want = [E._pre_lookahead(sd, "encoder.pre_lookahead_layer.", x.t()[None])[0].t() for x in xs]
    _cmp(from_rows(y, lay, 512), want)


def test_qkv_plan_and_embed_affine(w1):
    sd, w = w1
    lens = [11]
    lay = SeqLayout.build(lens, DEV)
    x = torch.randn(512, 11)
    D = emulate(w.layers[3].qkv, to_rows([x], lay, 512), lay)[:, 0]
    p = "encoder.encoders.3.self_attn."
    for i, n in enumerate("qkv"):
        want = F.linear(x.t(), sd[p + f"linear_{n}.weight"], sd[p + f"linear_{n}.bias"])
        assert (D[:11, 512 * i:512 * (i + 1)] - want).abs().max() < 2e-4
    # x * sqrt(d_model) of EspnetRelPositionalEncoding.forward folded into the LayerNorm affine
    h = torch.randn(4, 512)
    ref = F.layer_norm(h, (512,), sd["encoder.embed.out.1.weight"], sd["encoder.embed.out.1.bias"], 1e-5) * math.sqrt(512)
    got = F.layer_norm(h, (512,), w.embed.gamma, w.embed.beta, 1e-5)
    assert (got - ref).abs().max() < 1e-4


def test_regulator_conv_plans_on_128_wide_rows(w1):
    sd, w = w1
    lens = [60, 7]
    lay = SeqLayout.build(lens, DEV, align=256)
    xs = [torch.randn(80, l) for l in lens]
    plan, gamma, beta = w.reg[2]
    D = emulate(plan, to_rows(xs, lay, 128), lay)[:, 0]
    want = [F.conv1d(x[None], sd["length_regulator.model.6.weight"], sd["length_regulator.model.6.bias"], padding=1)[0] for x in xs]
    _cmp(from_rows(D, lay, 80), want)
    assert torch.equal(gamma, sd["length_regulator.model.7.weight"]) and torch.equal(beta, sd["length_regulator.model.7.bias"])
    D = emulate(w.reg_out, to_rows(xs, lay, 128), lay)[:, 0]
    want = [F.conv1d(x[None], sd["length_regulator.model.12.weight"], sd["length_regulator.model.12.bias"])[0] for x in xs]
    _cmp(from_rows(D, lay, 80), want)
