"""B200-native BigVGAN trunk (cosyvoice/BigVGAN/bigvgan.py:406-436): conv_pre -> speaker-conditioned
transposed-conv stages with three AMP blocks each -> anti-aliased SnakeBeta -> conv_post -> tanh.

The token embedding / Conformer encoders in front of the trunk (bigvgan.py:387-403) belong to the encoder
package and stay in PyTorch; `inference(feat, spk)` takes the encoder_proj output.
The reference class has no `.inference`; this one mirrors the trunk of `forward`.
Every conv is a tap GEMM; the anti-aliased activation (the reference's only native kernel,
alias_free_activation/cuda/anti_alias_activation_cuda.cu) is `anti_alias_act_rows`.
"""
from typing import Dict, List, Optional, Sequence

import torch

from . import _lib, ops
from .layout import SeqLayout
from .packing import conv1d_plan, conv_transpose1d_plan, fold_weight_norm

SD = Dict[str, torch.Tensor]


def _arch(topology: str) -> dict:
    if topology == "24k100":     # BASELINE config 4: NVIDIA 24 kHz / 100-band topology
        return dict(in_ch=100, rates=[4, 4, 2, 2, 2, 2], ks=[8, 8, 4, 4, 4, 4], ch0=1536, spk=512)
    if topology == "fork":       # examples/tts_vc/cosyvoice2/conf/cosyvoice_bigvgan_tts.yaml:43-61
        return dict(in_ch=1024, rates=[8, 5, 3, 2], ks=[16, 11, 7, 2], ch0=1536, spk=512)
    raise ValueError(topology)


def kaiser_sinc_filter12(device) -> torch.Tensor:
    """alias_free_activation/torch/filter.py:31-63 with cutoff 0.25, half_width 0.3, kernel 12 (the values the
    reference hard-codes in UpSample1d/DownSample1d for ratio 2, resample.py:13-21,44-51)."""
    import math
    ks, cutoff, half_width = 12, 0.25, 0.3
    half = ks // 2
    A = 2.285 * (half - 1) * math.pi * (4 * half_width) + 7.95
    beta = 0.1102 * (A - 8.7) if A > 50.0 else (0.5842 * (A - 21) ** 0.4 + 0.07886 * (A - 21.0) if A >= 21.0 else 0.0)
    window = torch.kaiser_window(ks, beta=beta, periodic=False)
    time = torch.arange(-half, half) + 0.5
    f = 2 * cutoff * window * torch.sinc(2 * cutoff * time)
    return (f / f.sum()).float().to(device).contiguous()


def _pad64(c: int) -> int:
    return (c + 63) // 64 * 64


class _AmpW:
    """AMPBlock1 (bigvgan.py:32-143): 3 x (act, dilated conv, act, conv) with residual; AMPBlock2 (bigvgan.py:146-232,
    selected by `resblock: "2"`): 3 x (act, dilated conv) with residual.  The block type is read off the state_dict keys."""

    def __init__(self, sd: SD, p: str, k: int, dtype, device):
        f32 = lambda key: sd[key].detach().float().to(device).contiguous()
        self.convs1, self.convs2, self.acts = [], [], []
        self.kind = 2 if f"{p}convs.0.bias" in sd else 1
        if self.kind == 2:
            for i, d in enumerate((1, 3, 5)):
                self.convs1.append(conv1d_plan(fold_weight_norm(sd, f"{p}convs.{i}."), sd[f"{p}convs.{i}.bias"], dtype, device,
                                               dilation=d))
                self.acts.append((f32(f"{p}activations.{i}.act.alpha"), f32(f"{p}activations.{i}.act.beta")))
            return
        for i, d in enumerate((1, 3, 5)):
            self.convs1.append(conv1d_plan(fold_weight_norm(sd, f"{p}convs1.{i}."), sd[f"{p}convs1.{i}.bias"], dtype,
                                           device, dilation=d))
            self.convs2.append(conv1d_plan(fold_weight_norm(sd, f"{p}convs2.{i}."), sd[f"{p}convs2.{i}.bias"], dtype,
                                           device))
        for a in range(6):
            self.acts.append((f32(f"{p}activations.{a}.act.alpha"), f32(f"{p}activations.{a}.act.beta")))


class _Program:
    def __init__(self, g: "B200BigVGAN", lens: Sequence[int]):
        self.g = g
        dev, dt, a = g.device, g.dtype, g.arch
        self.lens = [int(l) for l in lens]
        B = len(self.lens)
        lay0 = SeqLayout.build(self.lens, dev, align=128, halo=8)
        self.lay0 = lay0
        z = lambda *shape, dtype=dt: torch.zeros(*shape, dtype=dtype, device=dev)
        f32 = torch.float32
        self.feat = z(lay0.rows, _pad64(a["in_ch"]))
        self.spk = z(B, a["spk"], dtype=f32)
        chs = [a["ch0"] // 2 ** (i + 1) for i in range(len(a["rates"]))]
        self.cond0 = z(B, a["ch0"], dtype=f32)
        self.conds = [z(B, (c + 15) // 16 * 16, dtype=f32) for c in chs]
        self.ops: List[ops.Op] = []
        self.flops = 0
        # speaker conditioning: 1x1 convs on spk [B,512,1] = per-sequence biases (bigvgan.py:407,417-418)
        self.ops.append(lambda spk=self.spk, c0=self.cond0: ops.small_linear(spk, g.cond_w, g.cond_b, c0))
        for i in range(len(chs)):
            self.ops.append(lambda i=i, spk=self.spk, c=self.conds[i]: ops.small_linear(spk, g.conds_w[i], g.conds_b[i], c))
        x_prev = z(lay0.rows, a["ch0"])
        self._gemm(g.conv_pre, self.feat, lay0, x_prev, seq_bias=self.cond0)
        lay_prev, P = lay0, 1
        self.marks = {}
        for i, (u, c) in enumerate(zip(a["rates"], chs)):
            P *= u
            L = lay0.derive(P, 1, [l * P for l in self.lens])
            R, ld = L.rows, _pad64(c)
            xm, xs, xr = z(R, ld, dtype=f32), z(R, ld, dtype=f32), z(R, ld, dtype=f32)
            av, bv = z(R, ld), z(R, ld)
            self._gemm(g.ups[i], x_prev, lay_prev, xm, seq_bias=self.conds[i])
            last = i + 1 == len(chs)
            xb = None if last else z(R, ld)       # operand-dtype copy of the stage output for ups[i+1] (:413-415)
            for j, rb in enumerate(g.res[i]):
                fin = dict(out0=xs, out_scale=1.0 / 3.0, accumulate=j > 0)
                if j == 2 and not last:
                    fin.update(out1=xb, out1_mode=_lib.OUT1_COPY)
                self._amp(rb, L, c, xm, av, bv, xr, final=fin)
            self.marks[f"stage{i}"] = (xs, L)
            x_prev, lay_prev = (xs if last else xb), L
        c_last = chs[-1]
        self.ops.append(lambda xs=x_prev, av=av, c=c_last, L=lay_prev: ops.anti_alias_act_rows(
            xs, av, g.filt, g.post_alpha, g.post_beta, c, L))
        self.out = z(lay_prev.rows, 16, dtype=f32)
        self._gemm(g.conv_post, av, lay_prev, self.out, act=_lib.ACT_TANH)
        self.lay_out, self.P = lay_prev, P

    def _gemm(self, plan, A, L, out0, **kw):
        fl = plan.flops_per_row * sum(L.lens)
        self.ops.append(ops.gemm_op(plan, A, L, out0, engine=self.g.engine, flops=fl, **kw))
        self.flops += fl

    def _amp(self, rb: _AmpW, L, c, xin, av, bv, xr, final: dict):
        g = self.g
        x_cur = xin
        if rb.kind == 2:                      # AMPBlock2.forward (bigvgan.py:223-228): x = conv_d(act(x)) + x
            for j in range(3):
                a1, b1 = rb.acts[j]
                self.ops.append(lambda x=x_cur, a1=a1, b1=b1: ops.anti_alias_act_rows(x, av, g.filt, a1, b1, c, L))
                if j < 2:
                    self._gemm(rb.convs1[j], av, L, xr, res=x_cur)
                    x_cur = xr
                else:
                    self._gemm(rb.convs1[j], av, L, final.pop("out0"), res=x_cur, **final)
            return
        for j in range(3):
            a1, b1 = rb.acts[2 * j]
            a2, b2 = rb.acts[2 * j + 1]
            self.ops.append(lambda x=x_cur, a1=a1, b1=b1: ops.anti_alias_act_rows(x, av, g.filt, a1, b1, c, L))
            self._gemm(rb.convs1[j], av, L, bv)
            self.ops.append(lambda a2=a2, b2=b2: ops.anti_alias_act_rows(bv, av, g.filt, a2, b2, c, L))
            if j < 2:
                self._gemm(rb.convs2[j], av, L, xr, res=x_cur)
                x_cur = xr
            else:
                self._gemm(rb.convs2[j], av, L, final.pop("out0"), res=x_cur, **final)

    def run(self):
        for op in self.ops:
            op()


class B200BigVGAN:
    def __init__(self, state_dict: SD, topology: str = "24k100", dtype: torch.dtype = torch.bfloat16, device="cuda",
                 engine: Optional[int] = None):
        _lib.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("B200BigVGAN runs on CUDA only")
        if engine is None:
            engine = _lib.ENGINE_TCGEN05 if dtype == torch.bfloat16 else _lib.ENGINE_SIMT
        self.engine, self.dtype, self.arch = engine, dtype, _arch(topology)
        sd, a, dev = state_dict, self.arch, self.device
        f32 = lambda key: sd[key].detach().float().to(dev).contiguous()
        self.filt = kaiser_sinc_filter12(dev)
        self.conv_pre = conv1d_plan(fold_weight_norm(sd, "conv_pre."), sd["conv_pre.bias"], dtype, dev)
        self.cond_w, self.cond_b = f32("cond_layer.weight").reshape(a["ch0"], a["spk"]).contiguous(), f32("cond_layer.bias")
        self.ups, self.res, self.conds_w, self.conds_b = [], [], [], []
        for i, (u, k) in enumerate(zip(a["rates"], a["ks"])):
            c = a["ch0"] // 2 ** (i + 1)
            self.ups.append(conv_transpose1d_plan(fold_weight_norm(sd, f"ups.{i}.0."), sd[f"ups.{i}.0.bias"], dtype, dev,
                                                  stride=u, padding=(k - u) // 2))
            cp = (c + 15) // 16 * 16          # epilogue reads seq_bias in 16-column chunks: pad with zeros
            w = torch.zeros(cp, a["spk"], device=dev)
            b = torch.zeros(cp, device=dev)
            w[:c], b[:c] = f32(f"conds.{i}.weight").reshape(c, a["spk"]), f32(f"conds.{i}.bias")
            self.conds_w.append(w.contiguous())
            self.conds_b.append(b.contiguous())
            self.res.append([_AmpW(sd, f"resblocks.{i * 3 + j}.", rk, dtype, dev) for j, rk in enumerate((3, 7, 11))])
        self.post_alpha, self.post_beta = f32("activation_post.act.alpha"), f32("activation_post.act.beta")
        self.conv_post = conv1d_plan(fold_weight_norm(sd, "conv_post."), sd["conv_post.bias"], dtype, dev)
        self._programs: Dict[tuple, _Program] = {}

    def program(self, lens) -> _Program:
        key = tuple(int(l) for l in lens)
        if key not in self._programs:
            ops.trim_cache(self._programs, self.device, 4)          # LRU by count and memory; holders keep evicted programs alive
            self._programs[key] = _Program(self, key)
        return self._programs[key]

    @torch.inference_mode()
    def inference(self, feat: torch.Tensor, spk: torch.Tensor, lens: Optional[Sequence[int]] = None) -> torch.Tensor:
        """feat [B, C_in, T] (encoder_proj output, bigvgan.py:403), spk [B, 512, 1] -> wav [B, T * prod(rates)]
        (rows past lens[b] * prod(rates) are zero)."""
        B, _, T = feat.shape
        lens = [T] * B if lens is None else [int(l) for l in lens]
        prog = self.program(lens)
        ops.nct_to_rows(feat.to(self.device, torch.float32).contiguous(), prog.feat, prog.lay0)
        prog.spk.copy_(spk.reshape(B, -1).to(self.device, torch.float32))
        prog.run()
        wav = torch.zeros(B, 1, T * prog.P, dtype=torch.float32, device=self.device)
        ops.rows_to_nct(prog.out, wav, prog.lay_out, 1)
        return wav.squeeze(1)
