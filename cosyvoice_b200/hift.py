"""B200-native HiFT vocoder (cosyvoice/hifigan/generator.py:223-411 HiFTGenerator +
cosyvoice/hifigan/f0_predictor.py:19-55 ConvRNNF0Predictor).

`B200HiFTGenerator.inference(speech_feat, cache_source)` keeps the reference signature and RNG call
order (generator.py:149-150,163,219); `inference_batch` is the varlen batched entry.  Every conv /
transposed conv is an instance of the tap GEMM; Snake / LeakyReLU / residual / 1/3-average / source
fusion ride in its epilogue; the sine source, STFT and the exp-sin-iSTFT-OLA head are single kernels.
"""
import math
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.distributions.uniform import Uniform

from . import _lib, ops
from .layout import SeqLayout, copy_async_
from .packing import conv1d_plan, conv_transpose1d_plan, fold_weight_norm, linear_plan

SD = Dict[str, torch.Tensor]
N_FFT, HOP = 16, 4
NB_HARM = 8


def _arch(v2: bool) -> dict:
    """examples/tts_vc/cosyvoice/conf/cosyvoice.yaml:114-136 / cosyvoice2/conf/cosyvoice.yaml."""
    if v2:
        return dict(rates=[8, 5, 3], ks=[16, 11, 7], src_k=[7, 7, 11])
    return dict(rates=[8, 8], ks=[16, 16], src_k=[7, 11])


class _ResBlockW:
    """cosyvoice/hifigan/generator.py:43-103 ResBlock(channels, k, dilations (1,3,5))."""

    def __init__(self, sd: SD, p: str, k: int, dtype, device):
        f32 = lambda key: sd[key].detach().float().to(device).contiguous()
        self.convs1, self.convs2, self.alpha1, self.alpha2 = [], [], [], []
        for i, d in enumerate((1, 3, 5)):
            self.convs1.append(conv1d_plan(fold_weight_norm(sd, f"{p}convs1.{i}."), sd[f"{p}convs1.{i}.bias"], dtype,
                                           device, dilation=d))
            self.convs2.append(conv1d_plan(fold_weight_norm(sd, f"{p}convs2.{i}."), sd[f"{p}convs2.{i}.bias"], dtype,
                                           device))
            self.alpha1.append(f32(f"{p}activations1.{i}.alpha"))
            self.alpha2.append(f32(f"{p}activations2.{i}.alpha"))


def _split3_conv_plan(w: torch.Tensor, bias, device, cpad: int):
    """'same' Conv1d [N, C, k] on the bf16 split of its fp32 input ([rows, 3*cpad], ops.split3_bf16) with the same split
    of its weights: atoms (tap shift, term column block, weight term) for the cross terms of lowest order."""
    from .packing import build_plan
    w = w.detach().float()
    N, Cin, k = w.shape
    w1 = w.to(torch.bfloat16).float()
    w2 = (w - w1).to(torch.bfloat16).float()
    w3 = (w - w1 - w2).to(torch.bfloat16).float()
    terms = (w1, w2, w3)
    atoms = []
    for j in range(k):
        q = j - (k - 1) // 2
        pairs = ((0, 0), (0, 1), (1, 0), (0, 2), (1, 1), (2, 0))
        if os.environ.get("CVB200_F0_TERMS", "3") != "6":
            pairs = pairs[:3]
        for i, kk in pairs:
            atoms.append((q, i * cpad, terms[kk][:, :, j]))
    return build_plan([atoms], N, cpad, bias, torch.bfloat16, device)


class _Weights:
    def __init__(self, sd: SD, v2: bool, dtype, device):
        a = _arch(v2)
        self.v2, self.dtype, self.device = v2, dtype, device
        self.rates, self.ks = a["rates"], a["ks"]
        f32 = lambda key: sd[key].detach().float().to(device).contiguous()
        self.hop = int(np.prod(self.rates)) * HOP
        # The f0 predictor never runs on plain bf16 operands: its output is integrated over the whole utterance by
        # the sine source (generator.py:148), so operand rounding would show up as audible phase drift.
        #   fp32 engine: CUDA-core tap GEMM (parity mode).
        #   bf16 engine: activations and weights are split into bf16 terms (x = x1 + x2 [+ x3], ops.split3_bf16) and
        #   ONE tcgen05 tap GEMM sums the cross terms x1w1 + x1w2 + x2w1 in its fp32 accumulator (16-bit operand
        #   mantissas).  Measured f0 error against the fp32 engine: 3e-5 relative (tests/test_gpu_hift.py); the
        #   24-bit six-term variant (CVB200_F0_TERMS=6) measures 5e-5, i.e. the tensor core's accumulator rounding,
        #   not the operands, is what is left.  3x the MMA work of a bf16 conv, ~10x faster than the CUDA-core path
        #   (which was 3 % of the bench step).
        self.f0_convs, self.f0_split_convs = [], []
        cpad = 128
        for i in range(5):
            p = f"f0_predictor.condnet.{2 * i}."
            w_f0 = fold_weight_norm(sd, p)
            self.f0_convs.append(conv1d_plan(w_f0, sd[p + "bias"], torch.float32, device, c_pad=cpad))
            if dtype == torch.bfloat16:
                self.f0_split_convs.append(_split3_conv_plan(w_f0, sd[p + "bias"], device, cpad))
            cpad = 512
        self.f0_w = f32("f0_predictor.classifier.weight").reshape(-1)
        self.f0_b = f32("f0_predictor.classifier.bias").reshape(-1)
        self.src_w = f32("m_source.l_linear.weight").reshape(-1)
        self.src_b = f32("m_source.l_linear.bias").reshape(-1)
        self.conv_pre = conv1d_plan(fold_weight_norm(sd, "conv_pre."), sd["conv_pre.bias"], dtype, device, c_pad=128)
        n_up = len(self.rates)
        down = [1] + self.rates[::-1][:-1]
        self.src_stride = [int(v) for v in list(np.cumprod(down))[::-1]]      # generator.py:290-300
        self.ups, self.src_down, self.src_res, self.res = [], [], [], []
        for i, (u, k) in enumerate(zip(self.rates, self.ks)):
            last = i == n_up - 1
            self.ups.append(conv_transpose1d_plan(fold_weight_norm(sd, f"ups.{i}."), sd[f"ups.{i}.bias"], dtype, device,
                                                  stride=u, padding=(k - u) // 2, shift=1 if last else 0))
            su = self.src_stride[i]
            wsd, bsd = sd[f"source_downs.{i}.weight"], sd[f"source_downs.{i}.bias"]
            if su == 1:
                self.src_down.append(conv1d_plan(wsd, bsd, dtype, device, c_pad=64))
            else:
                self.src_down.append(conv1d_plan(wsd, bsd, dtype, device, stride=su, pad_left=su // 2, c_pad=64))
            self.src_res.append(_ResBlockW(sd, f"source_resblocks.{i}.", a["src_k"][i], dtype, device))
            self.res.append([_ResBlockW(sd, f"resblocks.{i * 3 + j}.", k3, dtype, device) for j, k3 in enumerate((3, 7, 11))])
        self.conv_post = conv1d_plan(fold_weight_norm(sd, "conv_post."), sd["conv_post.bias"], dtype, device)
        self.channels = [512 // 2 ** (i + 1) for i in range(n_up)]


class _Program:
    """Launch list + buffers of one batched HiFT forward over a fixed set of mel lengths."""

    def __init__(self, w: _Weights, lens: Sequence[int], engine: int, strides: Optional[Sequence[int]] = None,
                 sampling_rate: float = 24000.0, audio_limit: float = 0.99):
        self.w, self.engine = w, engine
        dev, dt = w.device, w.dtype
        self.lens = [int(l) for l in lens]
        B = len(self.lens)
        lay0 = (SeqLayout.build(self.lens, dev, align=128, halo=8) if strides is None
                else SeqLayout(self.lens, strides, dev))
        self.lay0 = lay0
        self.Lm = None
        z = lambda *shape, dtype=dt: torch.zeros(*shape, dtype=dtype, device=dev)
        f32 = torch.float32
        R0 = lay0.rows
        n_up = len(w.rates)
        # stage layouts; the last one carries the reflection-padded length T*P + 1 (= number of STFT frames)
        self.lays: List[SeqLayout] = []
        P = 1
        for i, u in enumerate(w.rates):
            P *= u
            extra = 1 if i == n_up - 1 else 0
            self.lays.append(lay0.derive(P, 1, [l * P + extra for l in self.lens]))
        self.P = P
        lay_f = self.lays[-1]
        # sample buffers are sized for the longest utterance the strides admit, so that the program can be re-bound
        # to other lengths of the same stride bucket (set_lens); l_cur = samples of the longest CURRENT utterance
        self.l_max = (max(lay0.strides) - 8) * w.hop
        self.l_cur = max(self.lens) * w.hop
        assert self.l_max % 4 == 0
        # ---- inputs / source path
        self.mel32 = z(R0, 128, dtype=f32)
        self.mel = self.mel32 if dt == f32 else z(R0, 128)
        self.fa, self.fb = z(R0, 512, dtype=f32), z(R0, 512, dtype=f32)
        self.f0 = z(R0, dtype=f32)
        self.prefix = z(R0, dtype=torch.float64)
        self.noise = z(B, NB_HARM + 1, self.l_max, dtype=f32)
        self.phase = z(B, NB_HARM + 1, dtype=f32)
        self.s = z(B, self.l_max, dtype=f32)
        self.sst = z(lay_f.rows, 64)
        self.wav = z(B, self.l_max, dtype=f32)
        self.ops: List[ops.Op] = []
        self.marks: Dict[str, Tuple[int, torch.Tensor, SeqLayout]] = {}
        self.flops = 0
        simt = _lib.ENGINE_SIMT

        # ---- f0 predictor: 5 x (conv k3 + ELU), Linear 512->1, abs (fp32 accuracy on either engine, see _Weights)
        self.ops_f0: List[ops.Op] = []
        if w.f0_split_convs and engine == _lib.ENGINE_TCGEN05:
            self.msplit = z(R0, 3 * 128)
            self.hsplit = z(R0, 3 * 512)
            self.ops_f0.append(lambda m=self.mel32, d=self.msplit: ops.split3_bf16(m, d, 80, 128))
            a_in = self.msplit
            for i, (plan, plan32) in enumerate(zip(w.f0_split_convs, w.f0_convs)):
                fl = plan32.flops_per_row * sum(lay0.lens)          # algorithmic FLOPs of the conv, not of the 6 terms
                self.ops_f0.append(ops.gemm_op(plan, a_in, lay0, self.fa, engine=_lib.ENGINE_TCGEN05, act=_lib.ACT_ELU,
                                               flops=fl))
                self.flops += fl
                if i + 1 < len(w.f0_split_convs):
                    self.ops_f0.append(lambda m=self.fa, d=self.hsplit: ops.split3_bf16(m, d, 512, 512))
                a_in = self.hsplit
            src = self.fa
        else:
            src, dst = self.mel32, self.fa
            for i, plan in enumerate(w.f0_convs):
                fl = plan.flops_per_row * sum(lay0.lens)
                self.ops_f0.append(ops.gemm_op(plan, src, lay0, dst, engine=simt, act=_lib.ACT_ELU, flops=fl))
                self.flops += fl
                src, dst = dst, (self.fb if dst is self.fa else self.fa)
        f0_in = src
        self.ops_f0.append(lambda f0=self.f0: ops.rowdot_abs(f0_in, w.f0_w, w.f0_b, f0, R0))
        # ---- source: f0 -> s (generator.py:404-406)
        self.ops_src: List[ops.Op] = [
            lambda f0=self.f0, nz=self.noise, ph=self.phase, pre=self.prefix, s=self.s: ops.sine_source(
                f0, lay0, w.hop, float(sampling_rate), nz, ph, w.src_w, w.src_b, pre, s)]
        # ---- decode (generator.py:349-381)
        self.ops.append(lambda s=self.s, sst=self.sst: ops.stft16(s, sst, lay_f))
        self.x0 = z(R0, 512)
        self._gemm(w.conv_pre, self.mel, lay0, self.x0, act=_lib.ACT_LRELU, act_slope=0.1)
        x_prev, lay_prev = self.x0, lay0
        self.stage_bufs = []
        for i in range(n_up):
            Cc, L = w.channels[i], self.lays[i]
            last = i == n_up - 1
            R = L.rows
            xm, xs, xr, sa = (z(R, Cc, dtype=f32) for _ in range(4))
            a, b, xl = z(R, Cc), z(R, Cc), z(R, Cc)
            self.stage_bufs.append(dict(xm=xm, xs=xs, xr=xr, sa=sa, a=a, b=b, xl=xl))
            # ups[i] (+ ReflectionPad1d((1,0)) on the last stage): M space = input rows
            if last:
                Lm = lay_prev.with_lens([l + 1 for l in lay_prev.lens])
                self.Lm, self._Lm_P = Lm, P // w.rates[-1]
                self._gemm(w.ups[i], x_prev, Lm, xm, out_limit=L.seq_len, flop_lens=lay_prev.lens)
                self.ops.append(lambda xm=xm, Cc=Cc, L=L: ops.reflect_row0(xm, Cc, L))
            else:
                self._gemm(w.ups[i], x_prev, lay_prev, xm)
            # source branch: source_downs[i] -> source_resblocks[i], fused into xm
            sr = w.src_res[i]
            self._gemm(w.src_down[i], self.sst, L, sa, out1=a, out1_mode=_lib.OUT1_SNAKE, out1_alpha=sr.alpha1[0])
            first_alpha = w.res[i][0].alpha1[0]
            self._resblock(sr, L, sa, a, b, xr, final=dict(out0=xm, accumulate=True, out1=a,
                                                          out1_mode=_lib.OUT1_SNAKE, out1_alpha=first_alpha))
            self.marks[f"fuse{i}"] = (len(self.ops), xm, L)
            # three parallel ResBlocks averaged (generator.py:366-372)
            for j, rb in enumerate(w.res[i]):
                if j > 0:
                    self.ops.append(lambda xm=xm, a=a, rb=rb, Cc=Cc, L=L: ops.snake(xm, a, rb.alpha1[0], Cc, L))
                fin = dict(out0=xs, out_scale=1.0 / 3.0, accumulate=j > 0)
                if j == 2:
                    fin.update(out1=xl, out1_mode=_lib.OUT1_LRELU, out1_slope=0.01 if last else 0.1)
                self._resblock(rb, L, xm, a, b, xr, final=fin)
            self.marks[f"stage{i}"] = (len(self.ops), xs, L)
            x_prev, lay_prev = xl, L
        self.cp = z(lay_f.rows, 32, dtype=f32)
        self._gemm(w.conv_post, x_prev, lay_f, self.cp)
        self.marks["conv_post"] = (len(self.ops), self.cp, lay_f)
        self.ops.append(lambda cp=self.cp, wav=self.wav: ops.istft16(cp, lay_f, wav, float(audio_limit)))

    def release(self):
        """Drop buffers and launch closures at once (explicit teardown for an owner that knows nobody else holds the
        program).  The launch closures bind buffers, never the program itself, so a program that merely falls out of
        the LRU cache is reclaimed by reference counting as soon as its last holder lets go."""
        self.__dict__.clear()

    def set_lens(self, lens: Sequence[int]):
        """Re-bind to other mel lengths under the same strides (every kernel reads lengths from device arrays and
        re-zeroes the pad rows it owns)."""
        lens = [int(l) for l in lens]
        if lens == self.lens:
            return
        self.lay0.set_lens(lens)
        n_up, P = len(self.lays), 1
        for i, u in enumerate(self.w.rates):
            P *= u
            self.lays[i].set_lens([l * P + (1 if i == n_up - 1 else 0) for l in lens])
        if self.Lm is not None:
            self.Lm.set_lens([l * self._Lm_P + 1 for l in lens], halo=0)
        self.lens = lens
        self.l_cur = max(lens) * self.w.hop

    def _gemm(self, plan, A, L, out0, flop_lens=None, **kw):
        fl = plan.flops_per_row * sum(flop_lens if flop_lens is not None else L.lens)
        self.ops.append(ops.gemm_op(plan, A, L, out0, engine=self.engine, flops=fl, **kw))
        self.flops += fl

    def _resblock(self, rb: _ResBlockW, L: SeqLayout, xin, a, b, xr, final: dict):
        """generator.py:91-98.  xin: fp32 block input, a = snake_{alpha1[0]}(xin) in the operand dtype."""
        x_cur = xin
        for j in range(3):
            self._gemm(rb.convs1[j], a, L, b, act=_lib.ACT_SNAKE, act_alpha=rb.alpha2[j])
            if j < 2:
                self._gemm(rb.convs2[j], b, L, xr, res=x_cur, out1=a, out1_mode=_lib.OUT1_SNAKE,
                           out1_alpha=rb.alpha1[j + 1])
                x_cur = xr
            else:
                self._gemm(rb.convs2[j], b, L, final.pop("out0"), res=x_cur, **final)

    def run(self, upto: Optional[int] = None):
        for op in (self.ops if upto is None else self.ops[:upto]):
            op()


class B200HiFTGenerator:
    """Drop-in for HiFTGenerator at inference (cosyvoice/cli/model.py:161,172,352,363 call sites)."""

    def __init__(self, state_dict: SD, v2: bool = False, dtype: torch.dtype = torch.bfloat16, device="cuda",
                 engine: Optional[int] = None, sampling_rate: int = 24000, audio_limit: float = 0.99):
        _lib.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("B200HiFTGenerator runs on CUDA only")
        sd = {k[len("generator."):] if k.startswith("generator.") else k: v for k, v in state_dict.items()}   # model.py:79
        if engine is None:
            engine = _lib.ENGINE_TCGEN05 if dtype == torch.bfloat16 else _lib.ENGINE_SIMT
        self.engine, self.dtype, self.v2 = engine, dtype, v2
        self.weights = _Weights(sd, v2, dtype, self.device)
        self.sampling_rate = sampling_rate      # SineGen's F = f0 / sampling_rate (hifigan/generator.py:146-148)
        self.audio_limit = audio_limit          # final clamp (generator.py:380)
        self.hop = self.weights.hop
        self._programs: Dict[tuple, _Program] = {}
        self.serial = ops.Serial()          # cached programs are shared scratch: one user (thread / stream) at a time

    def program(self, lens: Sequence[int], uniform_stride: bool = False) -> _Program:
        """Programs are cached per stride pattern (128-frame buckets) and re-bound to the lengths of each call.
        uniform_stride: every utterance gets the stride of the longest one, so that ragged batches of one length
        bucket share a program (shard.make_batches)."""
        lens = [int(l) for l in lens]
        key = tuple((l + 8 + 127) // 128 * 128 for l in lens)
        if uniform_stride:
            key = (max(key),) * len(key)
        p = self._programs.pop(key, None)
        if p is None:
            ops.trim_cache(self._programs, self.device, 8)          # LRU by count and memory; holders keep evicted programs alive
            p = _Program(self.weights, lens, self.engine, strides=key, sampling_rate=float(self.sampling_rate),
                         audio_limit=float(self.audio_limit))
        else:
            p.set_lens(lens)
        self._programs[key] = p
        return p

    def draw_source_rng(self, prog: _Program):
        """Same draws, shapes, devices and order as the reference (generator.py:149-150 CPU Uniform sample,
        :163 randn_like(sine_waves) and :219 randn_like(uv) on the model device)."""
        B, L = len(prog.lens), prog.l_cur
        u = Uniform(low=-np.pi, high=np.pi).sample(sample_shape=(B, NB_HARM + 1, 1))
        u[:, 0, :] = 0
        copy_async_(prog.phase, u.reshape(B, NB_HARM + 1))          # pinned staging: no stream synchronisation
        prog.noise[:, :, :L].copy_(torch.randn(B, NB_HARM + 1, L, device=self.device))
        torch.randn(B, L, 1, device=self.device)          # consumed and discarded by the reference (:219)

    @torch.inference_mode()
    def inference_batch(self, mels: Sequence[torch.Tensor], phase_vec: Optional[torch.Tensor] = None,
                        noise: Optional[torch.Tensor] = None, cache_sources: Optional[Sequence[torch.Tensor]] = None,
                        source: Optional[torch.Tensor] = None, uniform_stride: bool = False
                        ) -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
        """mels: per-utterance [1,80,T_u].  phase_vec [B,9,1] / noise [B,9,L_max]: injected RNG (tests);
        source [B,1,L_max]: bypass the source module entirely.  Returns ([wav [1,L_u]], [s [1,1,L_u]])."""
        with self.serial:
            return self._inference_batch_locked(mels, phase_vec, noise, cache_sources, source, uniform_stride)

    def _inference_batch_locked(self, mels, phase_vec, noise, cache_sources, source, uniform_stride):
        lens = [int(m.shape[-1]) for m in mels]
        prog = self.program(lens, uniform_stride)
        B, dev = len(lens), self.device
        T_max = max(lens)
        buf = torch.zeros(B, 80, T_max, dtype=torch.float32, device=dev)
        for b, m in enumerate(mels):
            buf[b, :, :lens[b]] = m.reshape(80, lens[b]).to(dev, torch.float32)
        ops.nct_to_rows(buf, prog.mel32, prog.lay0)
        if prog.mel is not prog.mel32:
            ops.nct_to_rows(buf, prog.mel, prog.lay0)
        if source is None:
            if phase_vec is None:
                self.draw_source_rng(prog)
            else:
                prog.phase.copy_(phase_vec.reshape(B, NB_HARM + 1).to(dev, torch.float32))
                prog.noise.zero_()
                prog.noise[:, :, :noise.shape[-1]].copy_(noise.to(dev, torch.float32))
            for op in prog.ops_f0:
                op()
            for op in prog.ops_src:
                op()
        else:
            prog.s.zero_()
            prog.s[:, :source.shape[-1]].copy_(source.reshape(B, -1).to(dev, torch.float32))
        if cache_sources is not None:
            for b, cs in enumerate(cache_sources):
                if cs is not None and cs.shape[2] != 0:
                    prog.s[b, :cs.shape[2]].copy_(cs.reshape(-1).to(dev, torch.float32))     # generator.py:408-409
        prog.run()
        wavs = [prog.wav[b:b + 1, :lens[b] * self.hop].clone() for b in range(B)]
        ss = [prog.s[b:b + 1, :lens[b] * self.hop].clone().unsqueeze(1) for b in range(B)]
        return wavs, ss

    @torch.inference_mode()
    def inference(self, speech_feat: torch.Tensor, cache_source: torch.Tensor = torch.zeros(1, 1, 0)):
        """generator.py:399-411: speech_feat [1,80,T] -> (speech [1,L], source [1,1,L])."""
        if speech_feat.shape[0] != 1:
            raise ValueError("hift.inference is batch-1 like the reference; use inference_batch")
        wavs, ss = self.inference_batch([speech_feat], cache_sources=[cache_source])
        return wavs[0], ss[0]
