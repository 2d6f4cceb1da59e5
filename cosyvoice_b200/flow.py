"""B200-native flow-matching decoder: the U-Net estimator and the Euler/CFG solver.

Host-side mirror of the reference classes (same constructor-level concepts, same call signatures,
same state_dict key names), composed from the kernels of libcvb200.so:

  B200Estimator         <- cosyvoice/flow/decoder.py:88-334          ConditionalDecoder (v1 and causal v2)
  B200ConditionalCFM    <- cosyvoice/flow/flow_matching.py:22-124    ConditionalCFM.forward / solve_euler
  B200CausalConditionalCFM <- cosyvoice/flow/flow_matching.py:209-240

Differences from the reference are in mechanism only: activations are time-major and varlen-packed
(so a batch of utterances gives exactly the per-utterance B=1 results, SURVEY.md section 7 "ragged
batches"), the CFG pair is built on device without the six staging copies of flow_matching.py:103-108,
the [2B,T,T] attention mask is never materialised (it is a softmax no-op at B=1, flow/decoder.py:258),
and a whole solve is a flat list of kernel launches that can be replayed from a CUDA graph.
"""
import math
import os
import weakref
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import _lib, ops
from .layout import SeqLayout
from .packing import TapPlan, conv1d_plan, conv_transpose1d_plan, folded_linear_plan, linear_plan

SD = Dict[str, torch.Tensor]
C = 256           # estimator width
HEADS, HEAD_DIM = 8, 64
TIME_DIM = 1024
FUSED_FFN_MIN_TILES = 48      # row tiles (128 rows) from which block_ffn beats LN + FF1 + FF2 + LN (see _Program._stage)
N_FEATS = 80


def _arch(causal: bool) -> dict:
    """Stage structure of the two shipped configs (examples/tts_vc/cosyvoice/conf/cosyvoice.yaml:91-112,
    cosyvoice2/conf/cosyvoice.yaml): channels [256,256] non-causal, [256] causal; 4 TBs per stage; 12 mid."""
    return dict(n_down=1 if causal else 2, n_mid=12, n_blocks=4)


class _Weights:
    """Packed parameters of one estimator (reference key names: decoder.estimator.* stripped)."""

    @staticmethod
    def _ffn_weights(sd: SD, p: str, dtype, device):
        """Operands of the fused second half of a block (bf16 engine only, csrc/block_ffn_tc.cu): norm3 folded into the
        FFN-in weights (ops.fold_layernorm), FFN-out in its nn.Linear layout."""
        if dtype != torch.bfloat16:
            return None
        w1, w2 = sd[p + "ff.net.0.proj.weight"], sd[p + "ff.net.2.weight"]
        if tuple(w1.shape) != (1024, 256) or tuple(w2.shape) != (256, 1024):
            return None
        dev = lambda t, dt: t.detach().to(device=device, dtype=dt).contiguous()
        w1f, u1, b1f = ops.fold_layernorm(w1.to(device), sd[p + "ff.net.0.proj.bias"], sd[p + "norm3.weight"],
                                          sd[p + "norm3.bias"], torch.bfloat16)
        wo, bo = sd[p + "attn1.to_out.0.weight"], sd[p + "attn1.to_out.0.bias"]
        outp = (dev(wo, torch.bfloat16), dev(bo, torch.float32)) if tuple(wo.shape) == (256, 512) else None
        return (w1f, u1, b1f, dev(w2, torch.bfloat16), dev(sd[p + "ff.net.2.bias"], torch.float32), outp)

    def __init__(self, sd: SD, causal: bool, dtype: torch.dtype, device):
        self.causal, self.dtype, self.device = causal, dtype, device
        a = _arch(causal)
        f32 = lambda k: sd[k].detach().float().to(device).contiguous()
        self.t_w1, self.t_b1 = f32("time_mlp.linear_1.weight"), f32("time_mlp.linear_1.bias")
        self.t_w2, self.t_b2 = f32("time_mlp.linear_2.weight"), f32("time_mlp.linear_2.bias")
        self.resnets: List[dict] = []
        self.stages: List[dict] = []
        norm_idx = 2 if causal else 1
        pad_left = 2 if causal else None

        def conv3(key_w, key_b, **kw):
            return conv1d_plan(sd[key_w], sd[key_b], dtype, device, pad_left=pad_left, **kw)

        def block1d(p):
            return dict(conv=conv3(p + "block.0.weight", p + "block.0.bias"),
                        gamma=f32(p + f"block.{norm_idx}.weight"), beta=f32(p + f"block.{norm_idx}.bias"))

        def resnet(p):
            r = dict(b1=block1d(p + "block1."), b2=block1d(p + "block2."),
                     res=linear_plan(sd[p + "res_conv.weight"], sd[p + "res_conv.bias"], dtype, device),
                     mlp_w=sd[p + "mlp.1.weight"].detach().float(), mlp_b=sd[p + "mlp.1.bias"].detach().float(),
                     index=len(self.resnets))
            self.resnets.append(r)
            return r

        def tblock(p):
            wqkv = torch.cat([sd[p + "attn1.to_q.weight"], sd[p + "attn1.to_k.weight"], sd[p + "attn1.to_v.weight"]], 0)
            return dict(ln1=(f32(p + "norm1.weight"), f32(p + "norm1.bias")),
                        qkv=linear_plan(wqkv, None, dtype, device),
                        # norm1 folded into the QKV projection: used when the previous block's fused kernel left the row
                        # statistics of the stream (bf16 engine)
                        qkv_f=(folded_linear_plan(wqkv, None, sd[p + "norm1.weight"], sd[p + "norm1.bias"], dtype, device)
                               if dtype == torch.bfloat16 else None),
                        out=linear_plan(sd[p + "attn1.to_out.0.weight"], sd[p + "attn1.to_out.0.bias"], dtype, device),
                        ln3=(f32(p + "norm3.weight"), f32(p + "norm3.bias")),
                        ff1=linear_plan(sd[p + "ff.net.0.proj.weight"], sd[p + "ff.net.0.proj.bias"], dtype, device),
                        # norm3 folded into FFN-in: small layouts, statistics from the out-projection's epilogue (row partials)
                        ff1_f=(folded_linear_plan(sd[p + "ff.net.0.proj.weight"], sd[p + "ff.net.0.proj.bias"],
                                                  sd[p + "norm3.weight"], sd[p + "norm3.bias"], dtype, device)
                               if dtype == torch.bfloat16 else None),
                        ff2=linear_plan(sd[p + "ff.net.2.weight"], sd[p + "ff.net.2.bias"], dtype, device),
                        ffw=self._ffn_weights(sd, p, dtype, device))

        def stage(p):
            return dict(resnet=resnet(p + "0."), tbs=[tblock(p + f"1.{i}.") for i in range(a["n_blocks"])])

        self.down, self.mid, self.up = [], [], []
        for i in range(a["n_down"]):
            p = f"down_blocks.{i}."
            s = stage(p)
            if i < a["n_down"] - 1:      # Downsample1D: Conv1d(k3, stride 2, pad 1) reading the skip half of a cat buffer
                s["tail"] = conv1d_plan(sd[p + "2.conv.weight"], sd[p + "2.conv.bias"], dtype, device, stride=2,
                                        pad_left=1, col_pitch=2 * C)
                s["tail_kind"] = "down"
            else:
                s["tail"] = conv3(p + "2.weight", p + "2.bias")
                s["tail_kind"] = "conv"
            self.down.append(s)
        for i in range(a["n_mid"]):
            self.mid.append(stage(f"mid_blocks.{i}."))
        for i in range(a["n_down"]):
            p = f"up_blocks.{i}."
            s = stage(p)
            if i < a["n_down"] - 1:      # Upsample1D: ConvTranspose1d(C, C, 4, 2, 1)
                s["tail"] = conv_transpose1d_plan(sd[p + "2.conv.weight"], sd[p + "2.conv.bias"], dtype, device,
                                                  stride=2, padding=1)
                s["tail_kind"] = "up"
            else:
                s["tail"] = conv3(p + "2.weight", p + "2.bias")
                s["tail_kind"] = "conv"
            self.up.append(s)
        self.final_block = block1d("final_block.")
        self.final_proj = linear_plan(sd["final_proj.weight"], sd["final_proj.bias"], dtype, device)
        # all per-resnet time projections (Mish -> Linear 1024->256, components/decoder.py:47,56) as one matrix
        self.mlp_w = torch.cat([r["mlp_w"] for r in self.resnets], 0).to(device).contiguous()
        self.mlp_b = torch.cat([r["mlp_b"] for r in self.resnets], 0).to(device).contiguous()
        self.n_res = len(self.resnets)


class _Program:
    """One estimator forward over a fixed varlen layout: buffers + a flat list of launches."""

    def __init__(self, w: _Weights, lay: SeqLayout, engine: int):
        self.w, self.lay, self.engine = w, lay, engine
        dev, dt = w.device, w.dtype
        causal = w.causal
        # Fused second half of a transformer block (csrc/block_ffn_tc.cu): norm3 + FeedForward + residual (+ the next
        # block's norm1) in one kernel on the bf16 residual stream.  CVB200_FUSED_FFN=0 restores the four separate launches.
        # CVB200_FUSED_FFN = auto (default: by layout size, see _stage) | 1 (every layout) | 0 (never).  Pinning it to 0 or 1
        # makes a batch member bit-identical to its own B = 1 call (same kernels at every size).
        mode = os.environ.get("CVB200_FUSED_FFN", "auto")
        if mode not in ("auto", "0", "1"):
            raise ValueError(f"CVB200_FUSED_FFN must be auto, 0 or 1 (got {mode!r})")
        self.fused_ffn = engine == _lib.ENGINE_TCGEN05 and mode != "0"
        self.fused_ffn_min_tiles = FUSED_FFN_MIN_TILES if mode == "auto" else 1
        # CVB200_FUSED_OUT=1: the attention out-projection as the first GEMM of the same kernel (cvb200_block_out_ffn).  The two
        # forms are bit-identical (tests/test_gpu_ops.py); measured on the bench workload the fused form is NOT faster (block_ffn
        # 43.9 -> 62.4 ms per step, tap GEMM 111.0 -> 91.7 ms, step 252.5 -> 255.0 ms): every CTA already streams the block's
        # 1 MB of FFN weights per 128-row tile from L2 (~10 TB/s over 148 SMs) and the out-projection adds 384 KB per tile to
        # that stream, so it stays off by default.
        self.fused_out = os.environ.get("CVB200_FUSED_OUT", "0") == "1"
        self.fold_parts = os.environ.get("CVB200_FOLD_PARTS", "1") == "1"
        R, S = lay.rows, lay.n_seq
        z = lambda *shape, dtype=dt: torch.zeros(*shape, dtype=dtype, device=dev)
        self.t = z(S, dtype=torch.float32)
        self.tsin = z(S, 320, dtype=torch.float32)
        self.th = z(S, TIME_DIM, dtype=torch.float32)
        self.temb = z(S, TIME_DIM, dtype=torch.float32)
        self.tbias = z(S, w.n_res * C, dtype=torch.float32)
        n_norm = 2 * w.n_res + 1
        self.gn_stats = None if causal else z(n_norm, S, 8, 2, dtype=torch.float64)
        self.A0 = z(R, 320)
        self.h, self.hn = z(R, C), z(R, C)
        # Residual stream of the transformer blocks in the engine's operand dtype: fp32 for the fp32 parity engine, bf16 for
        # the bf16 engine -- what the reference's own half-precision path does (flow.half(), cli/model.py:43-45).  The
        # out-projection / FFN-out GEMMs and both LayerNorms of a block are bound by this stream's HBM traffic: bf16 removes
        # 19 % of a block's bytes (step 298 -> 283 ms) for 0.3 % more rel-rms error on the mel (0.92 % against the 3 % bar;
        # profiles/r01_residual_bf16.md).  CVB200_RESIDUAL=f32 keeps the stream in fp32.
        self.X = z(R, C, dtype=torch.float32 if os.environ.get("CVB200_RESIDUAL", "act") == "f32" else dt)
        self.n = z(R, C)
        self.row_stats = z(R, 2, dtype=torch.float32)       # (rstd, -mean rstd) of the stream's rows, left by block_ffn
        self.row_parts = z(R, 8, 2, dtype=torch.float32)    # (sum, sum of squares) per 32-column chunk, left by a GEMM epilogue
        self.qkv, self.o, self.f = z(R, 3 * HEADS * HEAD_DIM), z(R, HEADS * HEAD_DIM), z(R, 4 * C)
        self.s_in = z(R, C)
        self.fin = z(R, C)
        self.v = z(R, N_FEATS, dtype=torch.float32)
        self.ops: List[ops.Op] = []
        self.taps: Dict[str, Tuple[torch.Tensor, SeqLayout]] = {}
        self.flops = 0
        self._norm_slot = 0

        n_down = len(w.down)
        lays = [lay]
        for i in range(1, n_down):
            lays.append(lays[-1].derive(1, 2, [(l + 1) // 2 for l in lays[-1].lens]))
        cats = [z(lays[i].rows, 2 * C) for i in range(n_down)]      # [h | skip] inputs of the up path
        self.cats = cats
        self.lays = lays

        cur_in, cur_lay = self.A0, lay
        # ---- down path (flow/decoder.py:254-279)
        for i, st in enumerate(w.down):
            L = lays[i]
            self._stage(st, cur_in[:L.rows], L, last_copy=cats[i][:, C:])
            self.taps[f"down{i}.tb"] = (self.X, L)
            skipv = cats[i][:, C:]
            if st["tail_kind"] == "down":
                Ln = lays[i + 1]
                self._gemm(st["tail"], skipv, Ln, self.s_in[:Ln.rows], a_cols=3 * C)
                cur_lay = Ln
            else:
                self._gemm(st["tail"], skipv, L, self.s_in[:L.rows])
                cur_lay = L
            cur_in = self.s_in
        # ---- mid blocks (:283-304)
        Lm = cur_lay
        for j, st in enumerate(w.mid):
            last = j == len(w.mid) - 1
            dst = cats[n_down - 1][:, :C] if last else self.s_in[:Lm.rows]
            self._stage(st, self.s_in[:Lm.rows], Lm, last_copy=dst)
        self.taps["mid"] = (self.X, Lm)
        # ---- up path (:306-331)
        for i, st in enumerate(w.up):
            lvl = n_down - 1 - i
            L = lays[lvl]
            self._stage(st, cats[lvl], L, last_copy=self.s_in[:L.rows])
            if st["tail_kind"] == "up":
                Lo = lays[lvl - 1]
                # M space = input rows; output rows 2t+phase, cropped to the skip length (flow/decoder.py:309)
                self._gemm(st["tail"], self.s_in[:L.rows], L, cats[lvl - 1][:, :C], out_limit=Lo.seq_len)
                self.taps[f"up{i}"] = (cats[lvl - 1][:, :C], Lo)
            else:
                self._gemm(st["tail"], self.s_in[:L.rows], L, self.fin)
                self.taps[f"up{i}"] = (self.fin, L)
        # ---- final block + projection (:332-334)
        fb = w.final_block
        stats = self._next_stats()
        self._gemm(fb["conv"], self.fin, lay, self.h, gn_stats=stats)
        self.ops.append(ops.norm_mish_op(self.h, self.hn, fb["gamma"], fb["beta"], lay, gn_stats=stats))
        self._gemm(w.final_proj, self.hn, lay, self.v)

    def release(self):
        """Drop buffers and launch closures (reference cycles would otherwise keep the HBM until a gc pass)."""
        self.__dict__.clear()

    def set_lens(self, lens: Sequence[int]):
        """Re-bind the program (and any CUDA graph captured over it) to other valid lengths under the same strides:
        every kernel reads the lengths from the layouts' device arrays and re-zeroes the pad rows it owns."""
        for i, L in enumerate(self.lays):
            L.set_lens(lens, halo=max(1, 8 >> i))       # a level-i stride is stride0 / 2^i: the zero halo shrinks with it
            lens = [(l + 1) // 2 for l in lens]

    # -- helpers
    def _next_stats(self):
        if self.gn_stats is None:
            return None
        s = self.gn_stats[self._norm_slot]
        self._norm_slot += 1
        return s

    def _gemm(self, plan: TapPlan, A, L: SeqLayout, out0, **kw):
        fl = plan.flops_per_row * sum(L.lens)
        self.ops.append(ops.gemm_op(plan, A, L, out0, engine=self.engine, flops=fl, **kw))
        self.flops += fl

    def _stage(self, st: dict, A: torch.Tensor, L: SeqLayout, last_copy: torch.Tensor):
        """ResnetBlock1D + 4 BasicTransformerBlocks on layout L; the last FFN epilogue also writes the
        operand-dtype copy of the stage output to `last_copy`."""
        w = self.w
        R = L.rows
        r = st["resnet"]
        h, hn, X, n = self.h[:R], self.hn[:R], self.X[:R], self.n[:R]
        tb = self.tbias[:, r["index"] * C:(r["index"] + 1) * C]
        s1, s2 = self._next_stats(), self._next_stats()
        # components/decoder.py:54-59
        self._gemm(r["b1"]["conv"], A, L, h, gn_stats=s1)
        self.ops.append(ops.norm_mish_op(h, hn, r["b1"]["gamma"], r["b1"]["beta"], L, gn_stats=s1, seq_bias=tb))
        self._gemm(r["b2"]["conv"], hn, L, h, gn_stats=s2)
        self.ops.append(ops.norm_mish_op(h, hn, r["b2"]["gamma"], r["b2"]["beta"], L, gn_stats=s2))
        # Small layouts on the bf16 tcgen05 engine (where block_ffn is not selected): both LayerNorms of every block are
        # folded into the GEMM that follows, with the row statistics handed over as partial sums by the GEMM that WROTE the
        # stream (epilogue.row_parts_out -> row_parts): 2 of a block's 7 launches go, which is what a streaming chunk is
        # made of (launch latency, not FLOPs).  CVB200_FOLD_PARTS=0 restores the LayerNorm launches.
        fuse = (self.fused_ffn and X.dtype == torch.bfloat16 and L.m_tiles >= self.fused_ffn_min_tiles
                and all(t["ffw"] is not None for t in st["tbs"]))
        parts_path = (not fuse and self.fold_parts and self.engine == _lib.ENGINE_TCGEN05 and X.dtype == torch.bfloat16
                      and all(t["qkv_f"] is not None and t["ff1_f"] is not None for t in st["tbs"]))
        parts = self.row_parts[:R] if parts_path else None
        self._gemm(r["res"], A, L, X, res=hn, **(dict(row_parts_out=parts) if parts_path else {}))
        if r["index"] == 0:
            self.taps["down0.resnet"] = (self.X, L)
        qkv, o, f = self.qkv[:R], self.o[:R], self.f[:R]
        n_tb = len(st["tbs"])
        # Fused second half (norm3 + FFN + residual, one kernel) from FUSED_FFN_MIN_TILES 128-row tiles up.  One CTA owns
        # a tile's whole FFN (27 us alone), while the separate GEMMs spread a small problem over all SMs by output
        # column: below ~40 tiles the four launches win inside a CUDA graph (streaming chunk 30.6 vs 34.3 ms,
        # profiles/r02_ffn_rows_sweep.txt).  The choice depends on the STRIDE pattern only (L.m_tiles), so every call of
        # a program takes the same kernels and batch members stay independent of their neighbours.
        stats = self.row_stats[:R]
        for k, t in enumerate(st["tbs"]):
            # components/transformer.py:262-316
            last = k == n_tb - 1
            if fuse and k > 0:
                # norm1 folded into the projection: the previous block's kernel left (rstd, -mean rstd) of every row
                pf = t["qkv_f"]
                self._gemm(pf, X, L, qkv, row_stats=stats, row_u=pf.extra["row_u"])
            elif parts_path:
                pf = t["qkv_f"]
                self._gemm(pf, X, L, qkv, row_parts=parts, row_u=pf.extra["row_u"])
            else:
                self.ops.append(ops.layernorm_op(X, n, t["ln1"][0], t["ln1"][1], R))
                self._gemm(t["qkv"], n, L, qkv, bias=False)
            fl = sum(4 * HEADS * HEAD_DIM * l * l for l in L.lens)
            self.ops.append(ops.attention_op(qkv, o, L, HEADS, HEAD_DIM ** -0.5, flops=fl))
            self.flops += fl
            if fuse:
                # to_out + residual, norm3, FeedForward, residual (+ the next block's norm1 statistics): one kernel
                w1f, u1, b1f, w2, b2, outp = t["ffw"]
                fl = (t["ff1"].flops_per_row + t["ff2"].flops_per_row) * sum(L.lens)
                kw = {}
                if outp is not None and self.fused_out:
                    fl += t["out"].flops_per_row * sum(L.lens)
                    kw = dict(attn_out=o, w_out=outp[0], b_out=outp[1])
                else:
                    self._gemm(t["out"], o, L, X, res=X)
                self.ops.append(ops.block_ffn_op(X, w1f, u1, b1f, w2, b2, L, flops=fl, out1=last_copy if last else None,
                                                 stats_out=None if last else stats, **kw))
                self.flops += fl
                continue
            if parts_path:
                self._gemm(t["out"], o, L, X, res=X, row_parts_out=parts)
                pf = t["ff1_f"]
                self._gemm(pf, X, L, f, act=_lib.ACT_GELU, row_parts=parts, row_u=pf.extra["row_u"])
                kw = dict(out1=last_copy, out1_mode=_lib.OUT1_COPY) if last else dict(row_parts_out=parts)
                self._gemm(t["ff2"], f, L, X, res=X, **kw)
                continue
            self._gemm(t["out"], o, L, X, res=X)
            self.ops.append(ops.layernorm_op(X, n, t["ln3"][0], t["ln3"][1], R))
            self._gemm(t["ff1"], n, L, f, act=_lib.ACT_GELU)
            if last:
                self._gemm(t["ff2"], f, L, X, res=X, out1=last_copy, out1_mode=_lib.OUT1_COPY)
            else:
                self._gemm(t["ff2"], f, L, X, res=X)

    def run_time_embedding(self, shared: bool = False):
        """t [S] (self.t) -> per-resnet time biases (flow/decoder.py:240-241, components/decoder.py:47,56).
        shared=True: every sequence has the same t (the CFM solver): evaluate one row and broadcast it."""
        w = self.w
        n = 1 if shared else self.t.shape[0]
        ops.time_sinusoid(self.t[:n], self.tsin[:n])
        ops.small_linear(self.tsin[:n], w.t_w1, w.t_b1, self.th[:n], post_act=_lib.ACT_SILU)
        ops.small_linear(self.th[:n], w.t_w2, w.t_b2, self.temb[:n])
        ops.small_linear(self.temb[:n], w.mlp_w, w.mlp_b, self.tbias[:n], pre_act=_lib.ACT_MISH)
        if shared and self.t.shape[0] > 1:
            self.tbias[1:].copy_(self.tbias[0:1].expand(self.t.shape[0] - 1, -1))

    def run(self):
        if self.gn_stats is not None:
            self.gn_stats.zero_()
        for op in self.ops:
            op()


class B200Estimator:
    """Drop-in for the reference estimator object (flow.decoder.estimator): callable
    (x, mask, mu, t, spks, cond) -> dphi_dt with the tensor names/shapes of bin/export_onnx.py:68-69."""

    def __init__(self, state_dict: SD, causal: bool = False, dtype: torch.dtype = torch.bfloat16, device="cuda",
                 engine: Optional[int] = None):
        _lib.load()
        self.causal = causal
        self.dtype = dtype
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("B200Estimator runs on CUDA only")
        if engine is None:
            engine = _lib.ENGINE_TCGEN05 if dtype == torch.bfloat16 else _lib.ENGINE_SIMT
        if engine == _lib.ENGINE_TCGEN05 and dtype != torch.bfloat16:
            raise ValueError("the tcgen05 engine takes bf16 operands")
        self.engine = engine
        self.weights = _Weights(state_dict, causal, dtype, self.device)
        self.static_chunk_size = 0          # attribute written by cli/model.py:50,315
        self._programs: Dict[tuple, _Program] = {}
        self._cfms = weakref.WeakSet()       # solvers on this estimator: their CUDA graphs pin programs (see program())
        self.serial = ops.Serial()          # programs and solver state are shared scratch: one user at a time

    def layout_for(self, lens: Sequence[int]) -> SeqLayout:
        align = 128 * (2 ** (len(self.weights.down) - 1))
        return SeqLayout.build(lens, self.device, align=align, halo=8)

    def program(self, lay: SeqLayout) -> _Program:
        """Programs are cached per STRIDE pattern (lengths are device data, re-bound in place), so a streaming
        session whose T grows chunk by chunk reuses one program / CUDA graph per 128-row bucket."""
        key = lay.stride_key()
        p = self._programs.pop(key, None)
        if p is None:
            # LRU, bounded by count and by memory (ops.trim_cache).  An evicted program is only dropped from the cache, not torn
            # down: holders of a reference (AcousticBackend.prepare_resident*, a captured CUDA graph) keep a consistent object
            # and its HBM is returned when the last reference goes.  Graphs pin their programs, so they are trimmed first.
            for cfm in list(self._cfms):
                ops.trim_cache(cfm._graphs, self.device, 33)
            ops.trim_cache(self._programs, self.device, 32)
            p = _Program(self.weights, SeqLayout(lay.lens, lay.strides, self.device), self.engine)
        else:
            p.set_lens(lay.lens)
        self._programs[key] = p
        return p

    @torch.inference_mode()
    def forward(self, x, mask, mu, t, spks, cond, streaming=False, taps: Optional[dict] = None):
        """x, mu, cond [N,80,T]; mask [N,1,T]; t [N]; spks [N,80] -> [N,80,T] (flow/decoder.py:222-334)."""
        N, _, T = x.shape
        lens = [int(v) for v in mask.reshape(N, T).sum(dim=1).round().tolist()]
        with self.serial:
            return self._forward_locked(x, mu, t, spks, cond, lens, taps)

    def _forward_locked(self, x, mu, t, spks, cond, lens, taps):
        N, _, T = x.shape
        lay = self.layout_for(lens)
        prog = self.program(lay)
        f32 = lambda a: a.detach().to(self.device, torch.float32)
        x, mu, cond = f32(x).contiguous(), f32(mu).contiguous(), f32(cond).contiguous()
        sp = f32(spks).unsqueeze(-1).expand(-1, -1, T).contiguous()
        ops.nct_to_rows(x, prog.A0[:, 0:80], lay)
        ops.nct_to_rows(mu, prog.A0[:, 80:160], lay)
        ops.nct_to_rows(sp, prog.A0[:, 160:240], lay)
        ops.nct_to_rows(cond, prog.A0[:, 240:320], lay)
        prog.t.copy_(f32(t).reshape(N))
        prog.run_time_embedding()
        prog.run()
        out = torch.zeros(N, N_FEATS, T, dtype=torch.float32, device=self.device)
        ops.rows_to_nct(prog.v, out, lay)
        if taps is not None:
            for name, (buf, L) in prog.taps.items():
                taps[name] = (buf, L)
        return out

    __call__ = forward


def cosine_t_span(n_timesteps: int, device, dtype=torch.float32) -> torch.Tensor:
    """flow_matching.py:67-69."""
    s = torch.linspace(0, 1, n_timesteps + 1, device=device, dtype=dtype)
    return 1 - torch.cos(s * 0.5 * torch.pi)


def euler_schedule(t_span: torch.Tensor) -> Tuple[List[float], List[float]]:
    """The (t, dt) pairs of solve_euler: t is a running fp32 sum, dt = t_span[k+1] - t
    (flow_matching.py:87,118-122).  Evaluated in fp32 on the host exactly as the reference does on device."""
    ts = t_span.detach().float().cpu()
    t, dt = ts[0], ts[1] - ts[0]
    out_t, out_dt = [], []
    for step in range(1, len(ts)):
        out_t.append(float(t))
        out_dt.append(float(dt))
        t = t + dt
        if step < len(ts) - 1:
            dt = ts[step + 1] - t
    return out_t, out_dt


class B200ConditionalCFM:
    """cosyvoice/flow/flow_matching.py:22-124 (ConditionalCFM) on the B200 kernels.

    `forward` keeps the reference signature and B=1 semantics (noise from torch.randn_like under the
    caller's seed, flow_cache splice, in-place mutation of mu).  `solve_batch` is the varlen batched
    entry used for throughput: per-utterance results identical to B=1 calls."""

    FLOW_CACHE_TAIL = 34        # hard-coded at flow_matching.py:63-64

    def __init__(self, estimator: B200Estimator, inference_cfg_rate: float = 0.7, t_scheduler: str = "cosine",
                 sigma_min: float = 1e-6):
        self.estimator = estimator
        estimator._cfms.add(self)
        self.inference_cfg_rate = inference_cfg_rate
        self.t_scheduler = t_scheduler
        self.sigma_min = sigma_min
        self.device = estimator.device
        self._graphs: Dict[tuple, tuple] = {}
        self.use_cuda_graph = True
        self.last_graph_launches = 0        # kernels of libcvb200 inside the last replayed graph

    # ---- batched core -----------------------------------------------------------------
    def _state(self, lay1: SeqLayout):
        """Solver state buffers for a B-sequence layout (cached with the estimator program)."""
        lay2 = lay1.paired()
        prog = self.estimator.program(lay2)
        if not hasattr(prog, "cfm_state"):
            z = lambda: torch.zeros(lay1.rows, N_FEATS, dtype=torch.float32, device=self.device)
            prog.cfm_state = dict(lay1=SeqLayout(lay1.lens, lay1.strides, self.device), x=z(), mu=z(), cond=z(),
                                  spks=torch.zeros(lay1.n_seq, N_FEATS, dtype=torch.float32, device=self.device))
        prog.cfm_state["lay1"].set_lens(lay1.lens)
        return prog, prog.cfm_state

    def _time_table(self, prog: _Program, ts: List[float]) -> torch.Tensor:
        """The per-resnet time biases of every Euler step, [n_steps, n_res * 256]: they depend on nothing but the weights and
        the schedule (SinusoidalPosEmb -> TimestepEmbedding -> Mish -> Linear per resnet, components/decoder.py:18-27, :71-115,
        :47,56 with t_span fixed at flow_matching.py:67-69), so they are evaluated once per (program, schedule) with the same
        kernels a step would run and a step only copies its row."""
        key = tuple(ts)
        tab = getattr(prog, "_time_tables", None)
        if tab is None:
            tab = prog._time_tables = {}
        if key not in tab:
            rows = []
            for t in ts:
                prog.t.fill_(t)
                prog.run_time_embedding(shared=True)
                rows.append(prog.tbias[0].clone())
            if len(tab) >= 4:
                tab.pop(next(iter(tab)))
            tab[key] = torch.stack(rows, 0).contiguous()
        return tab[key]

    def _steps(self, prog: _Program, st: dict, ts: List[float], dts: List[float], table: Optional[torch.Tensor] = None):
        for k, (t, dt) in enumerate(zip(ts, dts)):
            if table is not None:
                prog.tbias.copy_(table[k:k + 1].expand(prog.tbias.shape[0], -1))
            else:
                prog.t.fill_(t)
                prog.run_time_embedding(shared=True)
            ops.pack_cfg_input(st["x"], st["mu"], st["cond"], st["spks"], prog.A0, st["lay1"])
            prog.run()
            ops.euler_cfg(st["x"], prog.v, dt, self.inference_cfg_rate)

    def solve_rows(self, lay1: SeqLayout, n_timesteps: int = 10):
        """Runs the Euler solve on the state buffers of `lay1` (x holds the noise on entry, the mel on exit)."""
        prog, st = self._state(lay1)
        t_span = torch.linspace(0, 1, n_timesteps + 1, dtype=torch.float32)
        if self.t_scheduler == "cosine":
            t_span = 1 - torch.cos(t_span * 0.5 * torch.pi)
        ts, dts = euler_schedule(t_span)
        table = self._time_table(prog, ts)
        if not self.use_cuda_graph:
            self._steps(prog, st, ts, dts, table)
            return
        # the captured launches bake in the schedule and the CFG rate: both are part of the key, so that changing
        # cfm.inference_cfg_rate / t_scheduler after the first solve re-captures instead of being silently ignored
        key = (lay1.stride_key(), n_timesteps, float(self.inference_cfg_rate), self.t_scheduler)
        g = self._graphs.pop(key, None)
        if g is not None and g[1] is not prog:      # the program of this bucket was evicted and rebuilt
            g = None
        if g is None:
            # warm-up run on a side stream (also sets the kernels' function attributes), then capture
            keep = st["x"].clone()
            s = torch.cuda.Stream(device=self.device)
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                self._steps(prog, st, ts[:1], dts[:1], table)
            torch.cuda.current_stream().wait_stream(s)
            st["x"].copy_(keep)
            graph = torch.cuda.CUDAGraph()
            n0 = _lib.launch_count()
            # thread_local: other session threads (front-end, HiFT, the LLM thread) may allocate or copy while this
            # thread captures; in the default "global" mode their calls would fail and invalidate the capture
            with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                self._steps(prog, st, ts, dts, table)
            n_launch = _lib.launch_count() - n0
            st["x"].copy_(keep)
            ops.trim_cache(self._graphs, self.estimator.device, 32)     # least recently used first; count and memory
            g = (graph, prog, n_launch, table)       # the graph reads `table`: keep it alive with the graph
        self._graphs[key] = g                                        # (re-)insert at the most-recent end
        self.last_graph_launches = g[2]
        g[0].replay()

    @torch.inference_mode()
    def solve_batch(self, mus: Sequence[torch.Tensor], spks: Sequence[torch.Tensor], conds: Sequence[torch.Tensor],
                    zs: Sequence[torch.Tensor], n_timesteps: int = 10) -> List[torch.Tensor]:
        """mus/conds/zs: per-utterance [1,80,T_u] (or [80,T_u]) fp32; spks [1,80].  Returns per-utterance
        mel [1,80,T_u] fp32."""
        with self.estimator.serial:
            return self._solve_batch_locked(mus, spks, conds, zs, n_timesteps)

    def _solve_batch_locked(self, mus, spks, conds, zs, n_timesteps):
        B = len(mus)
        lens = [int(m.shape[-1]) for m in mus]
        lay1 = self.estimator.layout_for(lens)
        prog, st = self._state(lay1)
        dev = self.device
        T_max = max(lens)

        def stack(seq):
            buf = torch.zeros(B, N_FEATS, T_max, dtype=torch.float32, device=dev)
            for b, a in enumerate(seq):
                buf[b, :, :lens[b]] = a.reshape(N_FEATS, lens[b]).to(dev, torch.float32)
            return buf
        ops.nct_to_rows(stack(zs), st["x"], lay1)
        ops.nct_to_rows(stack(mus), st["mu"], lay1)
        ops.nct_to_rows(stack(conds), st["cond"], lay1)
        st["spks"].copy_(torch.cat([s.reshape(1, N_FEATS) for s in spks], 0).to(dev, torch.float32))
        self.solve_rows(lay1, n_timesteps)
        out = torch.zeros(B, N_FEATS, T_max, dtype=torch.float32, device=dev)
        ops.rows_to_nct(st["x"], out, lay1)
        return [out[b:b + 1, :, :lens[b]].clone() for b in range(B)]

    # ---- reference-compatible API ---------------------------------------------------------
    def _noise(self, mu):
        return torch.randn_like(mu)

    @torch.inference_mode()
    def forward(self, mu, mask, n_timesteps, temperature=1.0, spks=None, cond=None, prompt_len=0,
                flow_cache=None):
        """flow_matching.py:37-70.  mu [1,80,T] (mutated in place like the reference, :62)."""
        if mu.shape[0] != 1:
            raise ValueError("the reference path is batch-1 (cosyvoice/flow/flow.py:125); use solve_batch for batches")
        if flow_cache is None:
            flow_cache = torch.zeros(1, 80, 0, 2, device=mu.device, dtype=mu.dtype)
        z = self._noise(mu).to(mu.device).to(mu.dtype) * temperature
        cache_size = min(flow_cache.shape[2], mu.size(2))
        if cache_size != 0:
            z[:, :, :cache_size] = flow_cache[:, :, :cache_size, 0]
            mu[:, :, :cache_size] = flow_cache[:, :, :cache_size, 1]
        z_cache = torch.concat([z[:, :, :prompt_len], z[:, :, -self.FLOW_CACHE_TAIL:]], dim=2)
        mu_cache = torch.concat([mu[:, :, :prompt_len], mu[:, :, -self.FLOW_CACHE_TAIL:]], dim=2)
        new_cache = torch.stack([z_cache, mu_cache], dim=-1)
        if spks is None:
            spks = torch.zeros(1, N_FEATS, device=mu.device)
        if cond is None:
            cond = torch.zeros_like(mu)
        mel = self.solve_batch([mu], [spks], [cond], [z], n_timesteps)[0]
        return mel.float(), new_cache

    __call__ = forward


class B200CausalConditionalCFM(B200ConditionalCFM):
    """flow_matching.py:209-240: fixed noise buffer drawn once with seed 0, no flow cache."""

    def __init__(self, estimator: B200Estimator, inference_cfg_rate: float = 0.7, t_scheduler: str = "cosine",
                 sigma_min: float = 1e-6):
        super().__init__(estimator, inference_cfg_rate, t_scheduler, sigma_min)
        g = torch.Generator(device="cpu")
        g.manual_seed(0)            # set_all_random_seed(0) + torch.randn on the default CPU generator (:212-213)
        self.rand_noise = torch.randn([1, 80, 50 * 300], generator=g)

    @torch.inference_mode()
    def forward(self, mu, mask, n_timesteps, temperature=1.0, spks=None, cond=None):
        if mu.shape[0] != 1:
            raise ValueError("the reference path is batch-1 (cosyvoice/flow/flow.py:277)")
        z = self.rand_noise[:, :, :mu.size(2)].to(mu.device).to(mu.dtype) * temperature
        if spks is None:
            spks = torch.zeros(1, N_FEATS, device=mu.device)
        if cond is None:
            cond = torch.zeros_like(mu)
        mel = self.solve_batch([mu], [spks], [cond], [z], n_timesteps)[0]
        return mel.float(), None

    __call__ = forward
