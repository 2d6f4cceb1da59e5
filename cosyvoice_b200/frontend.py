"""Flow front-end on the B200 kernels: what `flow.inference` does before the CFM, and the
reference's outer boundary objects built on it.

Reference (all under /root/reference/cosyvoice):
  flow/flow.py:108-160       MaskedDiffWithXvec.inference           -> B200MaskedDiffWithXvec.inference
  flow/flow.py:258-319       CausalMaskedDiffWithXvec.inference     -> B200CausalMaskedDiffWithXvec.inference
  transformer/encoder.py:116-165,388-476           ConformerEncoder (6 rel-pos blocks, d 512, FFN 2048)
  transformer/upsample_encoder.py:33-96,254-319    UpsampleConformerEncoder (pre-lookahead, 6 blocks, x2 upsample, 4 blocks)
  transformer/attention.py:243-330                 RelPositionMultiHeadedAttention
  transformer/embedding.py:201-298                 EspnetRelPositionalEncoding
  flow/length_regulator.py:52-70                   InterpolateRegulator.inference

Every Linear / Conv1d is a tap GEMM (tcgen05 in bf16 mode, CUDA cores in fp32 parity mode); the rest are the kernels of
csrc/encoder.cu.  The host code below only packs weights, sizes buffers and marshals pointers; there is no torch
arithmetic on the data path and no CPU fallback.
"""
import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import _lib, ops
from .layout import SeqLayout, copy_async_
from .packing import TapPlan, build_plan, conv1d_plan, linear_plan

SD = Dict[str, torch.Tensor]
D_MODEL, HEADS, D_K, FF = 512, 8, 64, 2048
N_FEATS = 80
PAD80 = 128         # row pitch of the 80-channel buffers (the GEMM's K chunk is 64)


def rel_pos_table(half: int, d_model: int = D_MODEL) -> torch.Tensor:
    """The constant table of EspnetRelPositionalEncoding.extend_pe (transformer/embedding.py:224-254), built on the host in
    fp32 exactly as the reference builds it at construction: row half-1-r holds PE(r), r = half-1 ... -(half-1)."""
    position = torch.arange(0, half, dtype=torch.float32).unsqueeze(1)
    div_term = torch.exp(torch.arange(0, d_model, 2, dtype=torch.float32) * -(math.log(10000.0) / d_model))
    pos, neg = torch.zeros(half, d_model), torch.zeros(half, d_model)
    pos[:, 0::2] = torch.sin(position * div_term)
    pos[:, 1::2] = torch.cos(position * div_term)
    neg[:, 0::2] = torch.sin(-1 * position * div_term)
    neg[:, 1::2] = torch.cos(-1 * position * div_term)
    return torch.cat([torch.flip(pos, [0]), neg[1:]], dim=0)


def regulator_segments(n1: int, n2: int, mel_len1: int, mel_len2: int, input_frame_rate: int, sample_rate: int):
    """(src_start, src_len, dst_start, dst_len) pieces of InterpolateRegulator.inference (flow/length_regulator.py:56-68)."""
    segs = []
    if n1 != 0:
        segs.append((0, n1, 0, mel_len1))
    if n2 > 40:
        e = int(20 / input_frame_rate * sample_rate / 256)
        segs += [(n1, 20, mel_len1, e), (n1 + 20, n2 - 40, mel_len1 + e, mel_len2 - 2 * e),
                 (n1 + n2 - 20, 20, mel_len1 + mel_len2 - e, e)]
    else:
        segs.append((n1, n2, mel_len1, mel_len2))
    return segs


class _LayerW:
    def __init__(self, sd: SD, p: str, dtype, device):
        f32 = lambda k: sd[k].detach().float().to(device).contiguous()
        a = p + "self_attn."
        wqkv = torch.cat([sd[a + f"linear_{n}.weight"] for n in "qkv"], 0)
        bqkv = torch.cat([sd[a + f"linear_{n}.bias"] for n in "qkv"], 0)
        self.qkv = linear_plan(wqkv, bqkv, dtype, device)
        self.out = linear_plan(sd[a + "linear_out.weight"], sd[a + "linear_out.bias"], dtype, device)
        self.pos32 = linear_plan(sd[a + "linear_pos.weight"], None, torch.float32, device)      # table projection, fp32
        self.u, self.v = f32(a + "pos_bias_u").reshape(-1), f32(a + "pos_bias_v").reshape(-1)
        self.ln_mha = (f32(p + "norm_mha.weight"), f32(p + "norm_mha.bias"))
        self.ln_ff = (f32(p + "norm_ff.weight"), f32(p + "norm_ff.bias"))
        self.ff1 = linear_plan(sd[p + "feed_forward.w_1.weight"], sd[p + "feed_forward.w_1.bias"], dtype, device)
        self.ff2 = linear_plan(sd[p + "feed_forward.w_2.weight"], sd[p + "feed_forward.w_2.bias"], dtype, device)


class _EmbedW:
    def __init__(self, sd: SD, p: str, dtype, device):
        self.lin = linear_plan(sd[p + "out.0.weight"], sd[p + "out.0.bias"], dtype, device)
        s = math.sqrt(D_MODEL)      # x * xscale of EspnetRelPositionalEncoding.forward folded into the LayerNorm affine
        self.gamma = (sd[p + "out.1.weight"].detach().float() * s).to(device).contiguous()
        self.beta = (sd[p + "out.1.bias"].detach().float() * s).to(device).contiguous()


class _Weights:
    """Packed parameters of the front-end (state_dict keys of the flow module without `decoder.*`)."""

    def __init__(self, sd: SD, v2: bool, dtype: torch.dtype, device):
        self.v2, self.dtype, self.device = v2, dtype, torch.device(device)
        f32 = lambda k: sd[k].detach().float().to(device).contiguous()
        self.table = f32("input_embedding.weight")
        self.spk_w, self.spk_b = f32("spk_embed_affine_layer.weight"), f32("spk_embed_affine_layer.bias")
        e = "encoder."
        self.embed = _EmbedW(sd, e + "embed.", dtype, device)
        self.layers = [_LayerW(sd, f"{e}encoders.{i}.", dtype, device) for i in range(6)]
        self.after = (f32(e + "after_norm.weight"), f32(e + "after_norm.bias"))
        self.proj = linear_plan(sd["encoder_proj.weight"], sd["encoder_proj.bias"], dtype, device)
        if v2:
            p = e + "pre_lookahead_layer."
            self.pl1 = conv1d_plan(sd[p + "conv1.weight"], sd[p + "conv1.bias"], dtype, device, pad_left=0)
            self.pl2 = conv1d_plan(sd[p + "conv2.weight"], sd[p + "conv2.bias"], dtype, device, pad_left=2)
            # Upsample1D (upsample_encoder.py:55-59): nearest x2, left-pad 4, conv k5  ==  two output phases over x[m-2..m]:
            #   out[2m]   = x[m-2](w0+w1) + x[m-1](w2+w3) + x[m] w4,   out[2m+1] = x[m-2] w0 + x[m-1](w1+w2) + x[m](w3+w4)
            w = sd[e + "up_layer.conv.weight"].detach().float()
            t = [w[:, :, j] for j in range(5)]
            phases = [[(-2, 0, t[0] + t[1]), (-1, 0, t[2] + t[3]), (0, 0, t[4])],
                      [(-2, 0, t[0]), (-1, 0, t[1] + t[2]), (0, 0, t[3] + t[4])]]
            self.up = build_plan(phases, D_MODEL, D_MODEL, sd[e + "up_layer.conv.bias"], dtype, device)
            self.up_embed = _EmbedW(sd, e + "up_embed.", dtype, device)
            self.up_layers = [_LayerW(sd, f"{e}up_encoders.{i}.", dtype, device) for i in range(4)]
        else:
            r = "length_regulator.model."
            self.reg = []
            for i in range(4):
                self.reg.append((conv1d_plan(sd[f"{r}{3 * i}.weight"], sd[f"{r}{3 * i}.bias"], dtype, device, c_pad=PAD80),
                                 f32(f"{r}{3 * i + 1}.weight"), f32(f"{r}{3 * i + 1}.bias")))
            self.reg_out = linear_plan(sd[r + "12.weight"], sd[r + "12.bias"], dtype, device, c_pad=PAD80)
        self._pos: Dict[Tuple[int, int], List[torch.Tensor]] = {}

    def pos_tables(self, half: int) -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
        """P_l[center - r] = linear_pos_l(PE(r)) for |r| < half, one fp32 [2*half-1 (padded), 512] table per layer: input
        independent, so it is computed once per table size (on the device, fp32 CUDA-core GEMM) and shared by every
        sequence and call (the reference recomputes linear_pos(pos_emb) per layer per call, attention.py:292-294)."""
        key = half
        if key not in self._pos:
            n_rel = 2 * half - 1
            lay = SeqLayout.build([n_rel], self.device, halo=0)
            A = torch.zeros(lay.rows, D_MODEL, dtype=torch.float32, device=self.device)
            A[:n_rel] = rel_pos_table(half).to(self.device)
            tabs = []
            for group in ([self.layers] + ([self.up_layers] if self.v2 else [])):
                g = []
                for lw in group:
                    out = torch.zeros(lay.rows, D_MODEL, dtype=torch.float32, device=self.device)
                    ops.gemm_op(lw.pos32, A, lay, out, engine=_lib.ENGINE_SIMT, bias=False)()
                    g.append(out[:n_rel])
                tabs.append(g)
            if len(self._pos) >= 8:
                self._pos.pop(next(iter(self._pos)))
            self._pos[key] = tabs
        t = self._pos[key]
        return t[0], (t[1] if self.v2 else [])


class _Program:
    """Front-end forward over fixed strides: token-rate layout `tok` (+ the x2 layout for v2), mel-rate layout `mel`.
    Lengths, tokens and regulator pieces are device data re-bound per call."""

    def __init__(self, w: _Weights, tok: SeqLayout, mel: SeqLayout, engine: int, chunk: int):
        self.w, self.engine, self.chunk = w, engine, chunk
        dev, dt = w.device, w.dtype
        self.tok, self.mel = tok, mel
        S = tok.n_seq
        z = lambda *shape, dtype=dt: torch.zeros(*shape, dtype=dtype, device=dev)
        self.up_lay = tok.derive(2, 1, [2 * l for l in tok.lens]) if w.v2 else None
        Rt = tok.rows
        Rmax = 2 * Rt if w.v2 else Rt
        half = max(tok.strides) * (2 if w.v2 else 1)
        self.center = half - 1
        pos, pos_up = w.pos_tables(half)
        for L in (tok, self.up_lay):
            if L is not None:       # the attention grid covers the whole stride, so re-bound lengths stay covered
                L.max_len, L._c = max(L.strides), None
        self.tokens = z(S, max(tok.strides), dtype=torch.int32)
        self.segs = z(S, 4, 4, dtype=torch.int32)
        self.spk_in = z(S, w.spk_w.shape[1], dtype=torch.float32)
        self.spk_n = z(S, w.spk_w.shape[1], dtype=torch.float32)
        self.spks = z(S, N_FEATS, dtype=torch.float32)
        self.E = z(Rt, D_MODEL)
        self.X = z(Rmax, D_MODEL, dtype=torch.float32)
        self.n = z(Rmax, D_MODEL)
        self.qkv = z(Rmax, 3 * D_MODEL, dtype=torch.float32)
        self.o = z(Rmax, D_MODEL)
        self.f = z(Rmax, FF)
        self.mu = z(mel.rows, N_FEATS, dtype=torch.float32)
        self.ops: List[ops.Op] = []
        self.flops = 0          # algorithmic FLOPs of one run at the lengths of construction (bench.py)
        g = self._gemm
        X, n = self.X[:Rt], self.n[:Rt]
        # ---- token embedding + LinearNoSubsampling (flow.py:131-135, subsampling.py:105-107)
        self.ops.append(ops.embed_tokens_op(self.tokens, w.table, self.E, tok))
        g(w.embed.lin, self.E, tok, self.qkv[:Rt, :D_MODEL])
        self.ops.append(ops.layernorm_seq_op(self.qkv[:Rt, :D_MODEL], X, w.embed.gamma, w.embed.beta, tok, 1e-5))
        if w.v2:
            # PreLookaheadLayer (upsample_encoder.py:80-96): conv k4 over [t, t+3], leaky_relu, causal conv k3, + input
            if dt != torch.float32:
                self.ops.append(ops.layernorm_seq_op(self.qkv[:Rt, :D_MODEL], n, w.embed.gamma, w.embed.beta, tok, 1e-5))
            g(w.pl1, X if dt == torch.float32 else n, tok, self.o[:Rt], act=_lib.ACT_LRELU, act_slope=0.01)
            g(w.pl2, self.o[:Rt], tok, X, res=X)
        self._layers(w.layers, pos, tok, Rt, chunk, copy_last=w.v2 and dt != torch.float32)
        if w.v2:
            up, Ru = self.up_lay, 2 * Rt
            src = X if dt == torch.float32 else n
            self.U = z(Ru, D_MODEL)
            g(w.up, src, tok, self.U, out_limit=up.seq_len)
            Xu = self.X[:Ru]
            self.Y = z(Ru, D_MODEL, dtype=torch.float32)
            g(w.up_embed.lin, self.U, up, self.Y)
            self.ops.append(ops.layernorm_seq_op(self.Y, Xu, w.up_embed.gamma, w.up_embed.beta, up, 1e-5))
            self._layers(w.up_layers, pos_up, up, Ru, 2 * chunk, copy_last=False)
            self.ops.append(ops.layernorm_seq_op(Xu, self.n[:Ru], w.after[0], w.after[1], up, 1e-5))
            assert mel.strides == up.strides
            g(w.proj, self.n[:Ru], up, self.mu)
        else:
            self.ops.append(ops.layernorm_seq_op(X, n, w.after[0], w.after[1], tok, 1e-5))
            self.hp = z(Rt, PAD80, dtype=torch.float32)
            g(w.proj, n, tok, self.hp)
            # InterpolateRegulator.inference (length_regulator.py:52-70)
            Rm = mel.rows
            self.ga, self.gb, self.hc = z(Rm, PAD80), z(Rm, PAD80), z(Rm, PAD80, dtype=torch.float32)
            self.ops.append(ops.interp_rows_op(self.hp, tok, self.ga, mel, self.segs, N_FEATS))
            cur, nxt = self.ga, self.gb
            for plan, gamma, beta in w.reg:
                g(plan, cur, mel, self.hc)
                self.ops.append(ops.groupnorm1_mish_op(self.hc, nxt, gamma, beta, mel, 1e-5))
                cur, nxt = nxt, cur
            g(w.reg_out, cur, mel, self.mu)

    def _gemm(self, plan: TapPlan, A, L: SeqLayout, out0, **kw):
        fl = plan.flops_per_row * sum(L.lens)
        self.ops.append(ops.gemm_op(plan, A, L, out0, engine=self.engine, flops=fl, **kw))
        self.flops += fl

    def _layers(self, layers: List[_LayerW], pos: List[torch.Tensor], L: SeqLayout, R: int, chunk: int, copy_last: bool):
        """ConformerEncoderLayer x len(layers) (encoder_layer.py:191-236; macaron/cnn off, pre-norm eps 1e-12)."""
        X, n, qkv, o, f = self.X[:R], self.n[:R], self.qkv[:R], self.o[:R], self.f[:R]
        for i, (lw, P) in enumerate(zip(layers, pos)):
            self.ops.append(ops.layernorm_seq_op(X, n, lw.ln_mha[0], lw.ln_mha[1], L, 1e-12))
            self._gemm(lw.qkv, n, L, qkv)
            fl = sum(6 * HEADS * D_K * l * l for l in L.lens)        # (q+u).k, (q+v).p and p.v
            self.ops.append(ops.relpos_attention_op(qkv, P, self.center, lw.u, lw.v, o, L, HEADS, 1.0 / math.sqrt(D_K), chunk,
                                                    flops=fl))
            self.flops += fl
            self._gemm(lw.out, o, L, X, res=X)
            self.ops.append(ops.layernorm_seq_op(X, n, lw.ln_ff[0], lw.ln_ff[1], L, 1e-12))
            self._gemm(lw.ff1, n, L, f, act=_lib.ACT_SILU)
            if copy_last and i == len(layers) - 1:      # operand-dtype copy of the stage output for the up-sampling conv
                self._gemm(lw.ff2, f, L, X, res=X, out1=n, out1_mode=_lib.OUT1_COPY)
            else:
                self._gemm(lw.ff2, f, L, X, res=X)

    def bind(self, tok_lens: Sequence[int], mel_lens: Sequence[int]):
        self.tok.set_lens(tok_lens, halo=4)
        self.tok.max_len = max(self.tok.strides)
        if self.up_lay is not None:
            self.up_lay.set_lens([2 * l for l in tok_lens], halo=8)
            self.up_lay.max_len = max(self.up_lay.strides)
        self.mel.set_lens(mel_lens, halo=8)

    def run(self):
        w = self.w
        ops.l2_normalize(self.spk_in, self.spk_n)
        ops.small_linear(self.spk_n, w.spk_w, w.spk_b, self.spks)
        for op in self.ops:
            op()

    def release(self):
        self.__dict__.clear()


def _stage_to_device(tensors: Sequence[torch.Tensor], device) -> List[torch.Tensor]:
    """fp32 device copies of a list of tensors; the host ones travel together through one pinned buffer."""
    host = [i for i, t in enumerate(tensors) if not t.is_cuda]
    out = [t if not t.is_cuda else t.to(device, torch.float32) for t in tensors]
    if host:
        sizes = [tensors[i].numel() for i in host]
        flat = torch.empty(sum(sizes), dtype=torch.float32, pin_memory=True)
        o = 0
        for i, n in zip(host, sizes):
            flat[o:o + n].copy_(tensors[i].reshape(-1))
            o += n
        dev = flat.to(device, non_blocking=True)
        o = 0
        for i, n in zip(host, sizes):
            out[i] = dev[o:o + n].view(tensors[i].shape)
            o += n
    return out


class _EncoderHandle:
    """`flow.encoder` as the caller sees it: cli/model.py:314 writes `static_chunk_size` on it."""

    def __init__(self):
        self.static_chunk_size = 0


class _FlowBase:
    """Shared by the two boundary classes: owns the front-end weights/programs, the estimator and the CFM solver."""
    V2 = False

    def __init__(self, state_dict: SD, dtype: torch.dtype = torch.bfloat16, device="cuda", input_frame_rate: int = 25,
                 estimator_state_dict: Optional[SD] = None):
        from .flow import B200CausalConditionalCFM, B200ConditionalCFM, B200Estimator
        _lib.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("cosyvoice_b200 runs on CUDA only (there is no CPU path)")
        self.dtype = dtype
        self.engine = _lib.ENGINE_TCGEN05 if dtype == torch.bfloat16 else _lib.ENGINE_SIMT
        self.input_frame_rate = input_frame_rate
        self.fp16 = False                       # attribute written by cli/model.py:42-45,308; precision is `dtype` here
        self.output_size = N_FEATS
        est_sd = estimator_state_dict
        if est_sd is None:
            pre = "decoder.estimator."
            est_sd = {k[len(pre):]: v for k, v in state_dict.items() if k.startswith(pre)}
        if not est_sd:
            raise KeyError("state_dict holds no decoder.estimator.* parameters")
        self.weights = _Weights(state_dict, self.V2, dtype, self.device)
        est = B200Estimator(est_sd, causal=self.V2, dtype=dtype, device=self.device)
        self.decoder = (B200CausalConditionalCFM if self.V2 else B200ConditionalCFM)(est)
        self.encoder = _EncoderHandle()
        self._programs: Dict[tuple, _Program] = {}
        self.serial = ops.Serial()              # front-end programs are shared scratch: one user (thread / stream) at a time

    # the reference calls these on the module (cli/model.py:43-45,76-81)
    def half(self):
        return self

    def eval(self):
        return self

    def to(self, *a, **k):
        return self

    # ---- batched front-end ---------------------------------------------------------------------------------------
    def _layouts(self, tok_lens: Sequence[int], mel_lens: Sequence[int]) -> Tuple[SeqLayout, SeqLayout]:
        # one token stride per batch (that of its longest utterance): ragged batches of one length bucket share a program
        stride = (max(int(l) for l in tok_lens) + 8 + 127) // 128 * 128
        tok = SeqLayout(tok_lens, [stride] * len(tok_lens), self.device)
        if self.V2:
            return tok, tok.derive(2, 1, list(mel_lens))
        return tok, self.decoder.estimator.layout_for(mel_lens)

    def _program(self, tok: SeqLayout, mel: SeqLayout) -> _Program:
        chunk = int(self.encoder.static_chunk_size)
        key = (tok.stride_key(), mel.stride_key(), chunk)
        p = self._programs.pop(key, None)
        if p is None:
            ops.trim_cache(self._programs, self.device, 32)         # LRU by count and memory; holders keep evicted programs alive
            p = _Program(self.weights, tok, mel, self.engine, chunk)
        self._programs[key] = p
        return p

    @torch.inference_mode()
    def encode_batch(self, tokens: Sequence[torch.Tensor], n_prompt: Sequence[int], mel_lens: Sequence[Tuple[int, int]],
                     embeddings: Sequence[torch.Tensor], sample_rate: int = 22050) -> _Program:
        """tokens[b]: int [n_b] prompt + generated tokens of utterance b; mel_lens[b] = (mel_len1, mel_len2).
        Returns the bound program: `.mu` [rows, 80] fp32 over `.mel`, `.spks` [B, 80]."""
        tok_lens = [int(t.numel()) for t in tokens]
        tot = [a + b for a, b in mel_lens]
        tok, mel = self._layouts(tok_lens, tot)
        prog = self._program(tok, mel)
        prog.bind(tok_lens, tot)
        B = len(tokens)
        tk = torch.zeros(B, prog.tokens.shape[1], dtype=torch.int32)
        sg = torch.zeros(B, 4, 4, dtype=torch.int32)
        for b, t in enumerate(tokens):
            tk[b, :tok_lens[b]] = t.reshape(-1).to("cpu", torch.int32)
            if not self.V2:
                pieces = regulator_segments(int(n_prompt[b]), tok_lens[b] - int(n_prompt[b]), mel_lens[b][0], mel_lens[b][1],
                                            self.input_frame_rate, sample_rate)
                sg[b, :len(pieces)] = torch.tensor(pieces, dtype=torch.int32)
        copy_async_(prog.tokens, tk)
        copy_async_(prog.segs, sg)
        if all(not e.is_cuda for e in embeddings):
            copy_async_(prog.spk_in, torch.cat([e.reshape(1, -1).float() for e in embeddings], 0))
        else:
            prog.spk_in.copy_(torch.cat([e.reshape(1, -1).to(self.device, torch.float32) for e in embeddings], 0))
        prog.run()
        return prog

    def _mel_lens(self, n_prompt_tok: int, n_tok: int, prompt_frames: int, sample_rate: int) -> Tuple[int, int]:
        raise NotImplementedError

    @torch.inference_mode()
    def inference_batch(self, tokens: Sequence[torch.Tensor], prompt_tokens: Sequence[torch.Tensor],
                        prompt_feats: Sequence[torch.Tensor], embeddings: Sequence[torch.Tensor],
                        zs: Optional[Sequence[torch.Tensor]] = None, n_timesteps: int = 10, sample_rate: int = 22050
                        ) -> List[torch.Tensor]:
        """Throughput entry: B utterances through front-end + CFM in one varlen batch; per-utterance results equal the
        B = 1 `inference` calls with the same noise (zs[b] [1,80,T_b]; drawn with torch.randn on the device when None).
        Returns the generated mels [1, 80, mel_len2_b] fp32 (device)."""
        with self.serial, self.decoder.estimator.serial:
            return self._inference_batch_locked(tokens, prompt_tokens, prompt_feats, embeddings, zs, n_timesteps, sample_rate)

    def _inference_batch_locked(self, tokens, prompt_tokens, prompt_feats, embeddings, zs, n_timesteps, sample_rate):
        B = len(tokens)
        n_prompt = [int(p.numel()) for p in prompt_tokens]
        mel_lens = [self._mel_lens(n_prompt[b], int(tokens[b].numel()), int(prompt_feats[b].shape[-2]), sample_rate)
                    for b in range(B)]
        cat = [torch.cat([prompt_tokens[b].reshape(-1).to(torch.int32), tokens[b].reshape(-1).to(torch.int32)]) for b in range(B)]
        prog = self.encode_batch(cat, n_prompt, mel_lens, embeddings, sample_rate)
        lay1 = prog.mel
        cfm = self.decoder
        _, st = cfm._state(lay1)
        st["mu"].copy_(prog.mu)
        st["spks"].copy_(prog.spks)
        st["cond"].zero_()
        st["x"].zero_()
        # host inputs go up in ONE pinned, non-blocking copy each (see layout.to_device_async): per-utterance copies from
        # pageable memory would synchronise the stream B times per batch
        prompt_feats = _stage_to_device(prompt_feats, self.device)
        if zs is not None:
            zs = _stage_to_device(zs, self.device)
        for b in range(B):
            r0, (l1, l2) = lay1.row0[b], mel_lens[b]
            if l1:
                st["cond"][r0:r0 + l1].copy_(prompt_feats[b].reshape(-1, N_FEATS)[:l1])
            T = l1 + l2
            if zs is not None:
                z = zs[b].reshape(N_FEATS, T)
            elif self.V2:
                z = cfm.rand_noise[0, :, :T].to(self.device)
            else:
                z = torch.randn(N_FEATS, T, device=self.device)
            st["x"][r0:r0 + T].copy_(z.t())
        cfm.solve_rows(lay1, n_timesteps)
        out = []
        for b in range(B):
            r0, (l1, l2) = lay1.row0[b], mel_lens[b]
            out.append(st["x"][r0 + l1:r0 + l1 + l2].t().unsqueeze(0).contiguous())
        return out

    def _mu_nct(self, prog: _Program, T: int) -> torch.Tensor:
        mu = torch.zeros(1, N_FEATS, T, dtype=torch.float32, device=self.device)
        ops.rows_to_nct(prog.mu, mu, prog.mel)
        return mu


class B200MaskedDiffWithXvec(_FlowBase):
    """Drop-in for `MaskedDiffWithXvec` at inference (cosyvoice/flow/flow.py:25-160, called at cli/model.py:137-145):
    same `inference` signature and return, same attributes the caller touches (`input_frame_rate`, `fp16`, `half()`,
    `decoder.estimator`, `encoder.static_chunk_size`)."""
    V2 = False

    def _mel_lens(self, n_prompt_tok, n_tok, prompt_frames, sample_rate):
        return prompt_frames, int(n_tok / self.input_frame_rate * sample_rate / 256)          # flow.py:140

    @torch.inference_mode()
    def inference(self, token, token_len, prompt_token, prompt_token_len, prompt_feat, prompt_feat_len, embedding,
                  flow_cache, sample_rate, n_timesteps=10):
        if token.shape[0] != 1:
            raise AssertionError("token.shape[0] == 1 (cosyvoice/flow/flow.py:125); use inference_batch for batches")
        n1, n2 = int(prompt_token.shape[1]), int(token.shape[1])
        if int(token_len.reshape(-1)[0]) != n2 or int(prompt_token_len.reshape(-1)[0]) != n1:
            raise ValueError("padded token tensors are not supported: token_len must equal token.shape[1] (B = 1)")
        mel_len1, mel_len2 = self._mel_lens(n1, n2, int(prompt_feat.shape[1]), sample_rate)
        tok = torch.cat([prompt_token.reshape(-1), token.reshape(-1)]).to(torch.int32)
        T = mel_len1 + mel_len2
        with self.serial:
            prog = self.encode_batch([tok], [n1], [(mel_len1, mel_len2)], [embedding], sample_rate)
            mu, spk = self._mu_nct(prog, T), prog.spks[0:1].clone()
        conds = torch.zeros(1, T, N_FEATS, dtype=torch.float32, device=self.device)          # flow.py:144-146
        conds[:, :mel_len1] = prompt_feat.to(self.device, torch.float32)
        mask = torch.ones(1, 1, T, dtype=torch.float32, device=self.device)
        if flow_cache is not None:
            flow_cache = flow_cache.to(self.device, torch.float32)
        feat, flow_cache = self.decoder(mu=mu, mask=mask, spks=spk, cond=conds.transpose(1, 2).contiguous(),
                                        n_timesteps=n_timesteps, prompt_len=mel_len1, flow_cache=flow_cache)
        feat = feat[:, :, mel_len1:]
        assert feat.shape[2] == mel_len2
        return feat, flow_cache


class B200CausalMaskedDiffWithXvec(_FlowBase):
    """Drop-in for `CausalMaskedDiffWithXvec` at inference (cosyvoice/flow/flow.py:163-319, called at cli/model.py:335-342)."""
    V2 = True

    def __init__(self, state_dict: SD, dtype: torch.dtype = torch.bfloat16, device="cuda", input_frame_rate: int = 25,
                 token_mel_ratio: int = 2, pre_lookahead_len: int = 3, estimator_state_dict: Optional[SD] = None):
        super().__init__(state_dict, dtype, device, input_frame_rate, estimator_state_dict)
        if token_mel_ratio != 2 or pre_lookahead_len != 3:
            raise ValueError("UpsampleConformerEncoder is built with stride 2 and pre_lookahead_len 3 (upsample_encoder.py:215,231)")
        self.token_mel_ratio = token_mel_ratio
        self.pre_lookahead_len = pre_lookahead_len

    def _mel_lens(self, n_prompt_tok, n_tok, prompt_frames, sample_rate):
        return prompt_frames, self.token_mel_ratio * (n_prompt_tok + n_tok) - prompt_frames                # flow.py:297

    @torch.inference_mode()
    def inference(self, token, token_len, prompt_token, prompt_token_len, prompt_feat, prompt_feat_len, embedding,
                  flow_cache=None, sample_rate=24000, n_timesteps=10, begin=False, finalize=True):
        if token.shape[0] != 1:
            raise AssertionError("token.shape[0] == 1 (cosyvoice/flow/flow.py:277)")
        # flow.py:279-283: make the prompt a whole number of tokens (mutates the caller's length tensors like the reference)
        if prompt_feat_len[0] % self.token_mel_ratio != 0:
            prompt_feat_len[0] -= prompt_feat_len[0] % self.token_mel_ratio
            prompt_feat = prompt_feat[:, :prompt_feat_len[0], :]
            prompt_token_len[0] = prompt_feat_len[0] // self.token_mel_ratio
            prompt_token = prompt_token[:, :prompt_token_len[0]]
        n1, n2 = int(prompt_token.shape[1]), int(token.shape[1])
        if int(token_len.reshape(-1)[0]) != n2 or int(prompt_token_len.reshape(-1)[0]) != n1:
            raise ValueError("padded token tensors are not supported: token_len must equal token.shape[1] (B = 1)")
        mel_len1, mel_len2 = self._mel_lens(n1, n2, int(prompt_feat.shape[1]), sample_rate)
        tok = torch.cat([prompt_token.reshape(-1), token.reshape(-1)]).to(torch.int32)
        T = mel_len1 + mel_len2
        with self.serial:
            prog = self.encode_batch([tok], [n1], [(mel_len1, mel_len2)], [embedding], sample_rate)
            mu, spk = self._mu_nct(prog, T), prog.spks[0:1].clone()
        conds = torch.zeros(1, T, N_FEATS, dtype=torch.float32, device=self.device)          # flow.py:301-303
        conds[:, :mel_len1] = prompt_feat.to(self.device, torch.float32)
        mask = torch.ones(1, 1, T, dtype=torch.float32, device=self.device)
        feat, _ = self.decoder(mu=mu, mask=mask, spks=spk, cond=conds.transpose(1, 2).contiguous(),
                               n_timesteps=n_timesteps)
        feat = feat[:, :, mel_len1:]
        assert feat.shape[2] == mel_len2
        return feat.float(), None
