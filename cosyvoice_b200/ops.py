"""Thin Python wrappers over the C ABI (include/cvb200.h): they only marshal torch device
pointers, shapes and the current CUDA stream.  No arithmetic happens here.

`gemm_op` / the other `*_op` builders return zero-argument callables with every argument
pre-marshalled, so that a model forward is a flat list of launches (cheap to replay and to
capture in a CUDA graph).
"""
import ctypes as C
import threading
from typing import Callable, Optional

import torch

from . import _lib
from ._lib import BF16, F32
from .layout import SeqLayout
from .packing import TapPlan

Op = Callable[[], None]


# Test-only A/B switch (tests/test_gpu_concurrency.py monkeypatches it to show the race it guards against).  Deliberately
# NOT an environment variable: a leaked variable must not be able to switch the protection off in production.
_NO_SERIAL = False


class Serial:
    """Serialises the use of an object's cached programs (pre-sized buffers + pre-marshalled launches) across host threads
    AND CUDA streams.  The reference allocates per call, so several `tts()` sessions and the LLM thread's stream may be
    inside flow / hift at once (cosyvoice/cli/model.py:62-69,119,183-189; SURVEY.md 8b "Threading"); here a call holds the
    lock while it binds inputs and enqueues its launches, waits (on the device, not the host) for the previous user's
    last launch if that ran on another stream, and leaves an event behind.  Nothing synchronises the device."""

    def __init__(self):
        self._lock = threading.RLock()
        self._event: Optional[torch.cuda.Event] = None

    def __enter__(self):
        if _NO_SERIAL:
            return self
        self._lock.acquire()
        try:
            if self._event is not None:
                torch.cuda.current_stream().wait_event(self._event)
        except BaseException:           # e.g. the caller's stream is capturing: never leave the lock held
            self._lock.release()
            raise
        return self

    def __exit__(self, *exc):
        if _NO_SERIAL:
            return False
        try:
            ev = torch.cuda.Event()
            ev.record()
            self._event = ev
        finally:
            self._lock.release()
        return False

# Program / graph caches are LRU dicts bounded by a count AND by memory: a program for a 48 K-row batch holds ~1.5 GB of
# pre-sized buffers, and a ragged sweep walks through dozens of stride buckets (at a 96 K-row budget the count-only caches
# filled 178 GB).  Before a cache grows, its least-recently-used entries go while the process holds more than
# CACHE_MEMORY_FRACTION of the device's memory; holders of an evicted program keep it alive and consistent.
CACHE_MEMORY_FRACTION = 0.75


def trim_cache(cache: dict, device, max_entries: int, keep: int = 1) -> None:
    dev = torch.device(device)
    limit = CACHE_MEMORY_FRACTION * torch.cuda.get_device_properties(dev).total_memory if dev.type == "cuda" else float("inf")
    while len(cache) >= max_entries or (len(cache) > keep and dev.type == "cuda" and torch.cuda.memory_allocated(dev) > limit):
        if not cache:
            break
        cache.pop(next(iter(cache)))


# bench.py's roofline leg: when a list is installed here, every GEMM / attention launch is bracketed by CUDA
# events on the launching stream and (kind, algorithmic_flops, start, end) is appended to it.
PROFILE = None


def _profiled(kind: str, flops: int, fn):
    def run():
        if PROFILE is None:
            fn()
            return
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        PROFILE.append((kind, flops, e0, e1))
    run.kind = kind          # which kernel family this launch belongs to (tests and bench breakdowns read it)
    return run


def dt_of(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"unsupported dtype {t.dtype}")


def torch_dtype(dt: int) -> torch.dtype:
    return torch.bfloat16 if dt == BF16 else torch.float32


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _require_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("cosyvoice_b200 kernels need CUDA tensors (there is no CPU path)")


def gemm_op(plan: TapPlan, A: torch.Tensor, lay: SeqLayout, out0: torch.Tensor, *, engine: int,
            a_rows: Optional[int] = None, a_cols: Optional[int] = None, seq_len: Optional[torch.Tensor] = None,
            out_limit: Optional[torch.Tensor] = None, bias: bool = True, seq_bias: Optional[torch.Tensor] = None,
            act: int = _lib.ACT_NONE, act_alpha: Optional[torch.Tensor] = None, act_slope: float = 0.0,
            res: Optional[torch.Tensor] = None, out_scale: float = 1.0, accumulate: bool = False,
            out1: Optional[torch.Tensor] = None, out1_mode: int = _lib.OUT1_NONE,
            out1_alpha: Optional[torch.Tensor] = None, out1_slope: float = 0.0,
            gn_stats: Optional[torch.Tensor] = None, row_stats: Optional[torch.Tensor] = None,
            row_u: Optional[torch.Tensor] = None, row_parts: Optional[torch.Tensor] = None,
            row_parts_out: Optional[torch.Tensor] = None, row_eps: float = 1e-5, flops: int = 0) -> Op:
    """A: 2-D time-major tensor (possibly a column slice view) whose rows follow the layout of the
    INPUT signal; `lay` is the M-space layout (output rows / n_phase).  out0/out1/res: 2-D tensors
    whose row r, column block [phase*phase_col, +N) receive M-row r of that phase."""
    _require_cuda(A, out0, plan.W)
    assert A.dim() == 2 and A.stride(1) == 1 and out0.dim() == 2 and out0.stride(1) == 1
    g = _lib.Gemm()
    grp = plan.in_group
    lda = A.stride(0) * grp
    g.A, g.W = A.data_ptr(), plan.W.data_ptr()
    g.tile_seq, g.tile_row0 = lay.tile_seq.data_ptr(), lay.tile_row0.data_ptr()
    g.seq_len = (seq_len if seq_len is not None else lay.seq_len).data_ptr()
    g.out_limit = _ptr(out_limit)
    g.lda = lda
    g.a_rows = a_rows if a_rows is not None else A.shape[0] // grp
    g.a_cols = a_cols if a_cols is not None else (A.shape[1] if grp == 1 else lda)
    g.n_pad, g.N, g.n_phase, g.n_seg, g.seg_len = plan.n_pad, plan.N, plan.n_phase, plan.n_seg, plan.seg_len
    g.dt = dt_of(A)
    assert plan.W.dtype == A.dtype, (plan.W.dtype, A.dtype)
    g.m_tiles = lay.m_tiles
    g.bn = plan.bn
    for i, (r, c) in enumerate(zip(plan.seg_row, plan.seg_col)):
        g.seg_row[i], g.seg_col[i] = r, c
    e = g.epi
    e.bias = _ptr(plan.bias) if bias else None
    e.seq_bias = _ptr(seq_bias)
    e.seq_bias_ld = seq_bias.stride(0) if seq_bias is not None else 0
    e.act, e.act_alpha, e.act_slope = act, _ptr(act_alpha), act_slope
    if res is not None:
        assert res.stride(1) == 1
        e.res, e.ldr = res.data_ptr(), res.stride(0) * plan.n_phase
        e.res_kind = _lib.RES_F32 if res.dtype == torch.float32 else _lib.RES_ACT
        e.res_dt = dt_of(res)
    e.out_scale, e.accumulate = out_scale, int(accumulate)
    # for n_phase > 1 the output buffer [rows*n_phase, ld] is addressed as [rows, n_phase*ld]
    e.out0, e.ldo0, e.out0_dt = out0.data_ptr(), out0.stride(0) * plan.n_phase, dt_of(out0)
    e.phase_col = out0.stride(0) if plan.n_phase > 1 else 0
    if out1 is not None:
        assert out1.stride(1) == 1 and (plan.n_phase == 1 or out1.stride(0) == out0.stride(0))
        e.out1, e.ldo1, e.out1_dt = out1.data_ptr(), out1.stride(0) * plan.n_phase, dt_of(out1)
        e.out1_mode = out1_mode if out1_mode != _lib.OUT1_NONE else _lib.OUT1_COPY
        e.out1_alpha, e.out1_slope = _ptr(out1_alpha), out1_slope
    if res is not None and plan.n_phase > 1:
        assert res.stride(0) == out0.stride(0)
    if gn_stats is not None:
        assert gn_stats.dtype == torch.float64
        e.gn_stats, e.gn_groups = gn_stats.data_ptr(), plan.N // 32
    if row_stats is not None:        # folded LayerNorm (include/cvb200.h): A holds the raw rows, W = W * gamma
        assert row_stats.dtype == torch.float32 and row_stats.is_contiguous() and row_stats.shape[-1] == 2
        assert row_u is not None and row_u.dtype == torch.float32 and row_u.numel() >= plan.n_pad
        e.row_stats, e.row_u = row_stats.data_ptr(), row_u.data_ptr()
    if row_parts is not None:        # folded LayerNorm from the producing GEMM's partial sums [rows, 8, 2]
        assert row_stats is None and row_u is not None and row_u.dtype == torch.float32 and row_u.numel() >= plan.n_pad
        assert row_parts.dtype == torch.float32 and row_parts.is_contiguous() and row_parts.shape[-2:] == (8, 2)
        e.row_parts, e.row_u, e.row_eps = row_parts.data_ptr(), row_u.data_ptr(), row_eps
    if row_parts_out is not None:    # (sum, sum of squares) per 32-column chunk of the stored rows, for the consumer's fold
        assert row_parts_out.dtype == torch.float32 and row_parts_out.is_contiguous() and row_parts_out.shape[-2:] == (8, 2)
        assert plan.N == 256 and plan.n_phase == 1 and gn_stats is None
        e.row_parts_out = row_parts_out.data_ptr()
    lib = _lib.load()
    keep = (plan, A, lay, out0, seq_len, out_limit, seq_bias, act_alpha, res, out1, out1_alpha, gn_stats, row_stats, row_u,
            row_parts, row_parts_out)

    def run(_g=g, _keep=keep):
        _lib.check(lib.cvb200_tapgemm(C.byref(_g), engine, _stream()), "tapgemm")
    return _profiled("tapgemm_tc" if engine == _lib.ENGINE_TCGEN05 else "tapgemm_simt", flops, run)


def ffn_fused_op(a: torch.Tensor, w1: torch.Tensor, b1: torch.Tensor, w2: torch.Tensor, b2: torch.Tensor,
                 x: torch.Tensor, lay: SeqLayout, out1: Optional[torch.Tensor] = None, flops: int = 0) -> Op:
    """x += W2 . gelu(W1 . a + b1) + b2 in one tcgen05 kernel (include/cvb200.h cvb200_ffn_fused)."""
    _require_cuda(a, w1, b1, w2, b2, x, out1)
    assert a.dtype == torch.bfloat16 and w1.dtype == torch.bfloat16 and w2.dtype == torch.bfloat16
    assert x.dtype == torch.float32 and b1.dtype == torch.float32 and b2.dtype == torch.float32
    assert tuple(w1.shape) == (1024, 256) and tuple(w2.shape) == (256, 1024) and w1.is_contiguous() and w2.is_contiguous()
    assert out1 is None or out1.dtype == torch.bfloat16
    lib = _lib.load()
    layc = lay.c
    args = (a.data_ptr(), a.stride(0), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(), x.data_ptr(),
            x.stride(0), _ptr(out1), out1.stride(0) if out1 is not None else 0, C.byref(layc))
    keep = (a, w1, b1, w2, b2, x, out1, lay, layc)

    def launch(_keep=keep):
        _lib.check(lib.cvb200_ffn_fused(*args, _stream()), "ffn_fused")
    return _profiled("gemm", flops, launch)


def fold_layernorm(w: torch.Tensor, b: Optional[torch.Tensor], gamma: torch.Tensor, beta: torch.Tensor, dtype):
    """LayerNorm folded into the Linear that consumes it: W . LN(x) + b = rstd (Wf . x) - rstd mean u + bf with
    Wf = W * gamma (rounded to the operand dtype), u = rowsum(Wf) of the ROUNDED weights (so that the mean term cancels
    exactly), bf = b + W . beta.  Returns (Wf [N, K] in `dtype`, u [N] fp32, bf [N] fp32) on w's device."""
    w32, g32, be32 = w.detach().float(), gamma.detach().float().to(w.device), beta.detach().float().to(w.device)
    wf = (w32 * g32[None, :]).to(dtype)
    u = wf.float().sum(dim=1)
    bf = w32 @ be32
    if b is not None:
        bf = bf + b.detach().float().to(w.device)
    return wf.contiguous(), u.contiguous(), bf.contiguous()


def block_ffn_op(x: torch.Tensor, w1f: torch.Tensor, u1: torch.Tensor, b1f: torch.Tensor, w2: torch.Tensor, b2: torch.Tensor,
                 lay: SeqLayout, out1: Optional[torch.Tensor] = None, stats_out: Optional[torch.Tensor] = None,
                 eps: float = 1e-5, flops: int = 0, attn_out: Optional[torch.Tensor] = None,
                 w_out: Optional[torch.Tensor] = None, b_out: Optional[torch.Tensor] = None) -> Op:
    """x += W2 . gelu(W1 . LN(x) + b1) + b2 on the bf16 residual stream in one tcgen05 kernel, LayerNorm folded into
    (w1f, u1, b1f) by fold_layernorm; optionally stats_out[row] = (rstd, -mean rstd) of the new rows
    (include/cvb200.h cvb200_block_ffn)."""
    _require_cuda(x, w1f, u1, b1f, w2, b2, out1, stats_out)
    assert x.dtype == torch.bfloat16 and w1f.dtype == torch.bfloat16 and w2.dtype == torch.bfloat16
    assert all(t.dtype == torch.float32 and t.is_contiguous() for t in (u1, b1f, b2))
    assert tuple(w1f.shape) == (1024, 256) and tuple(w2.shape) == (256, 1024) and w1f.is_contiguous() and w2.is_contiguous()
    assert out1 is None or out1.dtype == torch.bfloat16
    assert stats_out is None or (stats_out.dtype == torch.float32 and stats_out.is_contiguous()
                                 and stats_out.shape[0] >= lay.rows and stats_out.shape[-1] == 2)
    lib = _lib.load()
    layc = lay.c
    tail = (eps, w1f.data_ptr(), u1.data_ptr(), b1f.data_ptr(), w2.data_ptr(), b2.data_ptr(),
            _ptr(out1), out1.stride(0) if out1 is not None else 0, _ptr(stats_out), C.byref(layc))
    keep = (x, w1f, u1, b1f, w2, b2, out1, stats_out, lay, layc, attn_out, w_out, b_out)
    if attn_out is not None:
        # the attention out-projection + residual add of the block runs first, inside the same kernel (cvb200_block_out_ffn)
        _require_cuda(attn_out, w_out, b_out)
        assert attn_out.dtype == torch.bfloat16 and attn_out.stride(1) == 1 and attn_out.shape[1] >= 512
        assert w_out.dtype == torch.bfloat16 and tuple(w_out.shape) == (256, 512) and w_out.is_contiguous()
        assert b_out.dtype == torch.float32 and b_out.numel() == 256 and b_out.is_contiguous()
        args = (x.data_ptr(), x.stride(0), attn_out.data_ptr(), attn_out.stride(0), w_out.data_ptr(), b_out.data_ptr()) + tail

        def launch(_keep=keep):
            _lib.check(lib.cvb200_block_out_ffn(*args, _stream()), "block_out_ffn")
        return _profiled("block_ffn", flops, launch)
    args = (x.data_ptr(), x.stride(0)) + tail

    def launch(_keep=keep):
        _lib.check(lib.cvb200_block_ffn(*args, _stream()), "block_ffn")
    return _profiled("block_ffn", flops, launch)


def layernorm_op(x, out, gamma, beta, rows: int, eps: float = 1e-5) -> Op:
    _require_cuda(x, out)
    lib = _lib.load()
    keep = (x, out, gamma, beta)
    if x.dtype != torch.float32:       # residual stream in the operand dtype
        args = (x.data_ptr(), x.stride(0), dt_of(x), out.data_ptr(), out.stride(0), dt_of(out), gamma.data_ptr(),
                beta.data_ptr(), rows, x.shape[1], eps)
        return lambda _k=keep: _lib.check(lib.cvb200_layernorm_act(*args, _stream()), "layernorm_act")
    args = (x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), dt_of(out), gamma.data_ptr(), beta.data_ptr(),
            rows, x.shape[1], eps)
    return lambda _k=keep: _lib.check(lib.cvb200_layernorm(*args, _stream()), "layernorm")


def norm_mish_op(h, out, gamma, beta, lay: SeqLayout, gn_stats=None, seq_bias=None, eps: float = 1e-5) -> Op:
    _require_cuda(h, out)
    lib = _lib.load()
    args = (h.data_ptr(), h.stride(0), dt_of(h), out.data_ptr(), out.stride(0), dt_of(out), gamma.data_ptr(),
            beta.data_ptr(), _ptr(gn_stats), _ptr(seq_bias), seq_bias.stride(0) if seq_bias is not None else 0,
            C.byref(lay.c), h.shape[1], eps)
    keep = (h, out, gamma, beta, lay, gn_stats, seq_bias)
    return lambda _k=keep: _lib.check(lib.cvb200_norm_mish(*args, _stream()), "norm_mish")


def attention_op(qkv, out, lay: SeqLayout, heads: int, scale: float, flops: int = 0,
                 impl: int = _lib.ATTN_TCGEN05_V2) -> Op:
    _require_cuda(qkv, out)
    lib = _lib.load()
    assert qkv.dtype == out.dtype
    args = (qkv.data_ptr(), qkv.stride(0), out.data_ptr(), out.stride(0), dt_of(qkv), C.byref(lay.c), heads, scale, impl)
    keep = (qkv, out, lay)
    return _profiled("attention", flops,
                     lambda _k=keep: _lib.check(lib.cvb200_attention_ex(*args, _stream()), "attention"))


def time_sinusoid(t: torch.Tensor, out: torch.Tensor, scale: float = 1000.0):
    _require_cuda(t, out)
    _lib.check(_lib.load().cvb200_time_sinusoid(t.data_ptr(), t.shape[0], out.shape[1], scale, out.data_ptr(), _stream()),
               "time_sinusoid")


def small_linear(x, W, b, y, pre_act=_lib.ACT_NONE, post_act=_lib.ACT_NONE):
    _require_cuda(x, W, y)
    assert x.dtype == W.dtype == y.dtype == torch.float32 and x.is_contiguous() and W.is_contiguous() and y.is_contiguous()
    _lib.check(_lib.load().cvb200_small_linear(x.data_ptr(), x.shape[0], x.shape[1], W.data_ptr(), _ptr(b), W.shape[0],
                                               pre_act, post_act, y.data_ptr(), _stream()), "small_linear")


def pack_cfg_input(x, mu, cond, spks, out, lay1: SeqLayout):
    _require_cuda(x, mu, cond, spks, out)
    _lib.check(_lib.load().cvb200_pack_cfg_input(x.data_ptr(), mu.data_ptr(), cond.data_ptr(), spks.data_ptr(),
                                                 out.data_ptr(), dt_of(out), C.byref(lay1.c), _stream()), "pack_cfg_input")


def euler_cfg(x, v, dt: float, cfg_rate: float):
    _require_cuda(x, v)
    _lib.check(_lib.load().cvb200_euler_cfg(x.data_ptr(), v.data_ptr(), x.numel(), dt, cfg_rate, _stream()), "euler_cfg")


def rowdot_abs(x, w, b, out, rows: int):
    _require_cuda(x, w, out)
    _lib.check(_lib.load().cvb200_rowdot_abs(x.data_ptr(), x.stride(0), dt_of(x), w.data_ptr(), b.data_ptr(), w.numel(),
                                             out.data_ptr(), rows, _stream()), "rowdot_abs")


def sine_source(f0, lay0: SeqLayout, hop: int, sr: float, noise, phase, lin_w, lin_b, prefix_ws, s):
    _require_cuda(f0, noise, phase, s)
    assert noise.is_contiguous() and s.is_contiguous() and noise.shape[-1] == s.shape[-1]
    _lib.check(_lib.load().cvb200_sine_source(f0.data_ptr(), C.byref(lay0.c), hop, sr, noise.data_ptr(), phase.data_ptr(),
                                              lin_w.data_ptr(), lin_b.data_ptr(), prefix_ws.data_ptr(), s.data_ptr(),
                                              s.shape[-1], _stream()), "sine_source")


def stft16(s, out, lay: SeqLayout):
    _require_cuda(s, out)
    _lib.check(_lib.load().cvb200_stft16(s.data_ptr(), s.shape[-1], out.data_ptr(), out.stride(0), dt_of(out),
                                         C.byref(lay.c), _stream()), "stft16")


def istft16(x, lay: SeqLayout, wav, limit: float):
    _require_cuda(x, wav)
    assert x.dtype == torch.float32
    _lib.check(_lib.load().cvb200_istft16(x.data_ptr(), x.stride(0), C.byref(lay.c), wav.data_ptr(), wav.shape[-1], limit,
                                          _stream()), "istft16")


def split3_bf16(x: torch.Tensor, out: torch.Tensor, C_: int, cpad: int):
    """x fp32 [rows, >=C] -> out bf16 [rows, 3*cpad]: three bf16 terms whose sum is x to 24 bits."""
    _require_cuda(x, out)
    assert x.dtype == torch.float32 and out.dtype == torch.bfloat16
    _lib.check(_lib.load().cvb200_split3_bf16(x.data_ptr(), x.stride(0), C_, cpad, out.data_ptr(), out.stride(0),
                                              x.shape[0], _stream()), "split3_bf16")


def snake(x, out, alpha, C_: int, lay: SeqLayout):
    _require_cuda(x, out)
    _lib.check(_lib.load().cvb200_snake(x.data_ptr(), x.stride(0), dt_of(x), out.data_ptr(), out.stride(0), dt_of(out),
                                        alpha.data_ptr(), C_, C.byref(lay.c), _stream()), "snake")


def reflect_row0(x, C_: int, lay: SeqLayout):
    _require_cuda(x)
    _lib.check(_lib.load().cvb200_reflect_row0(x.data_ptr(), x.stride(0), dt_of(x), C_, C.byref(lay.c), _stream()),
               "reflect_row0")


def anti_alias_act_nct(x, up_filter, down_filter, alpha, beta):
    """Drop-in for the reference's fused operator (alias_free_activation/cuda/activation1d.py:24-28):
    x [B, C, T] -> [B, C, T], output allocated here like the reference's torch::empty (.cu:224-225)."""
    _require_cuda(x)
    x = x.contiguous()
    y = torch.empty_like(x)
    B, C_, T = x.shape
    _lib.check(_lib.load().cvb200_anti_alias_act_nct(x.data_ptr(), y.data_ptr(), dt_of(x), B, C_, T,
                                                     up_filter.data_ptr(), down_filter.data_ptr(), alpha.data_ptr(),
                                                     beta.data_ptr(), _stream()), "anti_alias_act_nct")
    return y


def anti_alias_act_rows(x, out, filt, alpha, beta, C_: int, lay: SeqLayout):
    _require_cuda(x, out)
    _lib.check(_lib.load().cvb200_anti_alias_act_rows(x.data_ptr(), x.stride(0), dt_of(x), out.data_ptr(), out.stride(0),
                                                      dt_of(out), filt.data_ptr(), alpha.data_ptr(), beta.data_ptr(), C_,
                                                      C.byref(lay.c), _stream()), "anti_alias_act_rows")


def nct_to_rows(src, out, lay: SeqLayout, C_: Optional[int] = None):
    """src [B, C, T_max] fp32 (any batch/channel strides, unit time stride) -> rows of `out`."""
    _require_cuda(src, out)
    assert src.dtype == torch.float32 and src.stride(2) == 1
    _lib.check(_lib.load().cvb200_nct_to_rows(src.data_ptr(), src.stride(0), src.stride(1), C_ or src.shape[1],
                                              out.data_ptr(), out.stride(0), dt_of(out), C.byref(lay.c), _stream()),
               "nct_to_rows")


def rows_to_nct(src, out, lay: SeqLayout, C_: Optional[int] = None):
    _require_cuda(src, out)
    assert src.dtype == torch.float32 and out.dtype == torch.float32 and out.stride(2) == 1
    _lib.check(_lib.load().cvb200_rows_to_nct(src.data_ptr(), src.stride(0), C_ or out.shape[1], out.data_ptr(),
                                              out.stride(0), out.stride(1), C.byref(lay.c), _stream()), "rows_to_nct")


def fade_in_out(x: torch.Tensor, prev: torch.Tensor, window: torch.Tensor) -> torch.Tensor:
    """In place on x [..., L]: utils/common.py:149-157 with the float64 window [2n] resident on the device.
    x and prev are contiguous with the same leading shape (mel [1,80,T] or speech [1,L])."""
    _require_cuda(x, prev, window)
    assert x.dtype == torch.float32 and prev.dtype == torch.float32 and window.dtype == torch.float64
    assert x.is_contiguous() and prev.is_contiguous() and window.is_contiguous()
    assert x.shape[:-1] == prev.shape[:-1], (x.shape, prev.shape)
    L, Lp = x.shape[-1], prev.shape[-1]
    rows = x.numel() // L if L else 0
    _lib.check(_lib.load().cvb200_fade_in_out(x.data_ptr(), L, L, prev.data_ptr(), Lp, Lp, window.data_ptr(),
                                              window.numel(), rows, _stream()), "fade_in_out")
    return x


def interp_linear(x: torch.Tensor, t_out: int) -> torch.Tensor:
    """x fp32 [1, C, T] -> [1, C, t_out], F.interpolate(mode='linear', align_corners=False) (cli/model.py:171)."""
    _require_cuda(x)
    assert x.dtype == torch.float32 and x.dim() == 3
    x = x.contiguous()
    out = torch.empty(x.shape[0], x.shape[1], t_out, dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().cvb200_interp_linear(x.data_ptr(), x.shape[2], x.shape[2], out.data_ptr(), t_out, t_out,
                                                x.shape[0] * x.shape[1], _stream()), "interp_linear")
    return out


def pcm16(wav: torch.Tensor, scale: float = 32768.0) -> torch.Tensor:
    """fp32 waveform -> int16 PCM on the device (BigVGAN/utils.py:97-101 save_audio's packing)."""
    _require_cuda(wav)
    assert wav.dtype == torch.float32
    wav = wav.contiguous()
    out = torch.empty(wav.shape, dtype=torch.int16, device=wav.device)
    _lib.check(_lib.load().cvb200_pcm16(wav.data_ptr(), out.data_ptr(), scale, wav.numel(), _stream()), "pcm16")
    return out


# ---- flow front-end (csrc/encoder.cu) -------------------------------------------------------------------------------
def embed_tokens_op(tokens: torch.Tensor, table: torch.Tensor, out: torch.Tensor, lay: SeqLayout) -> Op:
    """tokens int32 [n_seq, tok_ld]; table fp32 [V, C]; out rows [rows, >=C] (flow/flow.py:131-135)."""
    _require_cuda(tokens, table, out)
    assert tokens.dtype == torch.int32 and tokens.dim() == 2 and tokens.stride(1) == 1 and table.dtype == torch.float32
    assert table.is_contiguous() and tokens.shape[0] == lay.n_seq
    lib = _lib.load()
    layc = lay.c
    args = (tokens.data_ptr(), tokens.stride(0), table.data_ptr(), table.shape[0], table.shape[1], out.data_ptr(),
            out.stride(0), dt_of(out), C.byref(layc))
    keep = (tokens, table, out, lay, layc)
    return lambda _k=keep: _lib.check(lib.cvb200_embed_tokens(*args, _stream()), "embed_tokens")


def layernorm_seq_op(x, out, gamma, beta, lay: SeqLayout, eps: float) -> Op:
    _require_cuda(x, out, gamma, beta)
    assert x.dtype == torch.float32 and gamma.dtype == torch.float32
    lib = _lib.load()
    layc = lay.c
    args = (x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), dt_of(out), gamma.data_ptr(), beta.data_ptr(),
            C.byref(layc), gamma.numel(), eps)
    keep = (x, out, gamma, beta, lay, layc)
    return lambda _k=keep: _lib.check(lib.cvb200_layernorm_seq(*args, _stream()), "layernorm_seq")


def relpos_attention_op(qkv, pos, center: int, bias_u, bias_v, out, lay: SeqLayout, heads: int, scale: float,
                        chunk: int = 0, flops: int = 0) -> Op:
    """qkv fp32 [rows, 3*heads*64]; pos fp32 [n_rel, heads*64] with PE(r) projected at row center - r."""
    _require_cuda(qkv, pos, bias_u, bias_v, out)
    assert qkv.dtype == torch.float32 and pos.dtype == torch.float32 and bias_u.dtype == torch.float32
    assert bias_u.is_contiguous() and bias_v.is_contiguous() and pos.stride(1) == 1
    lib = _lib.load()
    layc = lay.c
    args = (qkv.data_ptr(), qkv.stride(0), pos.data_ptr(), pos.stride(0), center, pos.shape[0], bias_u.data_ptr(),
            bias_v.data_ptr(), out.data_ptr(), out.stride(0), dt_of(out), C.byref(layc), heads, scale, chunk)
    keep = (qkv, pos, bias_u, bias_v, out, lay, layc)
    return _profiled("relpos_attention", flops,
                     lambda _k=keep: _lib.check(lib.cvb200_relpos_attention(*args, _stream()), "relpos_attention"))


def groupnorm1_mish_op(h, out, gamma, beta, lay: SeqLayout, eps: float = 1e-5) -> Op:
    _require_cuda(h, out, gamma, beta)
    assert h.dtype == torch.float32
    lib = _lib.load()
    layc = lay.c
    args = (h.data_ptr(), h.stride(0), out.data_ptr(), out.stride(0), dt_of(out), gamma.data_ptr(), beta.data_ptr(),
            gamma.numel(), C.byref(layc), eps)
    keep = (h, out, gamma, beta, lay, layc)
    return lambda _k=keep: _lib.check(lib.cvb200_groupnorm1_mish(*args, _stream()), "groupnorm1_mish")


def interp_rows_op(x, src_lay: SeqLayout, out, dst_lay: SeqLayout, segs: torch.Tensor, C_: int) -> Op:
    """segs int32 [n_seq, 4, 4] = (src_start, src_len, dst_start, dst_len) (flow/length_regulator.py:52-70)."""
    _require_cuda(x, out, segs)
    assert x.dtype == torch.float32 and segs.dtype == torch.int32 and segs.is_contiguous()
    assert tuple(segs.shape) == (src_lay.n_seq, 4, 4)
    lib = _lib.load()
    sc, dc = src_lay.c, dst_lay.c
    args = (x.data_ptr(), x.stride(0), C.byref(sc), out.data_ptr(), out.stride(0), dt_of(out), C.byref(dc), segs.data_ptr(), C_)
    keep = (x, out, segs, src_lay, dst_lay, sc, dc)
    return lambda _k=keep: _lib.check(lib.cvb200_interp_rows(*args, _stream()), "interp_rows")


def l2_normalize(x: torch.Tensor, out: torch.Tensor):
    _require_cuda(x, out)
    assert x.dtype == torch.float32 and out.dtype == torch.float32 and x.is_contiguous() and out.is_contiguous()
    _lib.check(_lib.load().cvb200_l2_normalize(x.data_ptr(), out.data_ptr(), x.shape[0], x.shape[1], _stream()), "l2_normalize")
