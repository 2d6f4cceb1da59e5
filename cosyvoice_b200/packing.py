"""Weight packing: turns the reference's Linear / Conv1d / ConvTranspose1d parameters into the
operand of the tap GEMM (include/cvb200.h: cvb200_gemm).

A convolution over a time-major buffer [rows, c_pad] is a sum of "atoms"
    out[m, n] += sum_c A[m + row_off, col_off + c] * w_tap[n, c]
(one per kernel tap).  Atoms that read adjacent column ranges of the same view row are merged
into segments of a common width `seg_len` (a multiple of 64, the K chunk of the tcgen05 engine);
the packed weight is [n_phase, n_pad, n_seg * seg_len] with zeros wherever a segment is wider
than the taps it carries.

Reference layers covered: nn.Linear, nn.Conv1d (k, dilation, padding, causal left-pad
cosyvoice/flow/decoder.py:59-85, stride cosyvoice/flow/components/decoder.py:65 and
cosyvoice/hifigan/generator.py:290-300) and nn.ConvTranspose1d
(components/decoder.py:142, hifigan/generator.py:268-283, BigVGAN/bigvgan.py:319-334).
"""
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch

from . import _lib

Atom = Tuple[int, int, torch.Tensor]      # (row_off, col_off, w_tap [N, c_in])


def fold_weight_norm(sd: Dict[str, torch.Tensor], prefix: str) -> torch.Tensor:
    """torch.nn.utils.weight_norm(dim=0) pair (old `weight_g/weight_v` or parametrized `original0/original1` keys, as
    the released hift.pt files carry) -> plain weight
    (cosyvoice/hifigan/generator.py:25-26; the reference's own remove_weight_norm is broken, :327)."""
    if prefix + "weight" in sd:
        return sd[prefix + "weight"]
    if prefix + "parametrizations.weight.original1" in sd:      # torch.nn.utils.parametrizations.weight_norm checkpoints
        v, g = sd[prefix + "parametrizations.weight.original1"], sd[prefix + "parametrizations.weight.original0"]
    else:
        v, g = sd[prefix + "weight_v"], sd[prefix + "weight_g"]
    n = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
    return v * (g / n)


@dataclass
class TapPlan:
    W: torch.Tensor                 # [n_phase, n_pad, K] in the run's operand dtype
    bias: Optional[torch.Tensor]    # fp32 [N]
    N: int
    n_pad: int
    n_phase: int
    n_seg: int
    seg_len: int
    bn: int
    seg_row: List[int]
    seg_col: List[int]
    in_group: int = 1               # A is viewed as [rows / in_group, in_group * lda]
    flops_per_row: int = 0          # algorithmic 2*MACs per M-space row (all phases), for the roofline
    extra: dict = field(default_factory=dict)


def _choose_bn(n_pad: int) -> int:
    for bn in range(min(256, n_pad), 15, -16):
        if n_pad % bn == 0:
            return bn
    raise ValueError(n_pad)


def build_plan(phases: List[List[Atom]], N: int, c_pad: int, bias: Optional[torch.Tensor], dtype: torch.dtype,
               device, in_group: int = 1, max_segs: int = _lib.MAX_SEGS) -> TapPlan:
    """phases[p] = atoms of output phase p.  c_pad = column pitch of one tap in the A view."""
    assert c_pad % 64 == 0, "activation buffers are channel-padded to multiples of 64"
    n_phase = len(phases)
    n_pad = (N + 15) // 16 * 16
    macs = 0
    # group atoms by row, require column contiguity at pitch c_pad
    grouped = []
    for atoms in phases:
        rows: Dict[int, List[Atom]] = {}
        for a in atoms:
            rows.setdefault(a[0], []).append(a)
            macs += a[2].shape[0] * a[2].shape[1]
        runs = []
        for r in sorted(rows):
            lst = sorted(rows[r], key=lambda a: a[1])
            run = [lst[0]]
            for a in lst[1:]:
                if a[1] == run[-1][1] + c_pad:
                    run.append(a)
                else:
                    runs.append(run)
                    run = [a]
            runs.append(run)
        grouped.append(runs)
    # smallest taps-per-segment that fits the segment table
    tps = 1
    while True:
        n_seg = max(sum((len(run) + tps - 1) // tps for run in runs) for runs in grouped)
        n_seg = max(n_seg, 1)
        if n_phase * n_seg <= max_segs:
            break
        tps += 1
        assert tps <= 64
    seg_len = tps * c_pad
    K = n_seg * seg_len
    W = torch.zeros(n_phase, n_pad, K, dtype=torch.float32)
    seg_row, seg_col = [], []
    for p, runs in enumerate(grouped):
        s = 0
        for run in runs:
            for i in range(0, len(run), tps):
                chunk = run[i:i + tps]
                seg_row.append(chunk[0][0])
                seg_col.append(chunk[0][1])
                for j, (_, _, w) in enumerate(chunk):
                    k0 = s * seg_len + j * c_pad
                    W[p, :N, k0:k0 + w.shape[1]] = w.float()
                s += 1
        while s < n_seg:                      # dummy zero-weight segments keep n_seg uniform
            seg_row.append(0)
            seg_col.append(0)
            s += 1
    assert all(-32768 <= v < 32768 for v in seg_row + seg_col)
    if bias is not None:                      # the epilogue reads whole 16-column chunks: pad to n_pad
        bias = torch.cat([bias.detach().float().cpu().reshape(-1), torch.zeros(n_pad - N)])
    return TapPlan(W=W.to(device=device, dtype=dtype).contiguous(),
                   bias=None if bias is None else bias.detach().float().to(device).contiguous(),
                   N=N, n_pad=n_pad, n_phase=n_phase, n_seg=n_seg, seg_len=seg_len, bn=_choose_bn(n_pad),
                   seg_row=seg_row, seg_col=seg_col, in_group=in_group, flops_per_row=2 * macs)


def _pad64(c: int) -> int:
    return (c + 63) // 64 * 64


def linear_plan(w: torch.Tensor, bias, dtype, device, c_pad: Optional[int] = None) -> TapPlan:
    """nn.Linear / 1x1 Conv1d: w [N, C] (or [N, C, 1])."""
    w = w.reshape(w.shape[0], -1)
    return build_plan([[(0, 0, w)]], w.shape[0], c_pad or _pad64(w.shape[1]), bias, dtype, device)


def folded_linear_plan(w: torch.Tensor, bias, gamma: torch.Tensor, beta: torch.Tensor, dtype, device) -> TapPlan:
    """nn.Linear that consumes LayerNorm(x; gamma, beta), with the norm folded in (include/cvb200.h, epilogue.row_stats):
    W . LN(x) + b = rstd (Wf . x) - rstd mean u + bf, Wf = W * gamma, u = rowsum of the ROUNDED Wf (the mean term then cancels
    exactly), bf = b + W . beta.  The plan's A operand is the RAW x; plan.extra['row_u'] is u padded to n_pad."""
    w = w.detach().float().reshape(w.shape[0], -1)
    wf = w * gamma.detach().float().reshape(1, -1)
    bf = w @ beta.detach().float().reshape(-1)
    if bias is not None:
        bf = bf + bias.detach().float().reshape(-1)
    plan = build_plan([[(0, 0, wf)]], w.shape[0], _pad64(w.shape[1]), bf, dtype, device)
    plan.extra["row_u"] = plan.W[0].float().sum(dim=1).contiguous()        # [n_pad] fp32, of the weights as rounded
    return plan


def conv1d_plan(w: torch.Tensor, bias, dtype, device, dilation: int = 1, pad_left: Optional[int] = None,
                stride: int = 1, c_pad: Optional[int] = None, col_pitch: Optional[int] = None) -> TapPlan:
    """nn.Conv1d weight [N, C, k].  pad_left defaults to 'same' ((k-1)*dilation/2); a causal conv
    passes (k-1)*dilation.  stride > 1 reads A through the grouped view [rows/stride, stride*col_pitch]
    (col_pitch = row pitch of the underlying buffer, c_pad unless A is a column slice)."""
    N, Cin, k = w.shape
    c_pad = c_pad or _pad64(Cin)
    if pad_left is None:
        pad_left = (k - 1) * dilation // 2
    atoms: List[Atom] = []
    for j in range(k):
        q = j * dilation - pad_left
        if stride == 1:
            atoms.append((q, 0, w[:, :, j]))
        else:
            assert dilation == 1
            atoms.append((q // stride, (q % stride) * (col_pitch or c_pad), w[:, :, j]))
    return build_plan([atoms], N, c_pad, bias, dtype, device, in_group=stride)


def conv_transpose1d_plan(w: torch.Tensor, bias, dtype, device, stride: int, padding: int, shift: int = 0,
                          c_pad: Optional[int] = None) -> TapPlan:
    """nn.ConvTranspose1d weight [C_in, N, k]: out[i*stride - padding + j] += x[i] * w[:, :, j].
    Output row o' = stride*t + r is produced by phase r at M-row t; `shift` delays the output by
    `shift` rows (used for ReflectionPad1d((1, 0)), cosyvoice/hifigan/generator.py:358-359)."""
    Cin, N, k = w.shape
    c_pad = c_pad or _pad64(Cin)
    phases: List[List[Atom]] = []
    for r in range(stride):
        atoms: List[Atom] = []
        # o = stride*t + r - shift = stride*i - padding + j  with i = t + e  ->  j = r - shift + padding - e*stride
        base = r - shift + padding
        e_hi = base // stride                  # j >= 0
        e = e_hi
        while True:
            j = base - e * stride
            if j >= k:
                break
            atoms.append((e, 0, w[:, :, j].t()))
            e -= 1
        phases.append(atoms)
    return build_plan(phases, N, c_pad, bias, dtype, device)
