"""Torch (CPU) emulation of the tap-GEMM addressing contract of include/cvb200.h -- TEST INFRASTRUCTURE.
It checks the host-side packing/layout logic without a GPU and is the fp32 checker for the GEMM kernels."""
import torch


def emulate(plan, A, lay, seq_len=None, out_limit=None, a_rows=None, a_cols=None):
    """A: [rows, ld] float tensor following the INPUT layout.  Returns D [M_rows, n_phase, N] with invalid rows zeroed
    (bias added, no activation)."""
    grp = plan.in_group
    A = A.float()
    ld = A.stride(0)                      # row pitch of the underlying buffer (A may be a column slice)
    rows_v = A.shape[0] // grp
    Ap = torch.randn(rows_v * grp, ld)    # columns past the slice: arbitrary finite data (must meet zero weights)
    Ap[:, :A.shape[1]] = A[:rows_v * grp]
    Av = Ap.reshape(rows_v, grp * ld)
    a_rows = a_rows if a_rows is not None else rows_v
    a_cols = a_cols if a_cols is not None else (A.shape[1] if grp == 1 else grp * ld)
    M = lay.rows
    W = plan.W.float().cpu()
    D = torch.zeros(M, plan.n_phase, plan.N)
    m_idx = torch.arange(M)
    for p in range(plan.n_phase):
        for s in range(plan.n_seg):
            r, c = plan.seg_row[p * plan.n_seg + s], plan.seg_col[p * plan.n_seg + s]
            src = m_idx + r
            ok = (src >= 0) & (src < a_rows)
            slab = torch.zeros(M, plan.seg_len)
            c_hi = min(c + plan.seg_len, a_cols, Av.shape[1])
            if c_hi > c:
                slab[ok, :c_hi - c] = Av[src[ok], c:c_hi]
            D[:, p] += slab @ W[p, :plan.N, s * plan.seg_len:(s + 1) * plan.seg_len].t()
    if plan.bias is not None:
        D += plan.bias.float().cpu()[:plan.N].reshape(1, 1, -1)
    slen = lay.lens if seq_len is None else [int(v) for v in seq_len]
    for b in range(lay.n_seq):
        olim = slen[b] * plan.n_phase if out_limit is None else int(out_limit[b])
        for lrow in range(lay.strides[b]):
            for p in range(plan.n_phase):
                if not (lrow < slen[b] and lrow * plan.n_phase + p < olim):
                    D[lay.row0[b] + lrow, p] = 0
    return D


def to_rows(xs, lay, ld):
    """xs: list of [C, T_b] tensors -> [rows, ld] zero padded."""
    out = torch.zeros(lay.rows, ld)
    for b, x in enumerate(xs):
        out[lay.row0[b]:lay.row0[b] + x.shape[1], :x.shape[0]] = x.t()
    return out


def from_rows(buf, lay, C, lens=None):
    lens = lens or lay.lens
    return [buf[lay.row0[b]:lay.row0[b] + lens[b], :C].t() for b in range(lay.n_seq)]
