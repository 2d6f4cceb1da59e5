"""Varlen time-major batch layout shared by every kernel (see include/cvb200.h).

Sequence b occupies rows [row0[b], row0[b] + stride[b]); stride is a multiple of `align`
(itself a multiple of the 128-row GEMM tile) and leaves at least `halo` zero rows after the
valid ones, so the shifted reads of a conv tap never see the neighbouring utterance
(the reference pads each utterance with zeros on its own: B is always 1 there,
cosyvoice/flow/flow.py:125).
"""
from typing import List, Optional, Sequence

import torch

from . import _lib

TILE = 128


def to_device_async(src: torch.Tensor, device, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """Host -> device copy that does NOT synchronise the stream.  A copy from pageable host memory makes CUDA wait for
    everything already queued on the stream before it starts, which serialises the host with the GPU (one such copy per
    batch was enough to keep the host from packing batch i+1 under batch i: 21 ms per batch in the ragged sweep).  The
    source is converted on the host, staged in pinned memory (PyTorch's caching host allocator keeps the block alive
    until the copy has run) and copied with non_blocking=True.  Device tensors pass through."""
    if dtype is not None and src.dtype != dtype:
        src = src.to(dtype)
    if src.is_cuda or torch.device(device).type != "cuda":
        return src.to(device)
    return src.contiguous().pin_memory().to(device, non_blocking=True)


def copy_async_(dst: torch.Tensor, src: torch.Tensor) -> None:
    """dst.copy_(src) for a host `src` through pinned staging (see to_device_async)."""
    if src.is_cuda or not dst.is_cuda:
        dst.copy_(src)
        return
    dst.copy_(src.to(dst.dtype).contiguous().pin_memory(), non_blocking=True)


def round_up(a: int, b: int) -> int:
    return (a + b - 1) // b * b


class SeqLayout:
    def __init__(self, lens: Sequence[int], strides: Sequence[int], device):
        self.lens: List[int] = [int(l) for l in lens]
        self.strides: List[int] = [int(s) for s in strides]
        assert len(self.lens) == len(self.strides) and len(self.lens) > 0
        for l, s in zip(self.lens, self.strides):
            assert s % TILE == 0 and 0 <= l <= s, (l, s)
        self.device = torch.device(device)
        self.n_seq = len(self.lens)
        self.row0: List[int] = []
        r = 0
        for s in self.strides:
            self.row0.append(r)
            r += s
        self.rows = r
        assert self.rows < 2 ** 31
        self.max_len = max(self.lens)
        tile_seq, tile_row0 = [], []
        for b, s in enumerate(self.strides):
            n = s // TILE
            tile_seq += [b] * n
            tile_row0 += [i * TILE for i in range(n)]
        self.m_tiles = len(tile_seq)
        i32 = lambda v: to_device_async(torch.tensor(v, dtype=torch.int32), self.device)
        self.tile_seq = i32(tile_seq)
        self.tile_row0 = i32(tile_row0)
        self.seq_len = i32(self.lens)
        self.seq_row0 = i32(self.row0)
        self._c: Optional[_lib.Layout] = None

    @classmethod
    def build(cls, lens: Sequence[int], device, align: int = TILE, halo: int = 8) -> "SeqLayout":
        assert align % TILE == 0
        return cls(lens, [round_up(int(l) + halo, align) for l in lens], device)

    def derive(self, num: int, den: int, lens: Sequence[int]) -> "SeqLayout":
        """Layout of a resampled signal: every stride scaled by num/den (must stay tile aligned)."""
        strides = []
        for s in self.strides:
            assert (s * num) % den == 0
            strides.append(s * num // den)
        return SeqLayout(lens, strides, self.device)

    def set_lens(self, lens: Sequence[int], halo: int = 8) -> None:
        """Re-bind the valid lengths IN PLACE (strides, row offsets and every device pointer stay): a program or
        CUDA graph built on this layout then serves any batch whose lengths fit the same strides.  The copy is
        stream-ordered on the current stream."""
        lens = [int(l) for l in lens]
        assert len(lens) == self.n_seq
        for l, s in zip(lens, self.strides):
            assert 0 <= l and l + halo <= s, (l, s)
        if lens == self.lens:
            return
        self.lens = lens
        self.max_len = max(lens)
        copy_async_(self.seq_len, torch.tensor(lens, dtype=torch.int32))
        self._c = None

    def with_lens(self, lens: Sequence[int]) -> "SeqLayout":
        return SeqLayout(lens, self.strides, self.device)

    def paired(self) -> "SeqLayout":
        """CFG pair layout: sequences [0,B) conditional followed by [B,2B) unconditional."""
        return SeqLayout(self.lens + self.lens, self.strides + self.strides, self.device)

    @property
    def c(self) -> _lib.Layout:
        if self._c is None:
            self._c = _lib.Layout(self.tile_seq.data_ptr(), self.tile_row0.data_ptr(), self.seq_len.data_ptr(),
                                  self.seq_row0.data_ptr(), self.n_seq, self.rows, self.max_len)
        return self._c

    def key(self):
        return (tuple(self.lens), tuple(self.strides))

    def stride_key(self):
        return tuple(self.strides)
