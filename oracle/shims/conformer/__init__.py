"""Stub of the `conformer` package (reference imports ConformerBlock at
cosyvoice/flow/components/decoder.py:6 for a class no config uses)."""
import torch
class ConformerBlock(torch.nn.Module):
    def __init__(self, *a, **k):
        super().__init__()
