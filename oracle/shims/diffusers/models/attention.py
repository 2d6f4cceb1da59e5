"""diffusers 0.27.2 models/activations.py::GELU semantics: proj Linear then F.gelu(approximate)."""
import torch.nn as nn
import torch.nn.functional as F
class GELU(nn.Module):
    def __init__(self, dim_in, dim_out, approximate="none", bias=True):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out, bias=bias)
        self.approximate = approximate
    def forward(self, x):
        return F.gelu(self.proj(x), approximate=self.approximate)
class GEGLU(nn.Module):
    def __init__(self, dim_in, dim_out):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)
    def forward(self, x):
        h, g = self.proj(x).chunk(2, dim=-1)
        return h * F.gelu(g)
class ApproximateGELU(nn.Module):
    def __init__(self, dim_in, dim_out):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out)
    def forward(self, x):
        x = self.proj(x)
        return x * (1.702 * x).sigmoid()
class AdaLayerNorm(nn.Module):
    def __init__(self, *a, **k):
        raise NotImplementedError
class AdaLayerNormZero(nn.Module):
    def __init__(self, *a, **k):
        raise NotImplementedError
