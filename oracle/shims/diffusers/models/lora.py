import torch.nn as nn
class LoRACompatibleLinear(nn.Linear):
    """At inference with no LoRA layer attached this is nn.Linear (diffusers 0.27.2 models/lora.py)."""
    pass
