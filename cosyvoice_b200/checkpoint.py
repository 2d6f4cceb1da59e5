"""Checkpoint wire formats of the reference, read directly.

`CosyVoiceModel.load` (cosyvoice/cli/model.py:72-81) does
    flow.load_state_dict(torch.load(flow_model), strict=False)
    hift.load_state_dict({k.replace('generator.', ''): v for k, v in torch.load(hift_model).items()}, strict=False)
so `flow.pt` is the state_dict of the whole MaskedDiffWithXvec (keys `input_embedding.*`, `spk_embed_affine_layer.*`,
`encoder.*`, `encoder_proj.*`, `length_regulator.*`, `decoder.estimator.*`) and `hift.pt` is either a HiFTGenerator
state_dict or a HiFiGan training wrapper's (`generator.*` + `discriminator.*`).  The functions here cut out what the
hot path needs, fold weight-norm pairs in either key style, check shapes against the architecture and hand the result
to the CUDA back-end.  Host-side logic only: no tensor math beyond reshapes; nothing here runs per utterance.
"""
import os
from typing import Dict, Mapping, Optional, Tuple, Union

import torch

SD = Dict[str, torch.Tensor]
ESTIMATOR_PREFIX = "decoder.estimator."
_WN_SUFFIXES = (("weight_g", "weight_v"), ("parametrizations.weight.original0", "parametrizations.weight.original1"))


def _read(obj: Union[str, os.PathLike, Mapping[str, torch.Tensor]]) -> Mapping[str, torch.Tensor]:
    if isinstance(obj, (str, os.PathLike)):
        obj = torch.load(obj, map_location="cpu", weights_only=True)
    if not isinstance(obj, Mapping):
        raise TypeError(f"expected a state_dict or a path to one, got {type(obj).__name__}")
    # training checkpoints of this fork nest the weights (utils/train_utils.py save_model writes extra scalar entries
    # next to the tensors: keep tensors only)
    return {k: v for k, v in obj.items() if torch.is_tensor(v)}


def flow_estimator_state(flow: Union[str, os.PathLike, Mapping[str, torch.Tensor]]) -> Tuple[SD, bool]:
    """flow.pt (or an already-cut estimator state_dict) -> (estimator state_dict, causal).
    causal (CosyVoice2's CausalConditionalDecoder, flow/decoder.py:59-85) is recognised by its single down block:
    v1 has down_blocks.0 and down_blocks.1 (channels [256, 256]), v2 only down_blocks.0 (channels [256])."""
    sd = _read(flow)
    if any(k.startswith(ESTIMATOR_PREFIX) for k in sd):
        sd = {k[len(ESTIMATOR_PREFIX):]: v for k, v in sd.items() if k.startswith(ESTIMATOR_PREFIX)}
    elif any(k.startswith("estimator.") for k in sd):          # a ConditionalCFM state_dict
        sd = {k[len("estimator."):]: v for k, v in sd.items() if k.startswith("estimator.")}
    if "final_proj.weight" not in sd or "time_mlp.linear_1.weight" not in sd:
        raise KeyError("no ConditionalDecoder weights found (looked for decoder.estimator.* / estimator.* / bare keys)")
    causal = not any(k.startswith("down_blocks.1.") for k in sd)
    want_in = 320
    got = tuple(sd["down_blocks.0.0.block1.block.0.weight"].shape)
    if got[1] != want_in or got[0] != 256:
        raise ValueError(f"estimator input conv has shape {got}; expected [256, {want_in}, 3] (x|mu|spks|cond of 80)")
    if tuple(sd["final_proj.weight"].shape)[:2] != (80, 256):
        raise ValueError(f"final_proj has shape {tuple(sd['final_proj.weight'].shape)}; expected [80, 256, 1]")
    return dict(sd), causal


def hift_state(hift: Union[str, os.PathLike, Mapping[str, torch.Tensor]]) -> Tuple[SD, bool]:
    """hift.pt -> (HiFTGenerator state_dict without the `generator.` prefix, v2).
    Discriminator weights of a training wrapper are dropped; weight-norm pairs stay as they are (packing.fold_weight_norm
    folds either key style).  v2 (upsample_rates [8,5,3], hop 480) is recognised by its third upsampling stage."""
    sd = _read(hift)
    if any(k.startswith("generator.") for k in sd):
        sd = {k.replace("generator.", ""): v for k, v in sd.items() if k.startswith("generator.")}    # model.py:79
    sd = {k: v for k, v in sd.items() if not k.startswith(("discriminator.", "mel_spec_transform."))}
    if not any(k.startswith("conv_pre.") for k in sd) or not any(k.startswith("f0_predictor.") for k in sd):
        raise KeyError("no HiFTGenerator weights found (conv_pre.* / f0_predictor.* missing)")
    v2 = any(k.startswith("ups.2.") for k in sd)
    for prefix in ("conv_pre.", "conv_post.", "ups.0."):
        if prefix + "weight" in sd:
            continue
        if not any(prefix + g in sd and prefix + v in sd for g, v in _WN_SUFFIXES):
            raise KeyError(f"{prefix}: neither a plain weight nor a weight-norm pair")
    return dict(sd), v2


def load_acoustic_backend(flow, hift, dtype: torch.dtype = torch.bfloat16, device="cuda",
                          v2: Optional[bool] = None):
    """flow.pt + hift.pt -> AcousticBackend (flow decoder + vocoder on the CUDA path)."""
    from .pipeline import AcousticBackend
    est_sd, causal = flow_estimator_state(flow)
    hsd, hv2 = hift_state(hift)
    if v2 is None:
        v2 = causal
    if causal != v2 or hv2 != v2:
        raise ValueError(f"checkpoint mismatch: estimator causal={causal}, hift v2={hv2}, requested v2={v2}")
    return AcousticBackend(est_sd, hsd, v2=v2, dtype=dtype, device=device)
