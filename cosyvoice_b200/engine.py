"""The estimator behind the reference's third plug-in branch: raw device pointers, ONNX tensor names.

`ConditionalCFM.forward_estimator` (cosyvoice/flow/flow_matching.py:126-165) dispatches on the type of
`self.estimator`: nn.Module, onnxruntime.InferenceSession (host round trip), or -- for anything else -- a TensorRT-style
object: `[context, stream], engine = estimator.acquire_estimator()`, then `context.set_input_shape(name, shape)` for the
six inputs, `context.set_tensor_address(engine.get_tensor_name(i), ptr)` for six inputs + one preallocated output,
`context.execute_async_v3(stream_handle)`, `estimator.release_estimator(context, stream)`.  In the reference that branch
cannot run (`acquire_estimator` is not defined anywhere in the tree).  `B200EstimatorEngine` supplies exactly that
surface on top of the CUDA estimator, with the tensor names and order of bin/export_onnx.py:68-69
(`x, mask, mu, t, spks, cond -> estimator_out`):

    flow.decoder.estimator = B200EstimatorEngine(B200Estimator(est_sd, ...))

Nothing is copied: the seven addresses are wrapped as tensors over the caller's memory (fp32, contiguous, as the
reference passes them) and the result is written into the caller's output buffer on the given stream.
"""
import threading
from typing import Dict, List, Tuple

import torch

TENSOR_NAMES = ("x", "mask", "mu", "t", "spks", "cond", "estimator_out")      # bin/export_onnx.py:68-69
_RANK = {"x": 3, "mask": 3, "mu": 3, "t": 1, "spks": 2, "cond": 3, "estimator_out": 3}


class _DevMem:
    """A caller-owned device buffer seen through __cuda_array_interface__ (fp32, C-contiguous)."""

    def __init__(self, ptr: int, shape: Tuple[int, ...]):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class _Engine:
    """What the reference reads from `trt_engine`: tensor names by binding index (flow_matching.py:159)."""

    num_io_tensors = len(TENSOR_NAMES)

    def get_tensor_name(self, i: int) -> str:
        return TENSOR_NAMES[i]


class ExecutionContext:
    """set_input_shape / set_tensor_address / execute_async_v3 of a TensorRT IExecutionContext, as used at
    flow_matching.py:144-162."""

    def __init__(self, estimator):
        self._est = estimator
        self._shape: Dict[str, Tuple[int, ...]] = {}
        self._addr: Dict[str, int] = {}

    def set_input_shape(self, name: str, shape) -> bool:
        if name not in TENSOR_NAMES[:-1]:
            raise KeyError(f"unknown input tensor '{name}' (inputs are {TENSOR_NAMES[:-1]})")
        shape = tuple(int(s) for s in shape)
        if len(shape) != _RANK[name]:
            raise ValueError(f"'{name}' has rank {_RANK[name]}, got shape {shape}")
        self._shape[name] = shape
        return True

    def set_tensor_address(self, name: str, ptr: int) -> bool:
        if name not in TENSOR_NAMES:
            raise KeyError(f"unknown tensor '{name}'")
        self._addr[name] = int(ptr)
        return True

    def _check(self):
        missing = [n for n in TENSOR_NAMES[:-1] if n not in self._shape] + [n for n in TENSOR_NAMES if n not in self._addr]
        if missing:
            raise RuntimeError(f"execute_async_v3 before all bindings were set: {sorted(set(missing))}")
        N, C, T = self._shape["x"]
        want = {"x": (N, C, T), "mask": (N, 1, T), "mu": (N, C, T), "t": (N,), "spks": (N, C), "cond": (N, C, T)}
        for n, s in want.items():
            if self._shape[n] != s:
                raise ValueError(f"'{n}' has shape {self._shape[n]}, expected {s} for x {self._shape['x']}")
        if C != 80:
            raise ValueError(f"x has {C} channels, the estimator takes 80")
        if any(a == 0 for a in self._addr.values()):
            raise ValueError("null device pointer in the bindings")
        return N, C, T

    def execute_async_v3(self, stream_handle: int) -> bool:
        N, C, T = self._check()
        if N * T == 0:
            return True
        dev = self._est.device
        wrap = lambda n, s: torch.as_tensor(_DevMem(self._addr[n], s), device=dev)
        stream = torch.cuda.ExternalStream(int(stream_handle), device=dev) if stream_handle else torch.cuda.default_stream(dev)
        with torch.cuda.stream(stream):
            out = self._est.forward(wrap("x", (N, C, T)), wrap("mask", (N, 1, T)), wrap("mu", (N, C, T)), wrap("t", (N,)),
                                    wrap("spks", (N, C)), wrap("cond", (N, C, T)))
            wrap("estimator_out", (N, C, T)).copy_(out)
        return True


class B200EstimatorEngine:
    """Deliberately NOT an nn.Module and not an ORT session, so that forward_estimator takes its last branch."""

    def __init__(self, estimator):
        self.estimator = estimator
        self.static_chunk_size = 0                       # written by cli/model.py:50,315
        self._engine = _Engine()
        self._lock = threading.Lock()
        self._free: List[ExecutionContext] = []

    def acquire_estimator(self):
        """-> ([context, stream], engine).  `with stream:` is entered by the caller (flow_matching.py:143): the caller's
        inputs were produced on its current stream, so that stream is handed back (a side stream would need an event
        the reference does not record)."""
        with self._lock:
            ctx = self._free.pop() if self._free else ExecutionContext(self.estimator)
        return [ctx, torch.cuda.stream(torch.cuda.current_stream(self.estimator.device))], self._engine

    def release_estimator(self, context, stream):
        with self._lock:
            self._free.append(context)
