"""Stub of onnxruntime: the reference only uses it for isinstance() dispatch
(cosyvoice/flow/flow_matching.py:19,129).  Test infrastructure only."""
class InferenceSession:  # noqa: D401
    pass
class SessionOptions:
    pass
class GraphOptimizationLevel:
    ORT_ENABLE_ALL = 99
