"""Stub of omegaconf.DictConfig: attribute-access dict
(cosyvoice/flow/flow.py:21, flow_matching.py:30-32)."""
class DictConfig(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e
    def __setattr__(self, k, v):
        self[k] = v
