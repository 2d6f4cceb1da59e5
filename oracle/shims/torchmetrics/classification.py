class MulticlassAccuracy:  # pragma: no cover - import-time placeholder only
    def __init__(self, *a, **k):
        raise NotImplementedError("torchmetrics is not installed; stub for import only")
