def figure(*a, **k):
    raise RuntimeError("matplotlib stub")
