"""Stub of matplotlib (cosyvoice/BigVGAN/utils.py:7-13 imports it for plotting only)."""
def use(*a, **k):
    pass
