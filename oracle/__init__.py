"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU (torch fp32/fp64) restatement of the reference's acoustic back-end hot path
(duj12/CosyVoice: flow-matching decoder + HiFT / BigVGAN vocoder), used as the
checker for the CUDA path.  Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import this package;
the product package `cosyvoice_b200` never does.

Parity status: the reference ships no golden vectors or tests for this path
(SURVEY.md section 4), so the restatement is pinned against outputs of the
reference's own modules imported from /root/reference in the build container
(`oracle/make_golden.py` -> `tests/golden/*.pt`, checked by
`tests/test_oracle_golden.py`).
"""
