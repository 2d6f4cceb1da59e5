"""Random-init state_dicts with the reference's parameter names (moved to cosyvoice_b200/synthetic.py so that
bench.py's product arm can build weights "of the named architecture" without importing the oracle).  The
generator code and its RNG call order are unchanged: tests/golden/*.pt were produced from these seeds."""
from cosyvoice_b200.synthetic import bigvgan_sd, flow_estimator_sd, hift_sd  # noqa: F401
from cosyvoice_b200.synthetic import flow_frontend_sd, frontend_inputs  # noqa: F401,E402
