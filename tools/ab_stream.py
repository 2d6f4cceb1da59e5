"""A/B of kernel choices on the streaming configuration (BASELINE config 3): runs tools/bench_configs.py `stream`
with the attention implementation forced (--attn 0|1|2) and/or the fused FFN switched off (CVB200_FUSED_FFN=0).
Builder tool; nothing in the product reads these switches except the documented CVB200_FUSED_FFN."""
import functools
import sys

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from cosyvoice_b200 import ops  # noqa: E402

if "--attn" in sys.argv:
    i = sys.argv.index("--attn")
    impl = int(sys.argv[i + 1])
    del sys.argv[i:i + 2]
    _orig = ops.attention_op
    ops.attention_op = functools.partial(_orig, impl=impl)

sys.argv[0] = "bench_configs.py"
import importlib.util as _u  # noqa: E402
_spec = _u.spec_from_file_location("bench_configs", __file__.rsplit("/", 1)[0] + "/bench_configs.py")
bench_configs = _u.module_from_spec(_spec)
_spec.loader.exec_module(bench_configs)

bench_configs.main()
