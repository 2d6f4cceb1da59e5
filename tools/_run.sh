python tools/bench_configs.py stream --from-tokens > gpurun_out/s29_stream25.jsonl 2>/dev/null
python tools/bench_configs.py stream --from-tokens --hop 50 --chunks 5 > gpurun_out/s29_stream50.jsonl 2>/dev/null
python - <<'PY'
import json
for n in ("25", "50"):
    d = json.loads(open(f"gpurun_out/s29_stream{n}.jsonl").read().strip().splitlines()[-1])
    print("hop", n, "p50", round(d["p50"], 2), "p99", round(d["p99"], 2), "audio ms", d["audio_ms_per_chunk"], "front", round(d["frontend_device_ms_last_chunk"], 2), [round(c["p50_ms"], 1) for c in d["per_chunk"]])
PY
