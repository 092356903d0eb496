#!/bin/bash
# async row IO test, the pipelined e2e bench leg, then latency A/B of library variants given as arguments
mkdir -p gpurun_out
python -m pytest tests/test_gpu_levels.py -m gpu -q -k "async" 2>&1 | tail -3
timeout 600 python bench.py --steps 4 --no-cpu --no-latency --no-shortint > gpurun_out/bench_e2e.json 2> gpurun_out/bench_e2e.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_e2e.err
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_e2e.json")); print("resident %.0f  e2e %s" % (d["value"], d["e2e"]))
except Exception as e:
    print("no line", e)
PY
if [ $# -gt 0 ]; then bash scripts/ab_latency.sh helm_b200/libhelm_b200.so "$@"; fi
