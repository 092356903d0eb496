#!/bin/bash
# usage: gpu_check.sh [bench args]   -- GPU parity tests, then a short bench run
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 4 --no-cpu --no-latency "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_quick.err
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_quick.json"))
    r = d["roofline"]
    print("PBS/s %.0f  e2e %.0f  pbs_ms %.1f ks_ms %.1f  frac %.3f of %.1f TF" % (d["value"], (d["e2e"] or {}).get("value", 0), r["kernel_ms"], r["ks_kernel_ms"], r["frac"], r["peak"]))
except Exception as e:
    print("no bench line:", e)
PY
