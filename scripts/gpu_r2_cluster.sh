#!/bin/bash
# cluster kernel with 1 / 2 / 3 gates per cluster: parity tests, width sweep per setting, latency legs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_levels.py tests/test_gpu_primitives.py tests/test_gpu_circuits.py tests/test_gpu_ready.py -m gpu -x -q > gpurun_out/pytest_cluster.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_cluster.log
for g in 1 2 3; do echo "HB_PBS3C_G=$g"; HB_PBS3C_G=$g python tests/tools/narrow_sweep.py 40 48 60 96 100 120 144 148 2>&1 | grep width; done
for g in 1 2 3; do HB_PBS3C_G=$g timeout 600 python bench.py --steps 2 --no-cpu --no-e2e --no-shortint > gpurun_out/lat_g$g.json 2> gpurun_out/lat_g$g.err
python - $g <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/lat_g%s.json" % sys.argv[1]))
    print("G<=%s: c7552 %.2f ms  c880 %.2f ms  c7552_lut_4 %.2f ms" % (sys.argv[1], d["c7552_gates_mode"]["evaluate_ms"], d["c880_gates_mode"]["evaluate_ms"], d["c7552_lut_mode"]["evaluate_ms"]))
except Exception as e:
    print("no line", e)
PY
done
