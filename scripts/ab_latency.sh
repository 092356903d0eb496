#!/bin/bash
# usage: ab_latency.sh lib1.so lib2.so ...   -- c7552 gates-mode / LUT-mode latency legs of bench.py per build variant (HB_LIB)
mkdir -p gpurun_out
for lib in "$@"; do
  name=$(basename "$lib" .so)
  HB_LIB="$PWD/$lib" timeout 600 python bench.py --steps 2 --no-cpu --no-e2e > gpurun_out/lat_$name.json 2> gpurun_out/lat_$name.err
  python - "$name" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/lat_%s.json" % sys.argv[1]))
    g, l = d["c7552_gates_mode"], d["c7552_lut_mode"]
    print("%-24s c7552 gates %.2f ms (verified %s)  c7552_lut_4 %.2f ms (verified %s)" % (sys.argv[1], g["evaluate_ms"], g["outputs_verified"], l["evaluate_ms"], l["outputs_verified"]))
except Exception as e:
    print(sys.argv[1], "no bench line:", e)
PY
done
