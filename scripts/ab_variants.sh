#!/bin/bash
# usage: ab_variants.sh lib1.so lib2.so ...   -- short device-timed bench of each build variant (HB_LIB), one line each
mkdir -p gpurun_out
for lib in "$@"; do
  name=$(basename "$lib" .so)
  HB_LIB="$PWD/$lib" timeout 600 python bench.py --steps 4 --no-cpu --no-latency --no-e2e > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err
  python - "$name" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/ab_%s.json" % sys.argv[1])); r = d["roofline"]
    print("%-24s PBS/s %.0f  pbs_ms %.2f ks_ms %.2f  frac %.4f of %.2f TF  verified %s" % (sys.argv[1], d["value"], r["kernel_ms"], r["ks_kernel_ms"], r["frac"], r["peak"], d.get("verified")))
except Exception as e:
    print(sys.argv[1], "no bench line:", e)
PY
done
