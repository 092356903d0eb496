#!/bin/bash
# usage: ab_latency.sh lib1.so lib2.so ...   -- c7552 gates-mode / LUT-mode latency legs of bench.py per build variant.
# The circuit surface (libhelm_host.so) links the engine by RUNPATH=$ORIGIN, so HB_LIB does not reach it: each variant is
# copied over helm_b200/libhelm_b200.so for its run and the original is restored at the end (run this on the scratch copy
# of a gpurun box).
mkdir -p gpurun_out
cp helm_b200/libhelm_b200.so gpurun_out/.libhelm_b200.orig
for lib in "$@"; do
  name=$(basename "$lib" .so)
  if [ "$lib" != helm_b200/libhelm_b200.so ]; then cp "$lib" helm_b200/libhelm_b200.so; else cp gpurun_out/.libhelm_b200.orig helm_b200/libhelm_b200.so; fi
  timeout 600 python bench.py --steps 2 --no-cpu --no-e2e > gpurun_out/lat_$name.json 2> gpurun_out/lat_$name.err
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
cp gpurun_out/.libhelm_b200.orig helm_b200/libhelm_b200.so; rm -f gpurun_out/.libhelm_b200.orig
