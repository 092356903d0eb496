#!/bin/bash
# u64 throughput-kernel A/B: LUT-mode legs of bench.py per build variant (HB_LIB), then the parity tests that run wide u64 levels
mkdir -p gpurun_out
for lib in helm_b200/libhelm_b200.so "$@"; do
  name=$(basename "$lib" .so)
  HB_LIB="$PWD/$lib" timeout 600 python bench.py --only-shortint > gpurun_out/u64_$name.json 2> gpurun_out/u64_$name.err
  python - "$name" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/u64_%s.json" % sys.argv[1]))
    for r in d["lut_mode_throughput"]:
        print("%-16s %-14s PBS/s %.0f pbs_ms %.2f ks_ms %.2f frac %.3f verified %s" % (sys.argv[1], r["params"], r["pbs_per_s"], r["pbs_kernel_ms"], r["ks_kernel_ms"], r["frac_of_fp64_peak"], r["outputs_verified"]))
except Exception as e:
    print(sys.argv[1], "no line:", e)
PY
done
timeout 900 python -m pytest tests/test_gpu_levels.py tests/test_gpu_noise_shortint.py tests/test_gpu_primitives.py tests/test_gpu_circuits.py tests/test_gpu_corpus.py -q -k "shortint or lut or u64 or toy64" > gpurun_out/pytest_u64.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_u64.log
