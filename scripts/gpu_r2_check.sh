#!/bin/bash
# round-2 check: smoke, GPU parity tests (with durations), then a short bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
nproc >> gpurun_out/smi.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -s --durations=25 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 4 --no-cpu "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?" >> gpurun_out/bench_quick.err
tail -4 gpurun_out/smoke.log; grep -E "passed|failed|error" gpurun_out/pytest_gpu.log | tail -5; tail -2 gpurun_out/bench_quick.err; cat gpurun_out/bench_quick.json | head -c 3000
