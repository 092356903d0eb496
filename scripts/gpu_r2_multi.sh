#!/bin/bash
# usage: gpu_r2_multi.sh N   -- multi-GPU tests, then the bench at N GPUs both ways (one process / one rank per GPU)
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py -q -s > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi.log
timeout 600 python bench.py --gpus $N --single-process --steps 4 --no-cpu --no-latency > gpurun_out/bench_sp_n$N.json 2> gpurun_out/bench_sp_n$N.err; echo "sp rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 4 --no-cpu --no-latency > gpurun_out/bench_tr_n$N.json 2> gpurun_out/bench_tr_n$N.err; echo "tr rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?"
tail -12 gpurun_out/pytest_multi.log; tail -3 gpurun_out/bench_sp_n$N.err
python - $N <<'PY'
import json, sys
n = sys.argv[1]
for k in ("sp", "tr", "ref"):
    try:
        d = json.load(open("gpurun_out/bench_%s_n%s.json" % (k, n)))
        r = d.get("roofline") or {}
        print(k, "PBS/s %.0f ms/step %.2f e2e %s kernel_ms %s multi %s" % (d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value"), r.get("kernel_ms"), d.get("multi")))
    except Exception as e:
        print(k, "no line:", e)
PY
