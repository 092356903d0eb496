#!/bin/bash
# usage: gpu_r2_e2e_n.sh N  -- the bench under torchrun at N GPUs (resident + pipelined e2e)
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 8 --no-cpu --no-latency > gpurun_out/bench_tr2_n$N.json 2> gpurun_out/bench_tr2_n$N.err; echo "rc=$?"; tail -2 gpurun_out/bench_tr2_n$N.err
python - $N <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/bench_tr2_n%s.json" % sys.argv[1]))
    print("N=%s resident %.0f (%.2f ms/step)  e2e %.0f (%.2f ms/step) verified %s" % (sys.argv[1], d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["verified"]))
except Exception as e:
    print("no line", e)
PY
