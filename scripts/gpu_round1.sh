# Round-end measurement set: smoke, GPU parity tests, the default bench line, the reference arm, then (each only after
# its plain command exited 0) the ncu launch list and one full capture of the dominant kernel.  Outputs in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_arm.json 2> gpurun_out/bench_reference_arm.err
timeout 300 python bench.py --width 4736 --steps 2 --no-cpu --no-latency > gpurun_out/plain_small.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --width 4736 --steps 2 --no-cpu --no-latency > gpurun_out/ncu_launch.log 2>&1
timeout 300 python bench.py --width 4736 --steps 2 --no-cpu --no-e2e --no-latency > gpurun_out/plain_small2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pbs16_kernel -s 3 -c 1 -f -o gpurun_out/prof_pbs16 python bench.py --width 4736 --steps 2 --no-cpu --no-e2e --no-latency > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/smoke.log gpurun_out/pytest_gpu.log gpurun_out/bench.err; cat gpurun_out/bench.json
