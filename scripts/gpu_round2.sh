#!/bin/bash
# Round-2 measurement set: the default bench line, the reference arm, then (each only after its plain command exited 0) the ncu launch list of
# the same bench command and full captures of the dominant kernel (full width), of the cluster kernel (a 40-gate level) and of the u64
# tensor-core keyswitch.  Outputs in gpurun_out/; summaries are made here afterwards with scripts/ncu_summary.py / ncu_traffic.py.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_arm.json 2> gpurun_out/bench_reference_arm.err
# launch list of the default bench command (short: 2 steps)
timeout 300 python bench.py --steps 2 --no-cpu --no-latency --no-shortint > gpurun_out/plain_a.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --no-cpu --no-latency --no-shortint > gpurun_out/ncu_launch.log 2>&1
# dominant kernel, full width, one launch
timeout 300 python bench.py --steps 2 --no-cpu --no-e2e --no-latency --no-shortint > gpurun_out/plain_b.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pbs16_kernel -s 3 -c 1 -f -o gpurun_out/prof_pbs16 python bench.py --steps 2 --no-cpu --no-e2e --no-latency --no-shortint > gpurun_out/ncu_full.log 2>&1
# cluster kernel: one 40-gate level of the boolean parameter set
timeout 120 python tests/tools/narrow_sweep.py 40 > gpurun_out/plain_c.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pbs3c_kernel -s 1 -c 1 -f -o gpurun_out/prof_pbs3c python tests/tools/narrow_sweep.py 40 > gpurun_out/ncu_pbs3c.log 2>&1
# u64 tensor-core keyswitch + the u64 throughput kernels: one LUT-mode level
timeout 300 python bench.py --only-shortint > gpurun_out/plain_d.log 2>&1 && \
ncu --set full --clock-control none -k regex:"ks_mma2_kernel|pbs3_kernel|pbs16_kernel" -s 4 -c 4 -f -o gpurun_out/prof_u64 python bench.py --only-shortint > gpurun_out/ncu_u64.log 2>&1
tail -2 gpurun_out/bench.err; head -c 600 gpurun_out/bench.json; ls -la gpurun_out/*.ncu-rep
