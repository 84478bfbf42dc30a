#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r02_final4_c3.json 2> gpurun_out/bench_r02_final4_c3.err; echo "bench c3 rc=$?"; tail -3 gpurun_out/bench_r02_final4_c3.err | cut -c1-200
timeout 100 python tools/bench_ops.py gemm --only 2 > gpurun_out/ncu_plain_gateup.log 2>&1
timeout 200 ncu --set full --clock-control none --import-source on -k regex:gemm_wide_kernel -s 3 -c 1 -o gpurun_out/prof_r02_gemm_wide_gateup -f python tools/bench_ops.py gemm --only 2 > gpurun_out/ncu_gateup.log 2>&1; echo "ncu rc=$?"
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv
