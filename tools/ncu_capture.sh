#!/bin/bash
# ncu evidence: launch lists of a bench run (decode + prefill windows) and full captures of the top kernels.
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-cpu"
$B > gpurun_out/ncu_plain_bench.json 2> gpurun_out/ncu_plain_bench.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 6000 -c 700 --csv --log-file gpurun_out/launches_decode.csv $B > gpurun_out/ncu_bench1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -s 500 -c 320 --csv --log-file gpurun_out/launches_prefill.csv $B > gpurun_out/ncu_bench2.log 2>&1
A="python tools/bench_ops.py attn"
$A > gpurun_out/ncu_plain_attn.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_decode -s 3 -c 2 -o gpurun_out/prof_attn_decode -f $A > gpurun_out/ncu_attn.log 2>&1
G="python tools/bench_ops.py gemm --only 2"
$G > gpurun_out/ncu_plain_gemm.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 3 -c 2 -o gpurun_out/prof_gemm_gateup -f $G > gpurun_out/ncu_gemm.log 2>&1
G="python tools/bench_ops.py gemm --only 0"
$G > gpurun_out/ncu_plain_gemm0.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 3 -c 2 -o gpurun_out/prof_gemm_qkv -f $G > gpurun_out/ncu_gemm0.log 2>&1
ls -la gpurun_out | tail -20
