#!/bin/bash
# ncu evidence for profiles/: the launch list of one C3 pass (every kernel, gpu__time_duration) and full captures
# of the four kernels that carry the step.  Each profiled command first exits 0 without ncu.
mkdir -p gpurun_out
J="python tools/ncu_job.py"
timeout 120 $J > gpurun_out/ncu_plain_job.json 2> gpurun_out/ncu_plain_job.err &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c3.csv $J > gpurun_out/ncu_job.log 2>&1
echo "launch list rc=$? rows=$(wc -l < gpurun_out/launches_c3.csv)"
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 100 "$@" > gpurun_out/ncu_plain_$name.log 2>&1 &&
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -o gpurun_out/prof_$name -f "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
}
cap gemm_pair_gateup gemm_tc 3 python tools/bench_ops.py gemm --only 10
cap attn_tc_prefill attn_tc 3 python tools/bench_ops.py prefill --only 0
cap attn_decode attn_decode 3 python tools/bench_ops.py attn --only 0
cap gemm_decode_gateup gemm_tc 3 python tools/bench_ops.py gemm --only 2
ls -la gpurun_out | grep -E "prof_|launches_c3"
