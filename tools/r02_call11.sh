#!/bin/bash
# Round 2 (1 GPU): final evidence - suite, library table (graph-timed, 64-token blocks), ncu of the final attention kernels, C3 bench.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_k.log 2>&1; echo "gpu tests rc=$?"; tail -6 gpurun_out/r02_gputests_k.log | cut -c1-300
timeout 700 python tools/bench_vs_libs.py --out gpurun_out/r02_vs_libs_final.json > gpurun_out/r02_vs_libs_final.log 2>&1; echo "vs libs rc=$?"; grep "^{" gpurun_out/r02_vs_libs_final.log | cut -c1-360
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 100 "$@" > gpurun_out/ncu_plain_$name.log 2>&1 &&
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -o gpurun_out/prof_$name -f "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
}
cap r02_attn_decode_tma_bs64 attn_decode_tma 3 python tools/bench_ops.py attn --only 0 --bs 64
cap r02_attn_tc_tma_bs64 attn_tc 3 python tools/bench_ops.py prefill --only 0 --bs 64
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_c3.csv python tools/ncu_job.py > gpurun_out/ncu_job_r02.log 2>&1; echo "launch list rc=$? rows=$(wc -l < gpurun_out/r02_launches_c3.csv)"
timeout 700 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r02_final_c3.json 2> gpurun_out/bench_r02_final_c3.err; echo "bench c3 rc=$?"; tail -4 gpurun_out/bench_r02_final_c3.err
