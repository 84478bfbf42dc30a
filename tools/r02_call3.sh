#!/bin/bash
# Round 2, third GPU call (1 GPU): TMA-staged attention kernels - tests, same-box table (graph-timed), C3 bench, ncu.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_c.log 2>&1; echo "gpu tests rc=$?"; tail -25 gpurun_out/r02_gputests_c.log
timeout 700 python tools/bench_vs_libs.py --out gpurun_out/r02_vs_libs.json > gpurun_out/r02_vs_libs.log 2>&1; echo "vs libs rc=$?"; grep -v "^Traceback\|^  File\|^    " gpurun_out/r02_vs_libs.log | tail -20
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r02_c_c3.json 2> gpurun_out/bench_r02_c_c3.err; echo "bench c3 rc=$?"; tail -3 gpurun_out/bench_r02_c_c3.err
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 100 "$@" > gpurun_out/ncu_plain_$name.log 2>&1 &&
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -o gpurun_out/prof_$name -f "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
}
cap r02_attn_decode_tma attn_decode_tma 3 python tools/bench_ops.py attn --only 0
cap r02_attn_tc_tma attn_tc 3 python tools/bench_ops.py prefill --only 0
ls -la gpurun_out | grep -E "prof_r02_attn"
