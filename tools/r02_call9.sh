#!/bin/bash
# Round 2 (1 GPU): stream-K GEMM - tests, graph-timed table, interleaved A/B on the real decode step, ncu; fp16 engines.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_fp16_gpu.py -m gpu -q --timeout 600 -k "gemm or float16_engine" > gpurun_out/r02_gputests_h.log 2>&1; echo "tests rc=$?"; tail -8 gpurun_out/r02_gputests_h.log | cut -c1-300
timeout 500 python tools/bench_vs_libs.py --skip prefill,decode --out gpurun_out/r02_vs_libs_gemm.json > gpurun_out/r02_vs_libs_gemm.log 2>&1; echo "vs libs rc=$?"; grep "^{" gpurun_out/r02_vs_libs_gemm.log | cut -c1-330
PROBE_OUT=33 timeout 700 python tools/probe_decode.py '{}' '{"gemm_ws":1}' '{}' '{"gemm_ws":1}' '{}' '{"gemm_ws":1}' > gpurun_out/r02_probe_streamk.jsonl 2> gpurun_out/r02_probe_streamk.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_streamk.jsonl
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 100 "$@" > gpurun_out/ncu_plain_$name.log 2>&1 &&
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -o gpurun_out/prof_$name -f "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
}
cap r02_gemm_streamk_qkv gemm_tc 3 python tools/bench_ops.py gemm --only 0 --knob gemm_ws=1
cap r02_gemm_streamk_gateup gemm_tc 3 python tools/bench_ops.py gemm --only 2 --knob gemm_ws=1
