#!/bin/bash
# Round 2, fifth GPU call (1 GPU): where the TMA attention loses (block-size sweep), interleaved A/B of the GEMM split
# schemes on the real decode step, ncu of the workspace-split GEMM, failing parity test re-run.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -q --timeout 600 -k "baseline or seeded or rejects" > gpurun_out/r02_gputests_e.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r02_gputests_e.log
timeout 400 python tools/probe_attn_bs.py > gpurun_out/r02_attn_block_size.jsonl 2> gpurun_out/r02_attn_block_size.err; echo "bs probe rc=$?"; cat gpurun_out/r02_attn_block_size.jsonl
PROBE_OUT=33 timeout 700 python tools/probe_decode.py '{}' '{"gemm_ws":0}' '{"attn_tma":0}' '{"gemm_ws":0,"attn_tma":0}' '{}' '{"gemm_ws":0}' '{"attn_tma":0}' '{"gemm_ws":0,"attn_tma":0}' '{}' '{"gemm_ws":0,"attn_tma":0}' > gpurun_out/r02_probe_knobs2.jsonl 2> gpurun_out/r02_probe_knobs2.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_knobs2.jsonl
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 100 "$@" > gpurun_out/ncu_plain_$name.log 2>&1 &&
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -o gpurun_out/prof_$name -f "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
}
cap r02_gemm_ws_qkv gemm_tc 3 python tools/bench_ops.py gemm --only 0
