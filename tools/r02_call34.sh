#!/bin/bash
# Round 2 (1 GPU): evidence after the split-K epilogue rewrite and the persistent decode attention: suite, library table,
# ncu of the persistent attention kernel and of the qkv split GEMM, launch list, C3 bench (with the HTTP leg).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_final2.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/r02_gputests_final2.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_final2.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02_smoke_final2.log
timeout 700 python tools/bench_vs_libs.py --out gpurun_out/r02_vs_libs_final2.json > gpurun_out/r02_vs_libs_final2.log 2>&1; echo "vs libs rc=$?"; grep "^{" gpurun_out/r02_vs_libs_final2.log | cut -c1-330
timeout 300 python tools/probe_attn_persistent.py > gpurun_out/r02_attn_persistent_final.jsonl 2>/dev/null; echo "attn probe rc=$?"
rm -f gpurun_out/r02_gemm_trace_final.jsonl
timeout 120 python tools/gemm_trace.py --knob pdl=0 --out gpurun_out/r02_gemm_trace_final.jsonl > /dev/null 2>&1; echo "gemm trace rc=$?"
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 100 "$@" > gpurun_out/ncu_plain_$name.log 2>&1 &&
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -o gpurun_out/prof_$name -f "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
}
cap r02_attn_decode_persistent attn_decode_tma_persistent 3 python tools/bench_ops.py attn --only 0 --bs 64
cap r02_gemm_split_qkv_v2 gemm_tc_kernel 3 python tools/bench_ops.py gemm --only 0
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_c3_v2.csv python tools/ncu_job.py > gpurun_out/ncu_job_r02_v2.log 2>&1; echo "launch list rc=$? rows=$(wc -l < gpurun_out/r02_launches_c3_v2.csv)"
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r02_final2_c3.json 2> gpurun_out/bench_r02_final2_c3.err; echo "bench c3 rc=$?"; tail -3 gpurun_out/bench_r02_final2_c3.err; cut -c1-200 gpurun_out/bench_r02_final2_c3.json
