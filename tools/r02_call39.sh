#!/bin/bash
# Round 2 (1 GPU): evidence after the split-K epilogue rewrite and the persistent decode attention: suite, library table,
# ncu of the persistent attention kernel and of the qkv split GEMM, launch list, C3 bench (with the HTTP leg).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_final3.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/r02_gputests_final3.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_final3.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02_smoke_final3.log
timeout 400 python tools/bench_vs_libs.py --out gpurun_out/r02_vs_libs_final3.json > gpurun_out/r02_vs_libs_final3.log 2>&1; echo "vs libs rc=$?"; grep "^{" gpurun_out/r02_vs_libs_final3.log | grep "\"T\": 64\|ctx" | cut -c1-330
timeout 300 python tools/probe_attn_persistent.py > gpurun_out/r02_attn_persistent_final3.jsonl 2>/dev/null; echo "attn probe rc=$?"
rm -f gpurun_out/r02_gemm_trace_final3.jsonl
timeout 120 python tools/gemm_trace.py --knob pdl=0 --out gpurun_out/r02_gemm_trace_final3.jsonl > /dev/null 2>&1; echo "gemm trace rc=$?"
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 100 "$@" > gpurun_out/ncu_plain_$name.log 2>&1 &&
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -o gpurun_out/prof_$name -f "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
}

cap r02_gemm_wide_qkv gemm_wide_kernel 3 python tools/bench_ops.py gemm --only 0
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_c3_v3.csv python tools/ncu_job.py > gpurun_out/ncu_job_r02_v3.log 2>&1; echo "launch list rc=$? rows=$(wc -l < gpurun_out/r02_launches_c3_v3.csv)"
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r02_final3_c3.json 2> gpurun_out/bench_r02_final3_c3.err; echo "bench c3 rc=$?"; tail -3 gpurun_out/bench_r02_final3_c3.err; cut -c1-200 gpurun_out/bench_r02_final3_c3.json
timeout 300 python bench.py --workload c2 --no-http --no-cpu > gpurun_out/bench_r02_final3_c2.json 2> gpurun_out/bench_r02_final3_c2.err; echo "bench c2 rc=$?"; cut -c1-160 gpurun_out/bench_r02_final3_c2.json
timeout 300 python bench.py --workload c1 --no-http --no-cpu > gpurun_out/bench_r02_final3_c1.json 2> gpurun_out/bench_r02_final3_c1.err; echo "bench c1 rc=$?"; cut -c1-160 gpurun_out/bench_r02_final3_c1.json
