#!/bin/bash
# Round 2, first GPU call: whole GPU suite (new BASELINE-config parity tests included), bench lines for C3/C2/C1,
# the unmeasured two-CTAs-per-SM GEMM knob, same-box table vs cuBLAS / FlashInfer, ncu --set full of the three
# cluster-split-K decode GEMM shapes.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/r02_gputests.log 2>&1; echo "gpu tests rc=$?"; tail -5 gpurun_out/r02_gputests.log
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r02_a_c3.json 2> gpurun_out/bench_r02_a_c3.err; echo "bench c3 rc=$?"
timeout 300 python bench.py --workload c2 --steps 3 --warmup 3 > gpurun_out/bench_r02_c2.json 2> gpurun_out/bench_r02_c2.err; echo "bench c2 rc=$?"
timeout 300 python bench.py --workload c1 --steps 3 --warmup 3 > gpurun_out/bench_r02_c1.json 2> gpurun_out/bench_r02_c1.err; echo "bench c1 rc=$?"
timeout 300 python tools/probe_decode.py '{}' '{"gemm_shallow":1}' > gpurun_out/r02_probe_shallow.jsonl 2> gpurun_out/r02_probe_shallow.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_shallow.jsonl
timeout 600 python tools/bench_vs_libs.py --out gpurun_out/r02_vs_libs.json > gpurun_out/r02_vs_libs.log 2>&1; echo "vs libs rc=$?"; tail -20 gpurun_out/r02_vs_libs.log
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 100 "$@" > gpurun_out/ncu_plain_$name.log 2>&1 &&
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -o gpurun_out/prof_$name -f "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
}
cap r02_gemm_split_qkv gemm_tc 3 python tools/bench_ops.py gemm --only 0
cap r02_gemm_split_o gemm_tc 3 python tools/bench_ops.py gemm --only 1
cap r02_gemm_split_down gemm_tc 3 python tools/bench_ops.py gemm --only 3
ls -la gpurun_out | grep -E "prof_"
