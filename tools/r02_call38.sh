#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_fp16_gpu.py -m gpu -q --timeout 300 -k "gemm" -x > gpurun_out/r02_gemm_tests_x.log 2>&1; echo "gemm tests rc=$?"; tail -3 gpurun_out/r02_gemm_tests_x.log | cut -c1-300
rm -f gpurun_out/r02_gemm_trace_hybrid.jsonl
timeout 120 python tools/gemm_trace.py --shapes decode --knob pdl=0 --out gpurun_out/r02_gemm_trace_hybrid.jsonl > /dev/null 2>&1; echo "trace rc=$?"
python - <<'P'
import json
for l in open('gpurun_out/r02_gemm_trace_hybrid.jsonl'):
    d=json.loads(l); u=d['us_from_first_entry']
    print(d['N'],d['K'],'ctas',d['ctas'],'period',d['graph_us_per_launch'],'acc_ready',u['acc_ready'],'stored',u['stored'][1])
P
timeout 600 python tools/probe_decode.py '{}' '{"gemm_wide":0}' '{}' '{"gemm_wide":0}' > gpurun_out/r02_probe_wide2.jsonl 2> gpurun_out/r02_probe_wide2.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_wide2.jsonl
