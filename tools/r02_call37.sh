#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_fp16_gpu.py -m gpu -q --timeout 300 -k "gemm" -x > gpurun_out/r02_gemm_tests_w.log 2>&1; echo "gemm tests rc=$?"; tail -12 gpurun_out/r02_gemm_tests_w.log | cut -c1-600
rm -f gpurun_out/r02_gemm_trace_wide.jsonl
for W in 0 1; do
timeout 120 python tools/gemm_trace.py --shapes decode --knob pdl=0 --knob gemm_wide=$W --out gpurun_out/r02_gemm_trace_wide.jsonl > gpurun_out/r02_gemm_trace_wide$W.log 2>&1; echo "trace wide=$W rc=$?"
done
python - <<'P'
import json
for l in open('gpurun_out/r02_gemm_trace_wide.jsonl'):
    d=json.loads(l); u=d['us_from_first_entry']
    print(d['knobs'][-1],d['N'],d['K'],'ctas',d['ctas'],'period',d['graph_us_per_launch'],'first_tma',u['first_tma'][1],'loads_issued',u['loads_issued'][1],'acc_ready',u['acc_ready'],'stored',u['stored'][1])
P
timeout 600 python tools/probe_decode.py '{}' '{"gemm_wide":0}' '{}' '{"gemm_wide":0}' > gpurun_out/r02_probe_wide.jsonl 2> gpurun_out/r02_probe_wide.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_wide.jsonl
