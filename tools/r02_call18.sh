#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_fp16_gpu.py -m gpu -q --timeout 600 -k "gemm" -x > gpurun_out/r02_gemm_tests_c.log 2>&1; echo "gemm tests rc=$?"; tail -5 gpurun_out/r02_gemm_tests_c.log | cut -c1-400
rm -f gpurun_out/r02_gemm_trace_rows.jsonl
for R in 0 1; do
timeout 300 python tools/gemm_trace.py --T 64 --knob pdl=0 --knob gemm_rows=$R --out gpurun_out/r02_gemm_trace_rows.jsonl > gpurun_out/r02_gemm_trace_rows$R.log 2>&1; echo "trace rows=$R rc=$?"
done
timeout 300 python tools/gemm_trace.py --T 128 --knob pdl=0 --knob gemm_rows=0 --out gpurun_out/r02_gemm_trace_rows.jsonl > gpurun_out/r02_gemm_trace_T128.log 2>&1; echo "trace T128 rc=$?"
python - <<'P'
import json
for l in open('gpurun_out/r02_gemm_trace_rows.jsonl'):
    d=json.loads(l); u=d['us_from_first_entry']
    print(d['knobs'][-1],d['T'],d['N'],d['K'],'ctas',d['ctas'],'period',d['graph_us_per_launch'],'first_tma',u['first_tma'][1],'first_tile',u['first_tile'][1],'loads_issued',u['loads_issued'],'acc_ready',u['acc_ready'],'stored',u['stored'])
P
timeout 900 python tools/probe_decode.py '{}' '{"gemm_rows":0}' '{}' '{"gemm_rows":0}' > gpurun_out/r02_probe_rows.jsonl 2> gpurun_out/r02_probe_rows.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_rows.jsonl
