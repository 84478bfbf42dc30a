#!/bin/bash
# four independent accumulators per k-block in the decode GEMMs: parity, trace, in-engine decode step
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_fp16_gpu.py -m gpu -q --timeout 300 -k "gemm" -x > gpurun_out/r02_gemm_tests_e.log 2>&1; echo "gemm tests rc=$?"; tail -5 gpurun_out/r02_gemm_tests_e.log | cut -c1-400
rm -f gpurun_out/r02_gemm_trace_nacc.jsonl
timeout 120 python tools/gemm_trace.py --knob pdl=0 --out gpurun_out/r02_gemm_trace_nacc.jsonl > gpurun_out/r02_gemm_trace_nacc.log 2>&1; echo "trace rc=$?"
python - <<'P'
import json
for l in open('gpurun_out/r02_gemm_trace_nacc.jsonl'):
    d=json.loads(l); u=d['us_from_first_entry']
    print(d['N'],d['K'],'ctas',d['ctas'],'period',d['graph_us_per_launch'],'first_tma',u['first_tma'][1],'first_tile',u['first_tile'][1],'loads_issued',u['loads_issued'],'acc_ready',u['acc_ready'],'stored',u['stored'][1])
P
timeout 600 python tools/probe_decode.py '{}' '{}' > gpurun_out/r02_probe_nacc.jsonl 2> gpurun_out/r02_probe_nacc.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_nacc.jsonl
