#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/r02_gemm_trace_T.jsonl
for T in 16 32 64; do
timeout 300 python tools/gemm_trace.py --T $T --knob pdl=0 --out gpurun_out/r02_gemm_trace_T.jsonl > gpurun_out/r02_gemm_trace_T$T.log 2>&1; echo "trace T=$T rc=$?"
done
python - <<'P'
import json
for l in open('gpurun_out/r02_gemm_trace_T.jsonl'):
    d=json.loads(l); u=d['us_from_first_entry']
    print(d['T'],d['N'],d['K'],d['ctas'],'period',d['graph_us_per_launch'],'first_tma',u['first_tma'][1],'first_tile',u['first_tile'][1],'loads_issued',u['loads_issued'],'acc_ready',u['acc_ready'],'stored',u['stored'])
P
