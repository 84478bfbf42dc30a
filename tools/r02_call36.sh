#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/r02_gemm_trace_shape.jsonl
for D in 0 8 0 8; do
timeout 120 python tools/gemm_trace.py --shapes decode --knob pdl=0 --knob gemm_debug=$D --out gpurun_out/r02_gemm_trace_shape.jsonl > gpurun_out/r02_gemm_trace_shape$D.log 2>&1; echo "trace dbg=$D rc=$?"
done
python - <<'P'
import json
for l in open('gpurun_out/r02_gemm_trace_shape.jsonl'):
    d=json.loads(l); u=d['us_from_first_entry']
    print(d['knobs'][-1],d['N'],d['K'],'ctas',d['ctas'],'period',d['graph_us_per_launch'],'first_tma',u['first_tma'][1],'loads_issued',u['loads_issued'][1],'acc_ready',u['acc_ready'],'stored',u['stored'][1])
P
