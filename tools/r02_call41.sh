#!/bin/bash
mkdir -p gpurun_out
for M in 1 2 3; do
timeout 60 python tools/gemm_trace.py --shapes decode --knob pdl=0 --knob gemm_wide=$M 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); u=d['us_from_first_entry']; print(d['knobs'][-1],d['N'],d['K'],'ctas',d['ctas'],'period',d['graph_us_per_launch'],'acc_ready',u['acc_ready'][1],'stored',u['stored'][1])"
done
timeout 900 python tools/probe_decode.py '{"gemm_wide":1}' '{"gemm_wide":2}' '{"gemm_wide":3}' '{"gemm_wide":0}' '{"gemm_wide":1}' '{"gemm_wide":2}' '{"gemm_wide":3}' > gpurun_out/r02_probe_wide3.jsonl 2> gpurun_out/r02_probe_wide3.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_wide3.jsonl
