#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_ops_gpu.py -m gpu -q --timeout 300 -k "gemm" -x 2>&1 | tail -2
timeout 60 python tools/gemm_trace.py --shapes decode --knob pdl=0 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); u=d['us_from_first_entry']; print(d['N'],d['K'],'ctas',d['ctas'],'period',d['graph_us_per_launch'],'first_tma',u['first_tma'][1],'acc_ready',u['acc_ready'][1],'stored',u['stored'][1])"
timeout 600 python tools/probe_decode.py '{}' '{}' > gpurun_out/r02_probe_relaxed.jsonl 2>/dev/null; cat gpurun_out/r02_probe_relaxed.jsonl
