#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_fp16_gpu.py -m gpu -q --timeout 300 -k "gemm" -x > gpurun_out/r02_gemm_tests_d.log 2>&1; echo "gemm tests rc=$?"; tail -5 gpurun_out/r02_gemm_tests_d.log | cut -c1-400
rm -f gpurun_out/r02_gemm_trace_pf.jsonl
for PF in 0 2 6 12 24; do
timeout 120 python tools/gemm_trace.py --shapes decode --knob pdl=0 --knob gemm_l2_prefetch=$PF --out gpurun_out/r02_gemm_trace_pf.jsonl > gpurun_out/r02_gemm_trace_pf$PF.log 2>&1; echo "trace pf=$PF rc=$?"
done
python - <<'P'
import json
for l in open('gpurun_out/r02_gemm_trace_pf.jsonl'):
    d=json.loads(l); u=d['us_from_first_entry']
    print(d['knobs'][-1],d['N'],d['K'],'ctas',d['ctas'],'period',d['graph_us_per_launch'],'first_tma',u['first_tma'][1],'first_tile',u['first_tile'][1],'loads_issued',u['loads_issued'],'acc_ready',u['acc_ready'],'stored',u['stored'][1])
P
timeout 600 python tools/probe_decode.py '{}' '{"gemm_l2_prefetch":6}' '{}' '{"gemm_l2_prefetch":12}' > gpurun_out/r02_probe_pf.jsonl 2> gpurun_out/r02_probe_pf.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_pf.jsonl
