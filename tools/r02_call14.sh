#!/bin/bash
# decode-GEMM phase trace (our kernel) + what cuBLAS launches for the same shapes
mkdir -p gpurun_out
rm -f gpurun_out/r02_gemm_trace.jsonl
timeout 300 python tools/gemm_trace.py --knob pdl=0 --out gpurun_out/r02_gemm_trace.jsonl > gpurun_out/r02_gemm_trace_pdl0.log 2>&1; echo "trace pdl0 rc=$?"
timeout 300 python tools/gemm_trace.py --out gpurun_out/r02_gemm_trace.jsonl > gpurun_out/r02_gemm_trace_pdl1.log 2>&1; echo "trace pdl1 rc=$?"
timeout 300 python tools/gemm_trace.py --knob pdl=0 --knob gemm_ws=1 --out gpurun_out/r02_gemm_trace.jsonl > gpurun_out/r02_gemm_trace_ws.log 2>&1; echo "trace ws rc=$?"
cat gpurun_out/r02_gemm_trace.jsonl | cut -c1-1500
timeout 600 ncu --profile-from-start off --set full --clock-control none --csv --page raw --log-file gpurun_out/r02_ncu_cublas_decode_gemms.csv python tools/cublas_probe.py > gpurun_out/r02_ncu_cublas.log 2>&1; echo "ncu cublas rc=$?"
tail -3 gpurun_out/r02_ncu_cublas.log
