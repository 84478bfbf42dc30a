#!/bin/bash
# split-K epilogue rewrite: parity (GEMM tests), phase trace, in-engine decode step
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_fp16_gpu.py -m gpu -q --timeout 600 -k "gemm" > gpurun_out/r02_gemm_tests_b.log 2>&1; echo "gemm tests rc=$?"; tail -3 gpurun_out/r02_gemm_tests_b.log | cut -c1-300
rm -f gpurun_out/r02_gemm_trace_b.jsonl
timeout 300 python tools/gemm_trace.py --knob pdl=0 --out gpurun_out/r02_gemm_trace_b.jsonl > gpurun_out/r02_gemm_trace_b.log 2>&1; echo "trace rc=$?"
cut -c1-1400 gpurun_out/r02_gemm_trace_b.jsonl
timeout 600 python tools/probe_decode.py '{}' '{}' > gpurun_out/r02_probe_epilogue.jsonl 2> gpurun_out/r02_probe_epilogue.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_epilogue.jsonl
