#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/probe_decode.py '{}' '{"gemm_shallow":1}' '{}' '{"gemm_shallow":1}' > gpurun_out/r02_probe_shallow2.jsonl 2> gpurun_out/r02_probe_shallow2.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_shallow2.jsonl
timeout 900 python bench.py --no-http > gpurun_out/bench_r02_epilogue_c3.json 2> gpurun_out/bench_r02_epilogue_c3.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/bench_r02_epilogue_c3.json
