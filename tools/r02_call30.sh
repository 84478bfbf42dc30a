#!/bin/bash
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/asp tools/probes/attn_stream_probe.cu -lcuda || exit 1
timeout 60 /tmp/asp > gpurun_out/r02_attn_stream_probe.jsonl; echo "rc=$?"; cat gpurun_out/r02_attn_stream_probe.jsonl
