#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/probe_attn_persistent.py 0 1 2 3 4 > gpurun_out/r02_attn_debug.jsonl 2> gpurun_out/r02_attn_debug.err; echo "probe rc=$?"; cat gpurun_out/r02_attn_debug.jsonl; tail -3 gpurun_out/r02_attn_debug.err
