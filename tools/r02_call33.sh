#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/probe_decode.py '{}' '{"attn_persistent":0}' '{}' '{"attn_persistent":0}' > gpurun_out/r02_probe_persistent4.jsonl 2> gpurun_out/r02_probe_persistent4.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_persistent4.jsonl
