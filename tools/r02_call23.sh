#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py -m gpu -q --timeout 600 -k "attention or attn" -x > gpurun_out/r02_attn_tests_p.log 2>&1; echo "attn tests rc=$?"; tail -5 gpurun_out/r02_attn_tests_p.log | cut -c1-400
timeout 300 python tools/probe_attn_persistent.py > gpurun_out/r02_attn_persistent.jsonl 2> gpurun_out/r02_attn_persistent.err; echo "probe rc=$?"; cat gpurun_out/r02_attn_persistent.jsonl
timeout 600 python tools/probe_decode.py '{}' '{"attn_persistent":1}' '{}' '{"attn_persistent":1}' > gpurun_out/r02_probe_persistent.jsonl 2> gpurun_out/r02_probe_persistent.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_persistent.jsonl
