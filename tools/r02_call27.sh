#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py -m gpu -q --timeout 600 -k "persistent or variants_give" -x > gpurun_out/r02_attn_tests_r.log 2>&1; echo "attn tests rc=$?"; tail -8 gpurun_out/r02_attn_tests_r.log | cut -c1-400
timeout 300 python tools/probe_attn_persistent.py > gpurun_out/r02_attn_persistent3.jsonl 2> gpurun_out/r02_attn_persistent3.err; echo "probe rc=$?"; cat gpurun_out/r02_attn_persistent3.jsonl; tail -3 gpurun_out/r02_attn_persistent3.err
timeout 600 python tools/probe_decode.py '{}' '{"attn_persistent":1}' '{}' '{"attn_persistent":1}' > gpurun_out/r02_probe_persistent3.jsonl 2> gpurun_out/r02_probe_persistent3.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_persistent3.jsonl
