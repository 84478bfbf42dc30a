#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py -m gpu -q --timeout 600 -k "persistent or variants_give" -x > gpurun_out/r02_attn_tests_q.log 2>&1; echo "attn tests rc=$?"; tail -5 gpurun_out/r02_attn_tests_q.log | cut -c1-400
timeout 300 python tools/probe_attn_persistent.py > gpurun_out/r02_attn_persistent2.jsonl 2> gpurun_out/r02_attn_persistent2.err; echo "probe rc=$?"; cat gpurun_out/r02_attn_persistent2.jsonl
