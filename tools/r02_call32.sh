#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py -m gpu -q --timeout 600 -k "persistent or block_size or bench_shapes" -x > gpurun_out/r02_attn_tests_s.log 2>&1; echo "attn tests rc=$?"; tail -4 gpurun_out/r02_attn_tests_s.log | cut -c1-300
timeout 300 python tools/probe_attn_persistent.py > gpurun_out/r02_attn_persistent4.jsonl 2> gpurun_out/r02_attn_persistent4.err; echo "probe rc=$?"; cat gpurun_out/r02_attn_persistent4.jsonl
