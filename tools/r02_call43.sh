#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py -m gpu -q --timeout 600 -k "persistent or attention or attn" -x > gpurun_out/r02_attn_tests_t.log 2>&1; echo "attn tests rc=$?"; tail -4 gpurun_out/r02_attn_tests_t.log | cut -c1-300
timeout 900 python -m pytest tests/test_engine_gpu.py tests/test_zz_no_alloc.py -m gpu -q --timeout 600 -x > gpurun_out/r02_engine_tests_t.log 2>&1; echo "engine tests rc=$?"; tail -3 gpurun_out/r02_engine_tests_t.log | cut -c1-300
timeout 200 python tools/probe_attn_persistent.py 2>/dev/null | head -2 | cut -c1-250
