#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_tp_gpu.py -m gpu -q --timeout 700 > gpurun_out/r02_tptests_o.log 2>&1; echo "tp tests rc=$?"; tail -5 gpurun_out/r02_tptests_o.log | cut -c1-400
