#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_tp_gpu.py -m gpu -q --timeout 900 > gpurun_out/r02_tptests_p4.log 2>&1; echo "tp tests rc=$?"; tail -5 gpurun_out/r02_tptests_p4.log | cut -c1-400
