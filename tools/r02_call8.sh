#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_g.log 2>&1; echo "gpu tests rc=$?"; tail -30 gpurun_out/r02_gputests_g.log | cut -c1-300
