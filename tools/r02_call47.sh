#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_final5.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/r02_gputests_final5.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_final5.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02_smoke_final5.log
