#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_n.log 2>&1; echo "gpu tests rc=$?"; tail -6 gpurun_out/r02_gputests_n.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02_smoke.log
