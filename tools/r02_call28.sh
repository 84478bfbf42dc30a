#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_o.log 2>&1; echo "gpu tests rc=$?"; tail -6 gpurun_out/r02_gputests_o.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02_smoke.log
timeout 900 python bench.py --no-http > gpurun_out/bench_r02_streamk_c3.json 2> gpurun_out/bench_r02_streamk_c3.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_r02_streamk_c3.json
