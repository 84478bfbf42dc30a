#!/bin/bash
# Round 2, sixth GPU call (1 GPU): float16 build, 64-token KV blocks by default, whole suite, C3 bench with the HTTP leg.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_f.log 2>&1; echo "gpu tests rc=$?"; tail -30 gpurun_out/r02_gputests_f.log | cut -c1-300
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r02_f_c3.json 2> gpurun_out/bench_r02_f_c3.err; echo "bench c3 rc=$?"; tail -4 gpurun_out/bench_r02_f_c3.err
timeout 400 python tools/http_bench.py --gpus 1 --workload c3 --jobs 2 --out gpurun_out/r02_http_c3_n1.json 2> gpurun_out/r02_http.err; echo "http rc=$?"; tail -3 gpurun_out/r02_http.err
