#!/bin/bash
# Round 2, fourth GPU call (1 GPU): workspace split-K GEMM + lane-parallel TMA producers: tests, library table, knob A/B on
# the real decode step, C3 bench.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_engine_gpu.py tests/test_zz_no_alloc.py -m gpu -q --timeout 600 -x > gpurun_out/r02_gputests_d.log 2>&1; echo "gpu tests rc=$?"; tail -12 gpurun_out/r02_gputests_d.log
timeout 700 python tools/bench_vs_libs.py --out gpurun_out/r02_vs_libs.json > gpurun_out/r02_vs_libs.log 2>&1; echo "vs libs rc=$?"; grep "^{" gpurun_out/r02_vs_libs.log | cut -c1-420
timeout 500 python tools/probe_decode.py '{}' '{"gemm_ws":0}' '{"attn_tma":0}' '{"gemm_ws":0,"attn_tma":0}' > gpurun_out/r02_probe_knobs.jsonl 2> gpurun_out/r02_probe_knobs.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_knobs.jsonl
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r02_d_c3.json 2> gpurun_out/bench_r02_d_c3.err; echo "bench c3 rc=$?"; tail -3 gpurun_out/bench_r02_d_c3.err
