#!/bin/bash
# Round 2 (1 GPU): whole suite, bench with the replay-of-recorded-steps value leg, shared-memory carve-out A/B.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_i.log 2>&1; echo "gpu tests rc=$?"; tail -12 gpurun_out/r02_gputests_i.log | cut -c1-300
PROBE_OUT=33 timeout 700 python tools/probe_decode.py '{}' '{"smem_carveout":0}' '{}' '{"smem_carveout":0}' '{}' '{"smem_carveout":0}' > gpurun_out/r02_probe_carveout.jsonl 2> gpurun_out/r02_probe_carveout.err; echo "probe rc=$?"; cat gpurun_out/r02_probe_carveout.jsonl
timeout 700 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r02_j_c3.json 2> gpurun_out/bench_r02_j_c3.err; echo "bench c3 rc=$?"; tail -6 gpurun_out/bench_r02_j_c3.err
