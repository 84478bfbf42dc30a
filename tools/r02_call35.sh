#!/bin/bash
# Round 2 (2 GPUs): what the driver launches at N = 2 after the GEMM epilogue / attention changes, and the TP tests.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_tp_gpu.py -m gpu -q --timeout 700 > gpurun_out/r02_tptests_n.log 2>&1; echo "tp tests rc=$?"; tail -6 gpurun_out/r02_tptests_n.log | cut -c1-400
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/bench_r02_final2_n2.json 2> gpurun_out/bench_r02_final2_n2.err; echo "bench n2 rc=$?"; tail -c 1200 gpurun_out/bench_r02_final2_n2.json; tail -4 gpurun_out/bench_r02_final2_n2.err
