#!/bin/bash
# Round 2 (2 GPUs): final validation of what the driver will launch at N = 2, server TP test, min_tokens test, L2-hint A/B.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py tests/test_ops_gpu.py -m gpu -q --timeout 600 -k "min_tokens or l2_eviction" > gpurun_out/r02_gputests_m.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r02_gputests_m.log | cut -c1-300
timeout 900 python -m pytest tests/test_tp_gpu.py -m gpu -q --timeout 700 -k "server or kernels" > gpurun_out/r02_tptests_m.log 2>&1; echo "tp tests rc=$?"; tail -8 gpurun_out/r02_tptests_m.log | cut -c1-400
( CUDA_VISIBLE_DEVICES=1 PROBE_OUT=33 timeout 600 python tools/probe_decode.py '{}' '{"gemm_l2_hints":1}' '{}' '{"gemm_l2_hints":1}' > gpurun_out/r02_probe_l2hints.jsonl 2> gpurun_out/r02_probe_l2hints.err; echo "probe rc=$?" ) 
cat gpurun_out/r02_probe_l2hints.jsonl
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/bench_r02_final_n2.json 2> gpurun_out/bench_r02_final_n2.err; echo "bench n2 rc=$?"; tail -c 900 gpurun_out/bench_r02_final_n2.json; tail -4 gpurun_out/bench_r02_final_n2.err
