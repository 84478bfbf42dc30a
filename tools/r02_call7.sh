#!/bin/bash
# Round 2, seventh GPU call (8 GPUs): tensor-parallel tests at 2/4/8 ranks, Llama-3.1-70B at TP=8 and TP=4 (config 4),
# config 5 over HTTP (8 replicas behind the router).
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/gpu8.txt 2>&1
timeout 1200 python -m pytest tests/test_tp_gpu.py -m gpu -q --timeout 700 -k "kernels[4] or kernels[8] or 4-tiny or 8-tiny or 2-tiny-opt-nccl" > gpurun_out/r02_tptests8.log 2>&1; echo "tp tests rc=$?"; tail -15 gpurun_out/r02_tptests8.log | cut -c1-400
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --workload c4 --steps 2 --warmup 3 > gpurun_out/bench_r02_tp8_c4_70b.json 2> gpurun_out/bench_r02_tp8_c4_70b.err; echo "tp8 rc=$?"; tail -c 600 gpurun_out/bench_r02_tp8_c4_70b.json; tail -3 gpurun_out/bench_r02_tp8_c4_70b.err
( CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 4 --workload c4 --steps 2 --warmup 3 > gpurun_out/bench_r02_tp4_c4_70b.json 2> gpurun_out/bench_r02_tp4_c4_70b.err; echo "tp4 rc=$?" ) &
( CUDA_VISIBLE_DEVICES=4,5,6,7 timeout 500 python tools/http_bench.py --gpus 4 --workload c3 --jobs 2 --out gpurun_out/r02_http_c3_n4.json 2> gpurun_out/r02_http4.err; echo "http4 rc=$?" ) &
wait
tail -c 400 gpurun_out/bench_r02_tp4_c4_70b.json; cat gpurun_out/r02_http_c3_n4.json
timeout 500 python tools/http_bench.py --gpus 8 --workload c3 --jobs 2 --out gpurun_out/r02_http_c3_n8.json 2> gpurun_out/r02_http8.err; echo "http8 rc=$?"; cat gpurun_out/r02_http_c3_n8.json
