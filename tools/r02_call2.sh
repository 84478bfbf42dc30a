#!/bin/bash
# Round 2, second GPU call (2 GPUs): single-GPU suite (the first call stopped at its first failure), tensor-parallel
# tests (exchange kernels + engines, peer memory and NCCL), then bench.py --gpus 2 as the driver launches it (replicas of
# C3 + the config-4 sub-record at TP=2).
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > gpurun_out/gpu2.txt 2>&1
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q --timeout 600 --deselect tests/test_tp_gpu.py > gpurun_out/r02_gputests_b.log 2>&1; echo "gpu tests rc=$?"; tail -15 gpurun_out/r02_gputests_b.log
timeout 900 python -m pytest tests/test_tp_gpu.py -m gpu -q --timeout 700 -x > gpurun_out/r02_tptests.log 2>&1; echo "tp tests rc=$?"; tail -30 gpurun_out/r02_tptests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/bench_r02_b_n2.json 2> gpurun_out/bench_r02_b_n2.err; echo "bench n2 rc=$?"; tail -c 1500 gpurun_out/bench_r02_b_n2.json; tail -5 gpurun_out/bench_r02_b_n2.err
