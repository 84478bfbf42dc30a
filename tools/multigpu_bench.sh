#!/bin/bash
# usage: scripts_multigpu.sh N  -> DP-N on C3 and TP-N on C4 (70B), one JSON line each into gpurun_out/
N=$1
mkdir -p gpurun_out
nvidia-smi -L | head -8
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 1 --warmup 3 --no-cpu > gpurun_out/bench_dp${N}_c3.json 2> gpurun_out/bench_dp${N}_c3.err; echo "dp$N rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --workload c4 --steps 1 --warmup 3 --no-cpu > gpurun_out/bench_tp${N}_c4.json 2> gpurun_out/bench_tp${N}_c4.err; echo "tp$N rc=$?"
tail -3 gpurun_out/bench_tp${N}_c4.err
