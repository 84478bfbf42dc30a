#!/bin/bash
# Runs the GPU test groups in separate processes so one faulting kernel cannot poison the rest.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
for grp in "tests/test_ops_gpu.py -k 'not gemm and not attention'" "tests/test_ops_gpu.py -k gemm" "tests/test_ops_gpu.py -k attention" "tests/test_engine_gpu.py"; do
  name=$(echo "$grp" | tr -c 'a-zA-Z0-9' '_' | cut -c1-60)
  eval timeout 600 python -m pytest $grp -m gpu -q --timeout 300 > gpurun_out/t_$name.log 2>&1
  echo "== $grp rc=$?"; tail -${TAILN:-25} gpurun_out/t_$name.log
done
