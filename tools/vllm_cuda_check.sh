#!/bin/bash
# GPU box only: vLLM's own CUDA path as a third opinion (parity fixtures) and as the beat-this row (C3 throughput).
mkdir -p gpurun_out
rm -f gpurun_out/vllm_cuda*.json gpurun_out/vllm_cuda*.log
timeout 420 python tests/golden/make_vllm_cuda_golden.py --timeout 150 > gpurun_out/vllm_cuda_run.log 2>&1
rc=$?
tail -8 gpurun_out/vllm_cuda_run.log
if [ $rc -eq 0 ]; then
  timeout 330 python tests/golden/make_vllm_cuda_golden.py --perf-only > gpurun_out/vllm_cuda_c3.log 2>&1
  echo "perf rc=$?"; tail -5 gpurun_out/vllm_cuda_c3.log
else
  for f in gpurun_out/vllm_cuda_tiny-llama-hd128.log; do tail -30 $f; done
fi
ls -la gpurun_out | grep vllm
