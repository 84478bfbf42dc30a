#!/bin/bash
for name in tiny-llama-tp tiny-llama; do
  rm -f /tmp/ncclid_$name
  for r in 0 1 2 3; do RANK=$r WORLD_SIZE=4 timeout 120 python tools/tp_check.py $name /tmp/ncclid_$name > gpurun_out/tp4_${name}_$r.log 2>&1 & done
  wait
  echo "== $name"; grep -h "^{" gpurun_out/tp4_${name}_0.log; tail -2 gpurun_out/tp4_${name}_0.log | cut -c1-200
done
