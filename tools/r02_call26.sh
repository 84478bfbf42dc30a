#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --profile-from-start off --set full --clock-control none --csv --page raw --log-file gpurun_out/r02_ncu_attn_persistent.csv python tools/ncu_attn_persistent.py > gpurun_out/r02_ncu_attn_persistent.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/r02_ncu_attn_persistent.log
