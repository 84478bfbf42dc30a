#!/bin/bash
mkdir -p gpurun_out
KGE_DEBUG_OCC=1 timeout 300 python tools/probe_attn_persistent.py 2>&1 | grep -m3 "occupancy"
