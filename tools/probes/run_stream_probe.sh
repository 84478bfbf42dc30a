#!/bin/bash
# builds and runs the streaming micro-benchmark on the GPU box; output: gpurun_out/r02_stream_probe.jsonl
set -e
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/stream_probe tools/probes/stream_probe.cu -lcuda
timeout 300 /tmp/stream_probe > gpurun_out/r02_stream_probe.jsonl
python - <<'P'
import json
rows=[json.loads(l) for l in open('gpurun_out/r02_stream_probe.jsonl')]
keys=[]
for x in rows:
    k=(x['mode'],x['rows'],x['stages'],x['x_rows'])
    if k not in keys: keys.append(k)
print("mode rows stages xrows | " + " | ".join(f"{g:>4} CTAs: TB/s  GB/s/cta" for g in (48,96,128,148)))
for k in keys:
    sel=[x for x in rows if (x['mode'],x['rows'],x['stages'],x['x_rows'])==k]
    print(f"{k[0]:4d} {k[1]:4d} {k[2]:6d} {k[3]:5d} | "+" | ".join(f"{x['TB/s_best']:16.2f} {x['GB/s_per_cta']:9.1f}" for x in sel))
P
