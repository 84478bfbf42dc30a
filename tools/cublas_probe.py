"""Which kernels cuBLAS (through torch.matmul) runs for the decode-sized projections, for an ncu capture:
  ncu --profile-from-start off --set full --clock-control none --csv --page raw --log-file out.csv python tools/cublas_probe.py
Only the matmul launches are inside the profiled range."""
import sys

import torch

BF = torch.bfloat16
T = int(sys.argv[1]) if len(sys.argv) > 1 else 64
shapes = [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)]
bufs = []
for (N, K) in shapes:
    w = [torch.randn(N, K, device="cuda", dtype=BF) * 0.02 for _ in range(3)]
    x = torch.randn(T, K, device="cuda", dtype=BF)
    y = torch.empty(T, N, device="cuda", dtype=BF)
    for i in range(3):
        torch.matmul(x, w[i].t(), out=y)           # heuristics + workspace set up outside the range
    bufs.append((w, x, y))
torch.cuda.synchronize()
torch.cuda.profiler.start()
for (w, x, y) in bufs:
    for i in range(2):
        torch.matmul(x, w[i].t(), out=y)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
