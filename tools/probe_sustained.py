"""Burst vs sustained rate of the tcgen05 prefill attention kernel (4 x 2048, llama-3.1-8b heads), alone and
interleaved with a prefill-sized GEMM as inside an engine step."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_ops as bo  # noqa: E402

L, kge, BF = bo.L, bo.kge, bo.BF
B, qlen, H, KV, hd, bs = 4, 2048, 32, 8, 128, 16
nb_seq = qlen // bs
nblk = B * nb_seq + 1
kc = torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF)
vc = torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF)
bt = (torch.arange(B * nb_seq, dtype=torch.int32).view(B, nb_seq) + 1).cuda()
sl = torch.full((B,), qlen, dtype=torch.int32, device="cuda")
qsl = (torch.arange(B + 1, dtype=torch.int32) * qlen).cuda()
qw = (H + 2 * KV) * hd
q = torch.randn(B * qlen, qw, device="cuda", dtype=BF)
out = torch.empty(B * qlen, H * hd, device="cuda", dtype=BF)
T, N, K = 8192, 28672, 4096
w = torch.randn(N, K, device="cuda", dtype=BF) * 0.02
x = torch.randn(T, K, device="cuda", dtype=BF)
y = torch.empty(T, N, device="cuda", dtype=BF)
flops = 4.0 * H * hd * B * qlen * (qlen + 1) / 2


def attn():
    kge.check(L.kge_paged_attn(out.data_ptr(), q.data_ptr(), qw, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb_seq, sl.data_ptr(),
                               qsl.data_ptr(), B, B * qlen, qlen, qlen, H, KV, hd, bs, hd ** -0.5, None, 0, None), "attn")


def gemm():
    kge.check(L.kge_gemm_bf16(y.data_ptr(), x.data_ptr(), w.data_ptr(), None, T, N, K, 0, None, 0, None), "gemm")


def ev():
    return torch.cuda.Event(enable_timing=True)


for iters in (10, 100, 1000, 3000):
    attn(); torch.cuda.synchronize()
    e0, e1 = ev(), ev()
    e0.record()
    for _ in range(iters):
        attn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(json.dumps({"attn alone iters": iters, "ms": round(ms, 4), "TFLOP/s": round(flops / ms / 1e9, 1)}), flush=True)
# interleaved: [gemm, attn] x n, attention timed by its own events
for n in (5, 50, 300):
    pairs = [(ev(), ev()) for _ in range(n)]
    g0, g1 = ev(), ev()
    torch.cuda.synchronize()
    g0.record()
    for a, b in pairs:
        gemm()
        a.record(); attn(); b.record()
    g1.record(); torch.cuda.synchronize()
    ams = sum(a.elapsed_time(b) for a, b in pairs) / n
    tot = g0.elapsed_time(g1) / n
    print(json.dumps({"interleaved n": n, "attn_ms": round(ams, 4), "attn TFLOP/s": round(flops / ams / 1e9, 1),
                      "gemm_ms": round(tot - ams, 4), "gemm TFLOP/s": round(2.0 * T * N * K / (tot - ams) / 1e9, 1)}), flush=True)
