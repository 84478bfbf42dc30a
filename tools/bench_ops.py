"""Micro-benchmarks of single kernels through the C ABI with CUDA events (torch events on the
default stream, which is the stream the kernels are launched on when stream=NULL).
  python tools/bench_ops.py gemm|attn [--cublas]
Weights / KV are rotated through buffers larger than L2 so every launch streams from HBM."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kge_loader import load  # noqa: E402

kge = load()
L = kge.lib()
BF = torch.bfloat16


def time_ms(fn, iters):
    for _ in range(3):
        fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def bench_gemm(shapes, cublas):
    rows = []
    for (T, N, K) in shapes:
        ncopies = max(2, int(400e6 // (N * K * 2)) + 1)
        ws = [torch.randn(N, K, device="cuda", dtype=BF) * 0.02 for _ in range(min(ncopies, 12))]
        x = torch.randn(T, K, device="cuda", dtype=BF)
        y = torch.empty(T, N, device="cuda", dtype=BF)
        nws = L.kge_gemm_workspace_bytes(T, N, K)
        wsb = torch.zeros(nws, dtype=torch.uint8, device="cuda")

        def ours(i):
            kge.check(L.kge_gemm_bf16(y.data_ptr(), x.data_ptr(), ws[i % len(ws)].data_ptr(), None, T, N, K, 0,
                                      wsb.data_ptr(), nws, None), "gemm")
        ms = time_ms(ours, 40)
        bytes_ = 2.0 * (N * K + T * K + T * N)
        row = {"T": T, "N": N, "K": K, "ms": round(ms, 4), "GB/s": round(bytes_ / ms / 1e6, 1),
               "TFLOP/s": round(2.0 * T * N * K / ms / 1e9, 1)}
        if cublas:
            ms2 = time_ms(lambda i: torch.matmul(x, ws[i % len(ws)].t(), out=y), 40)
            row.update({"cublas_ms": round(ms2, 4), "cublas_GB/s": round(bytes_ / ms2 / 1e6, 1),
                        "cublas_TFLOP/s": round(2.0 * T * N * K / ms2 / 1e9, 1)})
        rows.append(row)
        print(json.dumps(row), flush=True)
        del ws
    return rows


def bench_attn(cases, bs=16):
    for (B, ctx, H, KV, hd) in cases:
        nb_seq = (ctx + bs - 1) // bs
        nblk = B * nb_seq + 1
        layers = max(2, int(400e6 // (nblk * KV * bs * hd * 4)) + 1)
        kcs = [torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF) for _ in range(min(layers, 6))]
        vcs = [torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF) for _ in range(min(layers, 6))]
        bt = (torch.arange(B * nb_seq, dtype=torch.int32).view(B, nb_seq) + 1).cuda()
        sl = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
        qsl = torch.arange(B + 1, dtype=torch.int32, device="cuda")
        qw = (H + 2 * KV) * hd
        q = torch.randn(B, qw, device="cuda", dtype=BF)
        out = torch.empty(B, H * hd, device="cuda", dtype=BF)
        nws = L.kge_attn_workspace_bytes(B, H, hd, ctx)
        ws = torch.empty(nws, dtype=torch.uint8, device="cuda")

        def run(i):
            kge.check(L.kge_paged_attn(out.data_ptr(), q.data_ptr(), qw, kcs[i % len(kcs)].data_ptr(),
                                       vcs[i % len(vcs)].data_ptr(), bt.data_ptr(), nb_seq, sl.data_ptr(), qsl.data_ptr(),
                                       B, B, 1, ctx, H, KV, hd, bs, hd ** -0.5, ws.data_ptr(), nws, None), "attn")
        ms = time_ms(run, 40)
        bytes_ = B * ctx * KV * hd * 4.0
        print(json.dumps({"B": B, "ctx": ctx, "H": H, "KV": KV, "hd": hd, "ms": round(ms, 4),
                          "GB/s": round(bytes_ / ms / 1e6, 1)}), flush=True)


def bench_prefill(cases, bs=16):
    for (B, qlen, H, KV, hd) in cases:
        nb_seq = (qlen + bs - 1) // bs
        nblk = B * nb_seq + 1
        kc = torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF)
        vc = torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF)
        bt = (torch.arange(B * nb_seq, dtype=torch.int32).view(B, nb_seq) + 1).cuda()
        sl = torch.full((B,), qlen, dtype=torch.int32, device="cuda")
        qsl = (torch.arange(B + 1, dtype=torch.int32) * qlen).cuda()
        qw = (H + 2 * KV) * hd
        q = torch.randn(B * qlen, qw, device="cuda", dtype=BF)
        out = torch.empty(B * qlen, H * hd, device="cuda", dtype=BF)

        def run(i):
            kge.check(L.kge_paged_attn(out.data_ptr(), q.data_ptr(), qw, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb_seq,
                                       sl.data_ptr(), qsl.data_ptr(), B, B * qlen, qlen, qlen, H, KV, hd, bs, hd ** -0.5,
                                       None, 0, None), "attn")
        ms = time_ms(run, 10)
        flops = 4.0 * H * hd * B * qlen * (qlen + 1) / 2
        print(json.dumps({"prefill B": B, "q_len": qlen, "H": H, "KV": KV, "hd": hd, "ms": round(ms, 4),
                          "TFLOP/s": round(flops / ms / 1e9, 1)}), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["gemm", "attn", "prefill"])
    ap.add_argument("--cublas", action="store_true")
    ap.add_argument("--only", type=int, default=-1)
    ap.add_argument("--bs", type=int, default=16, help="KV block size for the attention cases")
    ap.add_argument("--knob", action="append", default=[], help="name=value launch-shape knob (kge_set_tunable), repeatable")
    a = ap.parse_args()
    kge.require_cuda()
    for kv in a.knob:
        k, v = kv.split("=")
        kge.check(L.kge_set_tunable(k.encode(), int(v)), "kge_set_tunable")
    if a.what == "gemm":
        shapes = [(64, 6144, 4096), (64, 4096, 4096), (64, 28672, 4096), (64, 4096, 14336), (64, 128256, 4096),
                  (32, 3072, 2048), (32, 16384, 2048), (32, 2048, 8192),
                  (8192, 6144, 4096), (8192, 4096, 4096), (8192, 28672, 4096), (8192, 4096, 14336)]
        if a.only >= 0:
            shapes = [shapes[a.only]]
        bench_gemm(shapes, a.cublas)
    elif a.what == "prefill":
        cases = [(4, 2048, 32, 8, 128), (1, 8192, 32, 8, 128), (32, 128, 32, 8, 64), (4, 2048, 12, 12, 64)]
        bench_prefill(cases if a.only < 0 else [cases[a.only]], a.bs)
    else:
        cases = [(64, 2176, 32, 8, 128), (64, 1024, 32, 8, 128), (32, 192, 32, 8, 64), (8, 8192, 32, 8, 128)]
        bench_attn(cases if a.only < 0 else [cases[a.only]], a.bs)
