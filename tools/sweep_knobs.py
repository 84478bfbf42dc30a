"""One-process sweep of the launch-shape knobs (kge_set_tunable) over the decode-sized kernels of C3, to pick
defaults from measurements.  Every line is a JSON object with the knob settings in "knobs".
  python tools/sweep_knobs.py > gpurun_out/sweep_knobs.jsonl
pdl=1 lets back-to-back launches of the SAME kernel overlap (prologue and tail hidden), pdl=0 is the kernel alone
plus launch gaps: inside a decode step the truth lies between the two."""
import contextlib
import io
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_ops as bo  # noqa: E402

L, kge, BF = bo.L, bo.kge, bo.BF


def knobs(**kw):
    for k, v in kw.items():
        kge.check(L.kge_set_tunable(k.encode(), int(v)), "kge_set_tunable")


def captured(fn, *a):
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        fn(*a)
    return [json.loads(x) for x in buf.getvalue().splitlines() if x.startswith("{")]


def gated(T, f, K):
    n = max(2, int(400e6 // (2 * f * K * 2)) + 1)
    ws = [torch.randn(2 * f, K, device="cuda", dtype=BF) * 0.02 for _ in range(min(n, 8))]
    x = torch.randn(T, K, device="cuda", dtype=BF)
    y = torch.empty(T, f, device="cuda", dtype=BF)

    def run(i):
        kge.check(L.kge_gemm_silu_gate_bf16(y.data_ptr(), x.data_ptr(), ws[i % len(ws)].data_ptr(), T, f, K, None), "gated")
    ms = bo.time_ms(run, 40)
    return {"gated T": T, "f": f, "K": K, "ms": round(ms, 4), "GB/s": round(2.0 * (2 * f * K + T * K + T * f) / ms / 1e6, 1)}


def main():
    kge.require_cuda()
    out = []

    def emit(rows, **kn):
        for r in rows:
            r["knobs"] = kn
            print(json.dumps(r), flush=True)
            out.append(r)
    for pdl in (1, 0):
        for even in (0, 1):
            knobs(pdl=pdl, gemm_even_grid=even)
            emit(captured(bo.bench_gemm, [(64, 28672, 4096), (64, 128256, 4096), (64, 6144, 4096), (64, 4096, 14336)], False),
                 pdl=pdl, gemm_even_grid=even)
            emit([gated(64, 14336, 4096)], pdl=pdl, gemm_even_grid=even)
    knobs(gemm_even_grid=0)
    for pdl in (1, 0):
        for splits in (0, 2, 3, 4, 5, 6, 8):
            knobs(pdl=pdl, attn_decode_splits=splits, attn_split_balance=0)
            emit(captured(bo.bench_attn, [(64, 2176, 32, 8, 128), (64, 2304, 32, 8, 128), (32, 192, 32, 8, 64)]),
                 pdl=pdl, attn_decode_splits=splits)
        knobs(pdl=pdl, attn_decode_splits=0, attn_split_balance=1)
        emit(captured(bo.bench_attn, [(64, 2176, 32, 8, 128), (64, 1024, 32, 8, 128), (32, 192, 32, 8, 64), (8, 8192, 32, 8, 128)]),
             pdl=pdl, attn_split_balance=1)
        knobs(attn_split_balance=0)
        emit(captured(bo.bench_prefill, [(4, 2048, 32, 8, 128), (1, 8192, 32, 8, 128)]), pdl=pdl)
    knobs(pdl=1)


if __name__ == "__main__":
    main()
