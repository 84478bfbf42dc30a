"""Where a decode-sized GEMM launch spends its time: per-CTA %globaltimer stamps written by the kernel itself
(kge_gemm_set_trace, csrc/gemm_tc.cu) for back-to-back launches on one stream, weights rotated through HBM.

  python tools/gemm_trace.py [--T 64] [--knob gemm_ws=1] [--out gpurun_out/trace.jsonl]

Slots (ns, relative to the earliest CTA entry of the launch):
  0 entry            first instruction of the CTA (thread 0)
  1 setup            barriers initialised, TMEM allocated, CTA synchronised
  2 w_issued         producer: first pipeline stages of W issued (before the dependency wait)
  3 dep              producer: griddepcontrol.wait returned
  4 loads_issued     producer: every TMA of the CTA issued
  5 first_tile       MMA thread: first stage's bytes have landed
  6 mma_issued       MMA thread: last MMA + commit issued
  7 acc_ready        epilogue: accumulator of the CTA's last segment complete
  8 partial_out      split variants: my partial is in the exchange buffer / workspace
  9 exchanged        cluster split: every rank's partial has arrived
 10 stored           epilogue: outputs stored
 11 exit             all warps joined
 12 producer_start   producer thread reaches its first load (after role dispatch, L2 policies)
 13 first_tma        producer: first TMA instruction issued
Per slot: min / median / max over the CTAs of one launch, the median over the traced launches of each."""
import argparse
import json
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kge_loader import load  # noqa: E402

kge = load()
L = kge.lib()
BF = torch.bfloat16
NAMES = ["entry", "setup", "w_issued", "dep", "loads_issued", "first_tile", "mma_issued", "acc_ready", "partial_out",
         "exchanged", "stored", "exit", "producer_start", "first_tma"]
MAX_CTAS = 2048


def trace_shape(T, N, K, gated, launches=8, resident=False):
    rows_w = 2 * N if gated else N
    ncopies = 1 if resident else min(12, max(2, int(400e6 // (rows_w * K * 2)) + 1))
    ws = [torch.randn(rows_w, K, device="cuda", dtype=BF) * 0.02 for _ in range(ncopies)]
    x = torch.randn(T, K, device="cuda", dtype=BF)
    y = torch.empty(T, N, device="cuda", dtype=BF)
    nws = L.kge_gemm_workspace_bytes(T, N, K)
    wsb = torch.zeros(max(nws, 16), dtype=torch.uint8, device="cuda")
    tr = torch.zeros(launches, MAX_CTAS, 16, dtype=torch.int64, device="cuda")

    def launch(i):
        if gated:
            kge.check(L.kge_gemm_silu_gate_ws_bf16(y.data_ptr(), x.data_ptr(), ws[i % ncopies].data_ptr(), T, N, K,
                                                   wsb.data_ptr(), nws, None), "gemm")
        else:
            kge.check(L.kge_gemm_bf16(y.data_ptr(), x.data_ptr(), ws[i % ncopies].data_ptr(), None, T, N, K, 0,
                                      wsb.data_ptr(), nws, None), "gemm")
    for i in range(4):
        launch(i)
    torch.cuda.synchronize()
    # one graph of `launches` traced launches: back to back on the device, no host launch cost between them
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for i in range(launches):
                L.kge_gemm_set_trace(tr[i].data_ptr())
                st = torch.cuda.current_stream().cuda_stream
                if gated:
                    kge.check(L.kge_gemm_silu_gate_ws_bf16(y.data_ptr(), x.data_ptr(), ws[i % ncopies].data_ptr(), T, N, K,
                                                           wsb.data_ptr(), nws, st), "gemm")
                else:
                    kge.check(L.kge_gemm_bf16(y.data_ptr(), x.data_ptr(), ws[i % ncopies].data_ptr(), None, T, N, K, 0,
                                              wsb.data_ptr(), nws, st), "gemm")
            L.kge_gemm_set_trace(None)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    per_launch_us = e0.elapsed_time(e1) * 1e3 / launches
    t = tr.cpu().numpy()
    out = {"T": T, "N": N, "K": K, "gated": gated, "graph_us_per_launch": round(per_launch_us, 2)}
    per_slot = {n: [] for n in NAMES}
    gaps, ctas = [], 0
    prev_exit = None
    for i in range(1, launches):                       # launch 0 follows the graph's own start-up
        a = t[i]
        live = a[:, 0] > 0
        ctas = int(live.sum())
        a = a[live]
        t0 = a[:, 0].min()
        for k, n in enumerate(NAMES):
            col = a[:, k]
            col = col[col > 0]
            if len(col):
                per_slot[n].append(((col.min() - t0) / 1e3, float(statistics.median(col - t0)) / 1e3, (col.max() - t0) / 1e3))
        if prev_exit is not None:
            gaps.append((t0 - prev_exit) / 1e3)
        prev_exit = a[:, 11].max()
    out["ctas"] = ctas
    out["sms_used"] = int(len(set(t[launches - 1][t[launches - 1][:, 0] > 0][:, 15].tolist())))
    out["us_from_first_entry"] = {n: [round(statistics.median(v[j] for v in vals), 2) for j in range(3)]
                                  for n, vals in per_slot.items() if vals}
    out["gap_prev_exit_to_first_entry_us"] = round(statistics.median(gaps), 2) if gaps else None
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--T", type=int, default=64)
    ap.add_argument("--knob", action="append", default=[])
    ap.add_argument("--out", default=None)
    ap.add_argument("--resident", action="store_true", help="one weight copy: it stays in L2 between launches")
    ap.add_argument("--shapes", default="all", help="all | decode (the four per-layer projections)")
    a = ap.parse_args()
    for kv in a.knob:
        k, v = kv.split("=")
        kge.check(L.kge_set_tunable(k.encode(), int(v)), "tunable")
    shapes = [(6144, 4096, False), (4096, 4096, False), (14336, 4096, True), (4096, 14336, False), (128256, 4096, False)]
    fh = open(a.out, "a") if a.out else None
    if a.shapes == "decode":
        shapes = shapes[:4]
    for (N, K, gated) in shapes:
        row = trace_shape(a.T, N, K, gated, resident=a.resident)
        row["knobs"] = a.knob
        row["resident"] = a.resident
        print(json.dumps(row), flush=True)
        if fh:
            fh.write(json.dumps(row) + "\n")
            fh.flush()


if __name__ == "__main__":
    main()
