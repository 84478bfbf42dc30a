"""Decode and prefill attention vs KV block size (TMA box = one block x 64 columns), TMA and cp.async staging, graph-timed.
  python tools/probe_attn_bs.py > gpurun_out/r02_attn_block_size.jsonl"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_vs_libs import time_ms, kge, L, BF  # noqa: E402


def decode(B, ctx, H, KV, hd, bs, splits=0):
    nb_seq = (ctx + bs - 1) // bs
    nblk = B * nb_seq + 1
    kcs = [torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF) for _ in range(3)]
    vcs = [torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF) for _ in range(3)]
    bt = (torch.arange(B * nb_seq, dtype=torch.int32).view(B, nb_seq) + 1).cuda()
    sl = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
    qsl = torch.arange(B + 1, dtype=torch.int32, device="cuda")
    qw = (H + 2 * KV) * hd
    q = torch.randn(B, qw, device="cuda", dtype=BF)
    out = torch.empty(B, H * hd, device="cuda", dtype=BF)
    nws = L.kge_attn_workspace_bytes(B, H, hd, ctx)
    ws = torch.empty(nws, dtype=torch.uint8, device="cuda")

    def run(i, st):
        kge.check(L.kge_paged_attn(out.data_ptr(), q.data_ptr(), qw, kcs[i % 3].data_ptr(), vcs[i % 3].data_ptr(), bt.data_ptr(), nb_seq,
                                   sl.data_ptr(), qsl.data_ptr(), B, B, 1, ctx, H, KV, hd, bs, hd ** -0.5, ws.data_ptr(), nws, st), "attn")
    row = {"decode": [B, ctx, H, KV, hd], "block_size": bs, "splits": splits}
    L.kge_set_tunable(b"attn_decode_splits", splits)
    for tma in (1, 0):
        L.kge_set_tunable(b"attn_tma", tma)
        ms = time_ms(run, 60)
        row["tma_us" if tma else "cp_async_us"] = round(ms * 1e3, 1)
        row["tma_GB/s" if tma else "cp_async_GB/s"] = round(B * ctx * KV * hd * 4.0 / ms / 1e6)
    L.kge_set_tunable(b"attn_tma", 1)
    L.kge_set_tunable(b"attn_decode_splits", 0)
    print(json.dumps(row), flush=True)


def prefill(B, qlen, H, KV, hd, bs):
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

    def run(i, st):
        kge.check(L.kge_paged_attn(out.data_ptr(), q.data_ptr(), qw, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb_seq, sl.data_ptr(),
                                   qsl.data_ptr(), B, B * qlen, qlen, qlen, H, KV, hd, bs, hd ** -0.5, None, 0, st), "attn")
    fl = 4.0 * H * hd * B * qlen * (qlen + 1) / 2
    row = {"prefill": [B, qlen, H, KV, hd], "block_size": bs}
    for tma in (1, 0):
        L.kge_set_tunable(b"attn_tma", tma)
        ms = time_ms(run, 10)
        row["tma_TFLOP/s" if tma else "cp_async_TFLOP/s"] = round(fl / ms / 1e9)
    L.kge_set_tunable(b"attn_tma", 1)
    print(json.dumps(row), flush=True)


if __name__ == "__main__":
    kge.require_cuda()
    for bs in (16, 32, 64, 128):
        decode(64, 2176, 32, 8, 128, bs)
    for sp in (3, 4):
        decode(64, 2176, 32, 8, 128, 16, sp)
        decode(64, 2176, 32, 8, 128, 64, sp)
    for bs in (16, 64):
        prefill(4, 2048, 32, 8, 128, bs)
