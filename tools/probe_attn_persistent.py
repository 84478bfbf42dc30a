"""Decode attention: one CTA per (request, kv head, split) against the persistent variant (knob attn_persistent), graph-timed,
interleaved.  python tools/probe_attn_persistent.py > gpurun_out/r02_attn_persistent.jsonl"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_vs_libs import time_ms, kge, L, BF  # noqa: E402


def decode(B, ctx, H, KV, hd, bs, splits):
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
    for rep in range(2):
        for pers in (0, 1):
            L.kge_set_tunable(b"attn_persistent", pers)
            ms = time_ms(run, 60)
            key = "persistent" if pers else "per_item"
            row.setdefault(key + "_us", []).append(round(ms * 1e3, 1))
            row.setdefault(key + "_GB/s", []).append(round(B * ctx * KV * hd * 4.0 / ms / 1e6))
    L.kge_set_tunable(b"attn_decode_splits", 0)
    print(json.dumps(row), flush=True)


if __name__ == "__main__":
    kge.require_cuda()
    for splits in (0, 2, 4, 8):
        decode(64, 2176, 32, 8, 128, 64, splits)
    decode(64, 1024, 32, 8, 128, 64, 0)
    decode(64, 2176, 64, 8, 128, 64, 0)      # 70B shapes: groups of 8
    decode(8, 8192, 32, 8, 128, 64, 0)
    decode(32, 192, 32, 8, 64, 64, 0)
