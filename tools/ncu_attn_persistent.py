"""One launch of each decode attention variant for an ncu capture (after two warm launches each)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kge_loader import load
kge = load(); L = kge.lib(); BF = torch.bfloat16
B, ctx, H, KV, hd, bs = 64, 2176, 32, 8, 128, 64
nb_seq = (ctx + bs - 1) // bs
nblk = B * nb_seq + 1
kc = torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF); vc = torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF)
bt = (torch.arange(B * nb_seq, dtype=torch.int32).view(B, nb_seq) + 1).cuda()
sl = torch.full((B,), ctx, dtype=torch.int32, device="cuda"); qsl = torch.arange(B + 1, dtype=torch.int32, device="cuda")
qw = (H + 2 * KV) * hd
q = torch.randn(B, qw, device="cuda", dtype=BF); out = torch.empty(B, H * hd, device="cuda", dtype=BF)
nws = L.kge_attn_workspace_bytes(B, H, hd, ctx); ws = torch.empty(nws, dtype=torch.uint8, device="cuda")
def run():
    kge.check(L.kge_paged_attn(out.data_ptr(), q.data_ptr(), qw, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb_seq, sl.data_ptr(),
                               qsl.data_ptr(), B, B, 1, ctx, H, KV, hd, bs, hd ** -0.5, ws.data_ptr(), nws, None), "attn")
for pers in (0, 1):
    L.kge_set_tunable(b"attn_persistent", pers)
    for _ in range(2):
        run()
    torch.cuda.synchronize()
    torch.cuda.profiler.start(); run(); torch.cuda.synchronize(); torch.cuda.profiler.stop()
