import os, sys, json, ctypes
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kge_loader import load
kge = load(); L = kge.lib(); BF = torch.bfloat16
raw = ctypes.CDLL(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "kubernetes-gpu-cluster_b200", "libkge.so"))
B, ctx, H, KV, hd, bs = 64, 2176, 32, 8, 128, 64
nb_seq = (ctx + bs - 1) // bs
nblk = B * nb_seq + 1
kcs = [torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF) for _ in range(3)]; vcs = [torch.randn(nblk, KV, bs, hd, device="cuda", dtype=BF) for _ in range(3)]
bt = (torch.arange(B * nb_seq, dtype=torch.int32).view(B, nb_seq) + 1).cuda()
sl = torch.full((B,), ctx, dtype=torch.int32, device="cuda"); qsl = torch.arange(B + 1, dtype=torch.int32, device="cuda")
qw = (H + 2 * KV) * hd
q = torch.randn(B, qw, device="cuda", dtype=BF); out = torch.empty(B, H * hd, device="cuda", dtype=BF)
nws = L.kge_attn_workspace_bytes(B, H, hd, ctx); ws = torch.empty(nws, dtype=torch.uint8, device="cuda")
def run(i):
    kge.check(L.kge_paged_attn(out.data_ptr(), q.data_ptr(), qw, kcs[i % 3].data_ptr(), vcs[i % 3].data_ptr(), bt.data_ptr(), nb_seq, sl.data_ptr(),
                               qsl.data_ptr(), B, B, 1, ctx, H, KV, hd, bs, hd ** -0.5, ws.data_ptr(), nws, None), "attn")
for i in range(5): run(i)
torch.cuda.synchronize()
run(5); torch.cuda.synchronize()
buf = np.zeros(512 * 8, dtype=np.uint64)
raw.kge_debug_attn_trace(ctypes.c_void_p(buf.ctypes.data))
t = buf.reshape(512, 8)[:296].astype(np.int64)
t0 = t[:, 0].min()
names = ["entry", "prefix_done", "first_tma", "first_tile", "last_tile_done", "exit"]
for k, n in enumerate(names):
    col = (t[:, k] - t0) / 1e3
    print(n, "min %.2f med %.2f max %.2f" % (col.min(), np.median(col), col.max()))
