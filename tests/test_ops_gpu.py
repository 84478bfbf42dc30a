"""Per-op parity on the GPU: every kernel is called THROUGH THE C ABI (ctypes, raw device
pointers) and compared with the CPU oracle on the same seeded inputs.  Tolerances:
  - elementwise/norm ops restate vLLM's cast points exactly, so they must be bit-equal except where
    an fp32 reduction order / libm difference crosses a bf16 rounding boundary: <= 1 bf16 ulp on
    < 0.5 % of elements;
  - GEMM and attention (different fp32 summation order, bf16 P in the PV product): vLLM's IR bf16
    tolerance atol 1e-3*scale / rtol 1.6e-2 (vllm:ir/tolerances.py:12-36).
"""
import os

import numpy as np
import pytest
import torch

from oracle import model as om
from oracle import sampler as osamp

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


@pytest.fixture(scope="module")
def L(kge):
    kge.require_cuda()
    return kge.lib()


def dev(t):
    return t.cuda().contiguous()


def p(t):
    return None if t is None else t.data_ptr()


def sync():
    torch.cuda.synchronize()


def ulp_report(got, ref):
    """bf16 tensors -> (fraction differing, max ulp distance)."""
    g = got.cpu().view(torch.int16).to(torch.int32)
    r = ref.cpu().view(torch.int16).to(torch.int32)
    # map sign-magnitude to a monotone integer line
    g = torch.where(g < 0, -(g & 0x7FFF), g)
    r = torch.where(r < 0, -(r & 0x7FFF), r)
    d = (g - r).abs()
    return (d != 0).float().mean().item(), int(d.max())


def assert_near_exact(got, ref, frac=5e-3, max_ulp=1):
    f, m = ulp_report(got, ref)
    assert m <= max_ulp and f <= frac, f"differs on {f:.4%} of elements, max {m} ulp"


def check_ok(kge, rc, what):
    kge.check(rc, what)


# ------------------------------------------------------------------------------ norms


@pytest.mark.parametrize("d", [768, 2048, 4096, 8192])
@pytest.mark.parametrize("T", [1, 7, 64])
def test_rmsnorm_and_fused_add(kge, L, T, d):
    g = torch.Generator().manual_seed(T * 100003 + d)
    x = torch.randn(T, d, generator=g).to(BF)
    r = (torch.randn(T, d, generator=g) * 2).to(BF)
    w = (1 + 0.1 * torch.randn(d, generator=g)).to(BF)
    xd, rd, wd = dev(x), dev(r), dev(w)
    y = torch.empty_like(xd)
    check_ok(kge, L.kge_rmsnorm(p(y), p(xd), p(wd), T, d, 1e-5, None), "rmsnorm")
    sync()
    assert_near_exact(y, om.rms_norm(x, w, 1e-5))
    y2 = torch.empty_like(xd)
    check_ok(kge, L.kge_add_rmsnorm(p(y2), p(rd), p(xd), p(wd), T, d, 1e-5, None), "add_rmsnorm")
    sync()
    ey, eres = om.fused_add_rms_norm(x, r, w, 1e-5)
    assert torch.equal(rd.cpu(), eres)          # the residual is a plain rounded sum: exact
    assert_near_exact(y2, ey)


def test_rmsnorm_golden_from_vllm(kge, L, golden_dir):
    z = np.load(os.path.join(golden_dir, "ops_vllm.npz"))
    fb = lambda a: torch.from_numpy(a.copy()).view(BF)
    for d in (768, 2048):
        x, r, w = fb(z[f"rms_x_{d}"]), fb(z[f"rms_r_{d}"]), fb(z[f"rms_w_{d}"])
        xd, rd, wd = dev(x), dev(r), dev(w)
        y = torch.empty_like(xd)
        check_ok(kge, L.kge_add_rmsnorm(p(y), p(rd), p(xd), p(wd), x.shape[0], d, 1e-5, None), "add_rmsnorm")
        sync()
        assert torch.equal(rd.cpu(), fb(z[f"frms_res_{d}"]))
        assert_near_exact(y, fb(z[f"frms_y_{d}"]))


@pytest.mark.parametrize("d", [256, 768])
@pytest.mark.parametrize("T", [1, 33])
def test_layernorm_modes(kge, L, T, d):
    g = torch.Generator().manual_seed(T + d)
    res = torch.randn(T, d, generator=g).to(BF)
    add = torch.randn(T, d, generator=g).to(BF)
    w = (1 + 0.1 * torch.randn(d, generator=g)).to(BF)
    b = (0.1 * torch.randn(d, generator=g)).to(BF)
    rd, ad, wd, bd = dev(res), dev(add), dev(w), dev(b)
    y = torch.empty_like(rd)
    check_ok(kge, L.kge_layernorm(p(y), p(rd), None, p(wd), p(bd), T, d, 1e-5, None), "layernorm")
    sync()
    assert_near_exact(y, om.layer_norm(res, w, b))
    # also against torch's own nn.LayerNorm arithmetic (what vLLM's OPT actually calls)
    assert_near_exact(y, torch.nn.functional.layer_norm(res, (d,), w, b, 1e-5), frac=2e-2)
    check_ok(kge, L.kge_layernorm(p(y), p(rd), p(ad), p(wd), p(bd), T, d, 1e-5, None), "layernorm+add")
    sync()
    h = res + add                                  # opt.py:181 - rounded bf16 sum
    assert torch.equal(rd.cpu(), h)
    assert_near_exact(y, om.layer_norm(h, w, b))


@pytest.mark.parametrize("hd", [64, 128])
def test_head_rmsnorm_qwen3(kge, L, hd):
    g = torch.Generator().manual_seed(hd)
    T, H, KV = 9, 4, 2
    qkv = torch.randn(T, (H + 2 * KV) * hd, generator=g).to(BF)
    qw = (1 + 0.1 * torch.randn(hd, generator=g)).to(BF)
    kw = (1 + 0.1 * torch.randn(hd, generator=g)).to(BF)
    qd, qwd, kwd = dev(qkv), dev(qw), dev(kw)
    check_ok(kge, L.kge_head_rmsnorm(p(qd), p(qwd), p(kwd), T, H, KV, hd, 1e-6, None), "head_rmsnorm")
    sync()
    q, k, v = qkv.split([H * hd, KV * hd, KV * hd], dim=-1)
    eq = om.rms_norm(q.reshape(T, H, hd), qw, 1e-6).reshape(T, -1)
    ek = om.rms_norm(k.reshape(T, KV, hd), kw, 1e-6).reshape(T, -1)
    got = qd.cpu()
    assert_near_exact(got[:, :H * hd], eq)
    assert_near_exact(got[:, H * hd:(H + KV) * hd], ek)
    assert torch.equal(got[:, (H + KV) * hd:], v)


def test_silu_mul(kge, L, golden_dir):
    z = np.load(os.path.join(golden_dir, "ops_vllm.npz"))
    x = torch.from_numpy(z["silu_x"].copy()).view(BF)
    xd = dev(x)
    T, f = x.shape[0], x.shape[1] // 2
    y = torch.empty(T, f, dtype=BF, device="cuda")
    check_ok(kge, L.kge_silu_mul(p(y), p(xd), T, f, None), "silu_mul")
    sync()
    assert_near_exact(y, torch.from_numpy(z["silu_y"].copy()).view(BF))
    g = torch.Generator().manual_seed(5)
    x = (torch.randn(130, 2 * 8192, generator=g) * 3).to(BF)
    xd = dev(x)
    y = torch.empty(130, 8192, dtype=BF, device="cuda")
    check_ok(kge, L.kge_silu_mul(p(y), p(xd), 130, 8192, None), "silu_mul")
    sync()
    assert_near_exact(y, om.silu_and_mul(x))


def test_embed(kge, L):
    g = torch.Generator().manual_seed(0)
    table = torch.randn(1000, 768, generator=g).to(BF)
    ptab = torch.randn(66, 768, generator=g).to(BF)
    ids = torch.randint(0, 1000, (37,), generator=g, dtype=torch.int32)
    pos = torch.randint(0, 64, (37,), generator=g, dtype=torch.int32)
    out = torch.empty(37, 768, dtype=BF, device="cuda")
    td, pd, idd, posd = dev(table), dev(ptab), dev(ids), dev(pos)
    check_ok(kge, L.kge_embed(p(out), p(idd), p(td), None, None, 0, 37, 768, None), "embed")
    sync()
    assert torch.equal(out.cpu(), table[ids.long()])
    check_ok(kge, L.kge_embed(p(out), p(idd), p(td), p(pd), p(posd), 2, 37, 768, None), "embed+pos")
    sync()
    assert torch.equal(out.cpu(), table[ids.long()] + ptab[pos.long() + 2])
    assert L.kge_embed(p(out), p(idd), p(td), None, None, 0, 37, 770, None) == -1   # d % 8


# ------------------------------------------------------------------------------ rope + kv write


@pytest.mark.parametrize("name,H,KV", [("llama-3.2-1b", 32, 8), ("llama-3.1-8b", 32, 8), ("llama-3.1-8b", 8, 1)])
def test_rope_kvwrite(kge, L, name, H, KV):
    c = om.CONFIGS[name]
    hd, bs, nblk, T = c.head_dim, 16, 12, 29
    g = torch.Generator().manual_seed(H + hd)
    cache = om.rope_cos_sin_cache(c, 512, BF)
    qkv = torch.randn(T, (H + 2 * KV) * hd, generator=g).to(BF)
    pos = torch.randint(0, 512, (T,), generator=g, dtype=torch.int32)
    perm = torch.randperm(nblk * bs, generator=g)[:T].to(torch.int64)
    slots = perm.clone()
    slots[3] = -1                                   # padded row: rotated but not stored
    qd, cd, pd, sd = dev(qkv), dev(cache), dev(pos), dev(slots)
    kc = torch.zeros(nblk, KV, bs, hd, dtype=BF, device="cuda")
    vc = torch.zeros_like(kc)
    check_ok(kge, L.kge_rope_kvwrite(p(qd), p(cd), p(pd), p(sd), p(kc), p(vc), T, H, KV, hd, bs, None), "rope_kvwrite")
    sync()
    cos, sin = cache.index_select(0, pos.long()).chunk(2, dim=-1)
    q, k, v = qkv.split([H * hd, KV * hd, KV * hd], dim=-1)
    eq = om.apply_rope_neox(q.reshape(T, H, hd), cos, sin).reshape(T, -1)
    ek = om.apply_rope_neox(k.reshape(T, KV, hd), cos, sin).reshape(T, -1)
    got = qd.cpu()
    assert torch.equal(got[:, :H * hd], eq)
    assert torch.equal(got[:, H * hd:(H + KV) * hd], ek)
    assert torch.equal(got[:, (H + KV) * hd:], v)
    ekc = torch.zeros(nblk, KV, bs, hd, dtype=BF)
    evc = torch.zeros_like(ekc)
    for t in range(T):
        s = int(slots[t])
        if s < 0:
            continue
        ekc[s // bs, :, s % bs] = ek[t].view(KV, hd)
        evc[s // bs, :, s % bs] = v[t].view(KV, hd)
    assert torch.equal(kc.cpu(), ekc) and torch.equal(vc.cpu(), evc)


def test_rope_golden_from_vllm(kge, L, golden_dir):
    z = np.load(os.path.join(golden_dir, "ops_vllm.npz"))
    fb = lambda a: torch.from_numpy(a.copy()).view(BF)
    for name in ("llama-3.2-1b", "llama-3.1-8b"):
        c = om.CONFIGS[name]
        hd, H, KV = c.head_dim, 4, 2
        q, k = fb(z[f"rope_q_{name}"]), fb(z[f"rope_k_{name}"])
        pos = torch.from_numpy(z[f"rope_pos_{name}"]).to(torch.int32)
        T = q.shape[0]
        qkv = torch.cat([q, k, torch.zeros(T, KV * hd, dtype=BF)], dim=-1)
        n = 4096
        host = np.zeros((n, hd), dtype=np.uint16)
        rs = c.rope_scaling
        assert L.kge_rope_cache_host(host.ctypes.data, n, hd, c.rope_theta, rs["factor"], rs["low_freq_factor"],
                                     rs["high_freq_factor"], rs["original_max_position_embeddings"]) == 0
        cache = torch.from_numpy(host.view(np.int16)).view(BF)
        qd, cd, pd = dev(qkv), dev(cache), dev(pos)
        sd = dev(torch.full((T,), -1, dtype=torch.int64))
        kc = torch.zeros(1, KV, 16, hd, dtype=BF, device="cuda")
        check_ok(kge, L.kge_rope_kvwrite(p(qd), p(cd), p(pd), p(sd), p(kc), p(kc), T, H, KV, hd, 16, None), "rope")
        sync()
        got = qd.cpu()
        # the product's own host-built cache vs vLLM's: <= 1 ulp on rare elements
        assert_near_exact(got[:, :H * hd], fb(z[f"rope_qo_{name}"]), frac=2e-2, max_ulp=2)
        assert_near_exact(got[:, H * hd:(H + KV) * hd], fb(z[f"rope_ko_{name}"]), frac=2e-2, max_ulp=2)


# ------------------------------------------------------------------------------ GEMM


def gemm_ref(x, w, b, act):
    y = torch.nn.functional.linear(x.float(), w.float(), None if b is None else b.float())
    if act:
        y = torch.relu(y)
    return y


GEMM_CASES = [
    # T, N, K, bias, act                      what it exercises
    (1, 768, 768, True, 0),                   # single token, split-K, bias
    (5, 2304, 768, True, 0),                  # opt qkv
    (16, 3072, 768, True, 1),                 # fc1 + relu
    (64, 6144, 4096, False, 0),               # 8B qkv decode: split-K 3
    (64, 4096, 14336, False, 0),              # 8B down decode: deep split-K
    (33, 4096, 4096, False, 0),               # ragged token tile
    (64, 28672, 4096, False, 0),              # 8B gate_up decode: no split
    (7, 50272, 768, False, 0),                # opt lm head: N % 128 != 0
    (130, 1024, 512, True, 0),                # 256-token tile, OOB token rows
    (300, 2048, 2048, False, 0),              # two token tiles
    (1000, 4096, 4096, True, 1),              # prefill-sized
    (2048, 1024, 8192, False, 0),             # long K
    (16, 128, 64, False, 0),                  # smallest: one k-block
    (20, 200, 72, True, 0),                   # K % 64 != 0 (TMA zero fill), N < 128
    (64, 2560, 8192, False, 0),               # 70B TP4 qkv decode: 20 tiles, cluster of 6 or 7
    (64, 3584, 8192, True, 0),                # 28 tiles: cluster of 5
    (64, 1024, 8192, False, 0),               # 8 tiles: cluster of 8
    (64, 8192, 2048, False, 0),               # 70B TP4 o_proj decode: 64 tiles, cluster of 2
    (8192, 8192, 2048, False, 0),             # 70B TP4 o_proj prefill: CTA pairs, short K
    (8192, 2560, 8192, False, 0),             # 70B TP4 qkv prefill: 20 row tiles -> 10 pairs
    (64, 1000, 2048, True, 1),                # N % 8 != 0: scalar stores in the split reduce, ragged last tile
    (48, 4104, 4096, False, 0),               # N a multiple of 8 but of no tile height
    (100, 6144, 4096, True, 0),               # 128-token tile, split K
    (64, 8192, 8192, True, 1),                # 32 wide tiles x 4 splits, bias + relu in the split reduce
    (3, 37888, 1024, False, 0),               # 148 wide tiles exactly, no split
    (50, 5120, 640, True, 0),                 # 20 wide tiles, K = 10 k-blocks: 2 splits of 5
]


@pytest.mark.parametrize("T,N,K,bias,act", GEMM_CASES)
def test_gemm_bf16(kge, L, T, N, K, bias, act):
    g = torch.Generator().manual_seed(T * 7 + N * 3 + K)
    x = torch.randn(T, K, generator=g).to(BF)
    w = (torch.randn(N, K, generator=g) * 0.05).to(BF)
    b = torch.randn(N, generator=g).to(BF) if bias else None
    xd, wd, bd = dev(x), dev(w), (dev(b) if bias else None)
    y = torch.full((T, N), float("nan"), dtype=BF, device="cuda")
    nws = L.kge_gemm_workspace_bytes(T, N, K)
    ws = torch.zeros(nws, dtype=torch.uint8, device="cuda")
    for _ in range(2):       # twice: back-to-back launches must not interfere
        y.fill_(float("nan"))
        check_ok(kge, L.kge_gemm_bf16(p(y), p(xd), p(wd), p(bd), T, N, K, act, p(ws), nws, None), "gemm")
        sync()
        ref = gemm_ref(x, w, b, act)
        got = y.cpu().float()
        assert torch.isfinite(got).all()
        scale = ref.abs().max().item()
        err = (got - ref).abs()
        tol = 1e-3 * scale + 1.6e-2 * ref.abs()
        assert (err <= tol).all(), f"max err {err.max().item():.4g} (scale {scale:.3g}) at {torch.nonzero(err > tol)[:4].tolist()}"
        # and tight on average: one bf16 rounding of an fp32 accumulation
        assert (err / (ref.abs() + 1e-2 * scale)).mean().item() < 3e-3


@pytest.mark.parametrize("T,N,K,bias,act", [c for c in GEMM_CASES if c[0] <= 64])
def test_gemm_128_row_tiles_for_small_batches(kge, L, T, N, K, bias, act):
    """Knob gemm_wide=0: batches of up to 64 tokens through the 128-row weight tiles (tokens as MMA N) instead of the
    256-row tiles with the tokens as MMA M - the same checks."""
    try:
        check_ok(kge, L.kge_set_tunable(b"gemm_wide", 0), "tunable")
        test_gemm_bf16(kge, L, T, N, K, bias, act)
    finally:
        L.kge_set_tunable(b"gemm_wide", 1)


@pytest.mark.parametrize("T,f,K", [(1, 512, 256), (7, 1024, 512), (64, 14336, 4096), (64, 1792, 4096), (32, 8192, 2048), (64, 7168, 8192)])
def test_gemm_silu_gate_128_row_tiles(kge, L, T, f, K):
    try:
        check_ok(kge, L.kge_set_tunable(b"gemm_wide", 0), "tunable")
        test_gemm_silu_gate_fused(kge, L, T, f, K)
    finally:
        L.kge_set_tunable(b"gemm_wide", 1)


@pytest.mark.parametrize("T,N,K,bias,act", [c for c in GEMM_CASES if c[0] <= 128] + [(64, 6144, 4096, True, 1), (33, 14336, 4096, False, 0), (100, 4096, 14336, False, 0)])
def test_gemm_stream_k_through_a_workspace(kge, L, T, N, K, bias, act):
    """Knob gemm_ws: decode-sized shapes as stream-K - one CTA per SM, each a contiguous share of all (tile, k-block) units,
    partial tiles reduced through the caller's workspace by the last CTA to arrive - against fp32 torch and against the
    default (cluster split-K / persistent tiles).  Partials are summed in k order, so the result is the same bits on every
    call; it differs from the default only by fp32 summation order.  A workspace that is too small must fall back, not fail."""
    g = torch.Generator().manual_seed(T + N + K)
    xh = torch.randn(T, K, generator=g).to(BF)
    wh = (torch.randn(N, K, generator=g) * 0.05).to(BF)
    bh = torch.randn(N, generator=g).to(BF) if bias else None
    x, w, b = dev(xh), dev(wh), (dev(bh) if bias else None)
    outs = []
    try:
        check_ok(kge, L.kge_set_tunable(b"gemm_ws", 1), "tunable")
        nws = L.kge_gemm_workspace_bytes(T, N, K)
        ws = torch.zeros(max(nws, 16), dtype=torch.uint8, device="cuda")
        for (wsp, nb) in ((p(ws), nws), (p(ws), nws), (None, 0), (p(ws), 2048 if nws > 2048 else 0)):
            y = torch.full((T, N), float("nan"), dtype=BF, device="cuda")
            check_ok(kge, L.kge_gemm_bf16(p(y), p(x), p(w), p(b) if bias else None, T, N, K, act, wsp, nb, None), "gemm")
            sync()
            outs.append(y.cpu())
    finally:
        L.kge_set_tunable(b"gemm_ws", 0)
    assert torch.equal(outs[0].view(torch.int16), outs[1].view(torch.int16))          # workspace path: same bits every time
    assert torch.equal(outs[2].view(torch.int16), outs[3].view(torch.int16))          # too-small scratch == no scratch
    assert int(ws[:1024].max().item()) == 0                                            # the counters were left at zero
    ref = gemm_ref(xh, wh, bh, act)
    for o in (outs[0], outs[2]):
        err = (o.float() - ref).abs()
        assert (err <= 1e-3 * ref.abs().max().item() + 1.6e-2 * ref.abs()).all(), err.max().item()


@pytest.mark.parametrize("T,f,K", [(64, 14336, 4096), (16, 3584, 8192), (64, 1024, 512), (128, 7168, 2048)])
def test_gemm_silu_gate_stream_k(kge, L, T, f, K):
    """The fused SiLU-gate projection through the stream-K path (knob gemm_ws): partial tiles carry gate AND up rows, the
    reducing CTA applies silu(gate) * up after the sum - same cast points, so the same tolerance as the direct epilogue."""
    g = torch.Generator().manual_seed(T + f + K)
    x = dev(torch.randn(T, K, generator=g).to(BF))
    w = dev((torch.randn(2 * f, K, generator=g) * 0.05).to(BF))
    outs = []
    try:
        check_ok(kge, L.kge_set_tunable(b"gemm_ws", 1), "tunable")
        nws = L.kge_gemm_workspace_bytes(T, f, K)
        assert nws > 1024
        ws = torch.zeros(nws, dtype=torch.uint8, device="cuda")
        for (wsp, nb) in ((p(ws), nws), (p(ws), nws), (None, 0)):
            y = torch.full((T, f), float("nan"), dtype=BF, device="cuda")
            check_ok(kge, L.kge_gemm_silu_gate_ws_bf16(p(y), p(x), p(w), T, f, K, wsp, nb, None), "gemm_silu_gate_ws")
            sync()
            outs.append(y.cpu())
    finally:
        L.kge_set_tunable(b"gemm_ws", 0)
    assert torch.equal(outs[0].view(torch.int16), outs[1].view(torch.int16))
    gu = (x.cpu().float() @ w.cpu().float().t()).to(BF)
    ref = om.silu_and_mul(gu).float()
    for o in outs:
        err = (o.float() - ref).abs()
        assert (err <= 4e-3 * ref.abs().max().item() + 3.2e-2 * ref.abs()).all(), err.max().item()


@pytest.mark.parametrize("T,N,K,bias,act", [(64, 4096, 4096, False, 0), (1000, 4096, 4096, True, 1), (8192, 2560, 8192, False, 0), (300, 2048, 2048, False, 0)])
def test_gemm_with_l2_eviction_hints(kge, L, T, N, K, bias, act):
    """Knob gemm_l2_hints only changes which L2 lines are kept (weights evict_first, activations evict_last): same bits."""
    g = torch.Generator().manual_seed(T + N)
    x = dev(torch.randn(T, K, generator=g).to(BF))
    w = dev((torch.randn(N, K, generator=g) * 0.05).to(BF))
    b = dev(torch.randn(N, generator=g).to(BF)) if bias else None
    outs = []
    try:
        for knob in (0, 1):
            check_ok(kge, L.kge_set_tunable(b"gemm_l2_hints", knob), "tunable")
            y = torch.full((T, N), float("nan"), dtype=BF, device="cuda")
            check_ok(kge, L.kge_gemm_bf16(p(y), p(x), p(w), p(b) if bias else None, T, N, K, act, None, 0, None), "gemm")
            sync()
            outs.append(y.cpu())
    finally:
        L.kge_set_tunable(b"gemm_l2_hints", 0)
    assert torch.isfinite(outs[0].float()).all()
    assert torch.equal(outs[0].view(torch.int16), outs[1].view(torch.int16))


@pytest.mark.parametrize("T,f,K", [(1, 512, 256), (7, 1024, 512), (64, 14336, 4096), (64, 1792, 4096), (32, 8192, 2048),
                                   (130, 512, 1024), (300, 3584, 8192), (1000, 14336, 1024),
                                   (64, 7168, 8192), (8192, 7168, 1024)])    # 70B TP4 gate_up: decode, prefill pairs
def test_gemm_silu_gate_fused(kge, L, T, f, K):
    """Fused gate_up GEMM + SiLU*mul against the oracle's two ops (F.linear then silu_and_mul), incl.
    cluster split-K shapes (f/64 tiles <= 74) and 256-token tiles."""
    g = torch.Generator().manual_seed(T + f + K)
    x = torch.randn(T, K, generator=g).to(BF)
    w = (torch.randn(2 * f, K, generator=g) * (2.0 / K ** 0.5)).to(BF)
    xd, wd = dev(x), dev(w)
    y = torch.full((T, f), float("nan"), dtype=BF, device="cuda")
    for _ in range(2):
        y.fill_(float("nan"))
        check_ok(kge, L.kge_gemm_silu_gate_bf16(p(y), p(xd), p(wd), T, f, K, None), "gemm_silu_gate")
        sync()
        gu = torch.nn.functional.linear(x.float(), w.float()).to(BF)       # bf16 GEMM output, fp32 accumulate
        ref = om.silu_and_mul(gu).float()
        got = y.cpu().float()
        assert torch.isfinite(got).all()
        scale = ref.abs().max().item()
        err = (got - ref).abs()
        tol = 2e-3 * scale + 2.5e-2 * ref.abs()      # two bf16-rounded GEMM outputs feed a product
        assert (err <= tol).all(), f"max err {err.max().item():.4g} (scale {scale:.3g})"
        assert (err / (ref.abs() + 1e-2 * scale)).mean().item() < 4e-3
    assert L.kge_gemm_silu_gate_bf16(p(y), p(xd), p(wd), T, f + 8, K, None) == -1      # f % 64


def test_gemm_rejects_bad_arguments(kge, L):
    x = torch.zeros(4, 64, dtype=BF, device="cuda")
    assert L.kge_gemm_bf16(p(x), p(x), p(x), None, 4, 4, 60, 0, None, 0, None) == -1       # K % 8
    assert L.kge_gemm_bf16(p(x), None, p(x), None, 4, 4, 64, 0, None, 0, None) == -1       # null x
    assert L.kge_gemm_bf16(p(x), p(x), p(x), None, 4, 4, 64, 7, None, 0, None) == -1       # unknown act
    assert b"act" in L.kge_last_error()
    assert L.kge_gemm_workspace_bytes(8192, 4096, 4096) == 0    # prefill-sized: CTA pairs, no scratch
    try:
        L.kge_set_tunable(b"gemm_ws", 1)
        assert L.kge_gemm_workspace_bytes(1, 768, 4096) > 1024      # decode-sized: stream-K through the workspace
        assert L.kge_gemm_workspace_bytes(64, 128256, 4096) == 0    # 1002 tiles: persistent tiles already balance
    finally:
        L.kge_set_tunable(b"gemm_ws", 0)


# ------------------------------------------------------------------------------ attention


def oracle_attention_chunked(q, k, v, scale, q_pos0, chunk=256):
    """om.causal_attention over query chunks (same function, same values): keeps the fp32 score matrix of a
    2048- or 8192-token prefill small enough for the host."""
    outs = []
    for c0 in range(0, q.shape[0], chunk):
        c1 = min(q.shape[0], c0 + chunk)
        outs.append(om.causal_attention(q[c0:c1], k[:q_pos0 + c1], v[:q_pos0 + c1], scale, q_pos0 + c0))
    return torch.cat(outs, 0)


def run_attn_case(kge, L, H, KV, hd, bs, q_lens, seq_lens, seed, ramp=False):
    g = torch.Generator().manual_seed(seed)
    B = len(q_lens)
    T = sum(q_lens)
    max_blocks = max((s + bs - 1) // bs for s in seq_lens)
    nblk = sum((s + bs - 1) // bs for s in seq_lens) + 3
    free = torch.randperm(nblk - 1, generator=g).add(1).tolist()      # block 0 = null
    kc = (torch.randn(nblk, KV, bs, hd, generator=g)).to(BF)
    vc = (torch.randn(nblk, KV, bs, hd, generator=g)).to(BF)
    bt = torch.zeros(B, max_blocks, dtype=torch.int32)
    for b, s in enumerate(seq_lens):
        for j in range((s + bs - 1) // bs):
            bt[b, j] = free.pop()
    qkv_w = (H + 2 * KV) * hd
    qkv = torch.randn(T, qkv_w, generator=g).to(BF)
    if ramp:
        # growing row maximum: every query has a positive component along a fixed direction u and key j carries
        # (j / ctx) * 40 * u, so the scaled score rises by ~85 (in log2 units) across the context and the running
        # maximum of every row grows tile after tile - the case a lazily rescaled accumulator has to get right
        u = torch.nn.functional.normalize(torch.randn(hd, generator=g), dim=0)
        qkv[:, :H * hd] = (qkv[:, :H * hd].float().view(T, H, hd) + 6.0 * hd ** 0.5 * 0.25 * u).reshape(T, H * hd).to(BF)
        for b, s_len in enumerate(seq_lens):
            for j in range((s_len + bs - 1) // bs):
                pos = (torch.arange(bs) + j * bs).float() / max(1, s_len)
                kc[bt[b, j]] = (kc[bt[b, j]].float() + (40.0 * pos).view(1, bs, 1) * u.view(1, 1, hd)).to(BF)
    qsl = [0]
    for ql in q_lens:
        qsl.append(qsl[-1] + ql)
    out = torch.full((T, H * hd), float("nan"), dtype=BF, device="cuda")
    nws = L.kge_attn_workspace_bytes(B, H, hd, max(seq_lens))
    ws = torch.empty(max(nws, 16), dtype=torch.uint8, device="cuda")
    qd, kd, vd, btd = dev(qkv), dev(kc), dev(vc), dev(bt)
    sld, qsld = dev(torch.tensor(seq_lens, dtype=torch.int32)), dev(torch.tensor(qsl, dtype=torch.int32))
    scale = hd ** -0.5
    check_ok(kge, L.kge_paged_attn(p(out), p(qd), qkv_w, p(kd), p(vd), p(btd), max_blocks, p(sld), p(qsld), B, T,
                                   max(q_lens), max(seq_lens), H, KV, hd, bs, scale, p(ws), nws, None), "paged_attn")
    sync()
    got = out.cpu().float()
    assert torch.isfinite(got).all()
    for b in range(B):
        s, ql = seq_lens[b], q_lens[b]
        blocks = bt[b, :(s + bs - 1) // bs].long()
        k = kc[blocks].permute(0, 2, 1, 3).reshape(-1, KV, hd)[:s]
        v = vc[blocks].permute(0, 2, 1, 3).reshape(-1, KV, hd)[:s]
        q = qkv[qsl[b]:qsl[b + 1], :H * hd].reshape(ql, H, hd)
        ref = oracle_attention_chunked(q, k, v, scale, s - ql).float()
        err = (got[qsl[b]:qsl[b + 1]] - ref).abs()
        tol = 1e-3 * ref.abs().max() + 1.6e-2 * ref.abs() + 4e-3
        assert (err <= tol).all(), f"request {b}: max err {err.max().item():.4g}"
    return out.cpu()


ATTN_CASES = [
    # H, KV, hd, bs, q_lens, seq_lens
    (32, 8, 128, 16, [1] * 5, [1, 17, 64, 65, 300]),                    # decode, ragged, no split
    (32, 8, 128, 16, [1] * 3, [2176, 1500, 700]),                       # decode, split-KV + merge
    (32, 8, 64, 16, [1] * 4, [40, 128, 129, 1000]),                     # 1B shape decode
    (12, 12, 64, 16, [1] * 3, [5, 130, 257]),                           # OPT MHA decode (G = 1)
    (64, 8, 128, 32, [1, 1], [900, 33]),                                # 70B group of 8, block 32
    (32, 8, 128, 16, [37], [37]),                                       # pure prefill
    (32, 8, 128, 16, [200, 64], [200, 64]),                             # two prefills, multi q tiles
    (32, 8, 128, 16, [50], [300]),                                      # chunk continuing a context
    (32, 8, 64, 16, [1, 30, 1, 7], [77, 30, 256, 135]),                 # mixed decode + prefill
    (12, 12, 64, 16, [70, 1], [70, 200]),                               # OPT mixed
    (4, 1, 128, 16, [33, 1], [40, 90]),                                 # tiny-llama-hd128 shape
    (28, 4, 128, 16, [20, 1], [20, 300]),                               # group of 7 (Qwen2.5-7B shape)
]


@pytest.mark.parametrize("H,KV,hd,bs,q_lens,seq_lens", ATTN_CASES)
def test_paged_attention(kge, L, H, KV, hd, bs, q_lens, seq_lens):
    run_attn_case(kge, L, H, KV, hd, bs, q_lens, seq_lens, seed=H * 31 + sum(seq_lens))


# The shapes the C3 bench runs the tcgen05 prefill kernel at (32 and 128 key tiles per row, where its K/V pipeline and
# its lazy accumulator rescale actually cycle), the mixed step vLLM's scheduling policy produces, and both head sizes.
ATTN_LONG_CASES = [
    # H, KV, hd, q_lens, seq_lens, growing row maximum
    (32, 8, 128, [2048] * 4, [2048] * 4, False),                         # a C3 prefill step: four whole prompts
    (32, 8, 128, [8192], [8192], False),                                 # one 8192-token prompt
    (32, 8, 128, [1] * 7 + [2037, 4], [2100 + 7 * i for i in range(7)] + [2037, 2048], False),   # 7 decodes + head + tail chunk
    (32, 8, 64, [2048, 2048], [2048, 2048], False),                      # 1B head size
    (32, 8, 64, [1] * 3 + [1000, 48], [300, 2000, 77, 1000, 2048], False),
    (32, 8, 128, [2048], [2048], True),                                  # row maximum grows through every key tile
    (32, 8, 64, [1500, 1], [1500, 1800], True),
    (8, 1, 128, [512], [4096], True),                                    # a late chunk of a long context, group of 8
]


@pytest.mark.parametrize("H,KV,hd,q_lens,seq_lens,ramp", ATTN_LONG_CASES)
def test_paged_attention_bench_shapes(kge, L, H, KV, hd, q_lens, seq_lens, ramp):
    run_attn_case(kge, L, H, KV, hd, 16, q_lens, seq_lens, seed=H * 17 + sum(q_lens) + int(ramp), ramp=ramp)


PERSISTENT_DEFAULT = 1          # the library's default for knob attn_persistent

DECODE_SPLIT_CASES = [
    # H, KV, hd, q_lens, seq_lens, forced split-KV factor (0 = heuristic)
    (32, 8, 128, [1] * 3, [2176, 1500, 700], 0),
    (32, 8, 128, [1] * 3, [2176, 1500, 700], 3),
    (32, 8, 128, [1] * 2, [4100, 257], 8),
    (32, 8, 64, [1] * 4, [40, 128, 129, 1000], 2),
    (12, 12, 64, [1] * 3, [5, 130, 600], 2),
    (64, 8, 128, [1, 1], [900, 33], 3),
    (28, 4, 128, [1] * 2, [300, 2000], 5),
    (32, 8, 128, [1], [6000], 0),                # heuristic picks > 8 splits: workspace + merge kernel path
]


@pytest.mark.parametrize("H,KV,hd,q_lens,seq_lens,splits", DECODE_SPLIT_CASES)
def test_decode_attention_variants_give_the_same_bits(kge, L, H, KV, hd, q_lens, seq_lens, splits):
    """Three ways to stage K/V and merge the split-KV partials - TMA boxes + merge by the last CTA (default), per-thread
    cp.async + merge kernel (knob attn_tma 0), cp.async + cluster/DSMEM merge (attn_cluster_merge 1) - run the same
    arithmetic in the same order: they must give the same BITS, and each must match the oracle."""
    seed = H * 7 + sum(seq_lens) + splits
    try:
        check_ok(kge, L.kge_set_tunable(b"attn_decode_splits", splits), "tunable")
        a = run_attn_case(kge, L, H, KV, hd, 16, q_lens, seq_lens, seed)
        a2 = run_attn_case(kge, L, H, KV, hd, 16, q_lens, seq_lens, seed)        # arrival counters were reset by the first launch
        check_ok(kge, L.kge_set_tunable(b"attn_tma", 0), "tunable")
        b = run_attn_case(kge, L, H, KV, hd, 16, q_lens, seq_lens, seed)
        check_ok(kge, L.kge_set_tunable(b"attn_cluster_merge", 1), "tunable")
        c = run_attn_case(kge, L, H, KV, hd, 16, q_lens, seq_lens, seed)
    finally:
        L.kge_set_tunable(b"attn_decode_splits", 0)
        L.kge_set_tunable(b"attn_cluster_merge", 0)
        L.kge_set_tunable(b"attn_tma", 1)
    for other in (a2, b, c):
        assert torch.equal(a.view(torch.int16), other.view(torch.int16))


PERSISTENT_CASES = [
    # H, KV, hd, block size, seq_lens, forced split factor
    (32, 8, 128, 64, [2100 + 3 * i for i in range(48)], 0),        # 48 x 8 x 2 items > 296 resident CTAs: several items per CTA
    (32, 8, 128, 64, [1 + 97 * i for i in range(40)], 3),          # ragged: empty splits, one-tile items
    (32, 8, 128, 32, [700, 64, 65, 2176, 5], 2),
    (64, 8, 128, 64, [900 + 11 * i for i in range(20)], 2),        # groups of 8 q heads: two merge passes
    (32, 8, 64, 64, [40, 128, 129, 1000] * 12, 2),
    (12, 12, 64, 128, [5, 130, 600] * 30, 1),                      # no split: results straight from the cross-warp merge
    (16, 1, 128, 64, [3000, 17], 4),                               # groups of 16 q heads: four merge passes
    (32, 8, 128, 64, [130 + 7 * i for i in range(180)], 0),        # 180 requests: the per-request prefix is at its largest
    (32, 8, 64, 64, [1 + i for i in range(192)], 0),
]


@pytest.mark.parametrize("H,KV,hd,bs,seq_lens,splits", PERSISTENT_CASES)
def test_decode_attention_persistent(kge, L, H, KV, hd, bs, seq_lens, splits):
    """Knob attn_persistent: 2 x 148 CTAs each take an equal, contiguous share of all key tiles (a group is cut where the
    shares end, not into equal splits).  Same tile arithmetic, other split points: against the oracle (run_attn_case),
    within rounding of the one-CTA-per-split kernel, and the same bits on every run (partials merged in split order)."""
    seed = H * 5 + sum(seq_lens) + splits
    q_lens = [1] * len(seq_lens)
    try:
        check_ok(kge, L.kge_set_tunable(b"attn_decode_splits", splits), "tunable")
        check_ok(kge, L.kge_set_tunable(b"attn_persistent", 0), "tunable")
        a = run_attn_case(kge, L, H, KV, hd, bs, q_lens, seq_lens, seed)
        check_ok(kge, L.kge_set_tunable(b"attn_persistent", 1), "tunable")
        b = run_attn_case(kge, L, H, KV, hd, bs, q_lens, seq_lens, seed)
        c = run_attn_case(kge, L, H, KV, hd, bs, q_lens, seq_lens, seed)
    finally:
        L.kge_set_tunable(b"attn_decode_splits", 0)
        L.kge_set_tunable(b"attn_persistent", PERSISTENT_DEFAULT)
    assert torch.equal(b.view(torch.int16), c.view(torch.int16))
    assert (a.float() - b.float()).abs().max().item() <= 2e-2 * max(1.0, a.float().abs().max().item())


@pytest.mark.parametrize("H,KV,hd,bs,q_lens,seq_lens", [c for c in ATTN_CASES if max(c[4]) > 1][:5] + [ATTN_LONG_CASES[2][:3] + (16,) + ATTN_LONG_CASES[2][3:5]])
def test_prefill_attention_tma_and_cp_async_agree(kge, L, H, KV, hd, bs, q_lens, seq_lens):
    """The tcgen05 prefill kernel with TMA-staged K/V (default) and with the round-1 cp.async gathers computes the same
    tiles in the same order: same bits."""
    seed = H * 13 + sum(seq_lens)
    try:
        a = run_attn_case(kge, L, H, KV, hd, bs, q_lens, seq_lens, seed)
        check_ok(kge, L.kge_set_tunable(b"attn_tma", 0), "tunable")
        b = run_attn_case(kge, L, H, KV, hd, bs, q_lens, seq_lens, seed)
    finally:
        L.kge_set_tunable(b"attn_tma", 1)
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))


def test_attention_block_sizes(kge, L):
    """Block sizes from 8 (one TMA box per block, eight boxes per tile) to 128 (two tiles per block) and 4 (below the TMA
    box granularity: cp.async path), decode and prefill."""
    for bs in (4, 8, 32, 64, 128):
        run_attn_case(kge, L, 8, 2, 128, bs, [1, 1, 1], [700, 129, 64], seed=bs)
        run_attn_case(kge, L, 8, 2, 64, bs, [150, 1, 40], [150, 300, 200], seed=bs + 1)


# ------------------------------------------------------------------------------ sampling


def test_argmax_ties_and_padding(kge, L):
    g = torch.Generator().manual_seed(3)
    B, V, ld = 9, 50272, 50272
    lg = torch.randn(B, ld, generator=g).to(BF)
    lg[0, 100] = lg[0, 40000] = 9.0           # tie: lowest index wins
    lg[1, V - 1] = 50.0                       # last column
    lg[2, 0] = 60.0
    ld_ = dev(lg)
    out = torch.empty(B, dtype=torch.int32, device="cuda")
    check_ok(kge, L.kge_argmax(p(out), p(ld_), B, V, ld, None), "argmax")
    big = torch.randn(3, 128256, generator=g).to(BF)
    bigd = dev(big)
    ob = torch.empty(3, dtype=torch.int32, device="cuda")
    check_ok(kge, L.kge_argmax(p(ob), p(bigd), 3, 128256, 128256, None), "argmax")
    sync()
    assert ob.cpu().tolist() == osamp.greedy(big).tolist()
    sync()
    assert out.cpu().tolist() == osamp.greedy(lg).tolist()
    assert out[0].item() == 100
    # padded leading dimension: columns >= V are ignored
    V2, ld2 = 1000, 1024
    lg2 = torch.randn(4, ld2, generator=g).to(BF)
    lg2[:, V2:] = 100.0
    out2 = torch.empty(4, dtype=torch.int32, device="cuda")
    lg2d = dev(lg2)
    check_ok(kge, L.kge_argmax(p(out2), p(lg2d), 4, V2, ld2, None), "argmax")
    sync()
    assert out2.cpu().tolist() == lg2[:, :V2].float().argmax(-1).tolist()


def test_sample_topk_topp_with_given_noise(kge, L):
    """Same draw as vLLM's sampler when the Exp(1) noise is supplied."""
    g = torch.Generator().manual_seed(11)
    B, V = 12, 4096
    lg = (torch.randn(B, V, generator=g) * 2.5).to(BF)
    temp = torch.tensor([0.0, 1.0, 0.7, 1.3, 1.0, 1.0, 0.5, 1.0, 2.0, 1.0, 1.0, 0.0])
    k = torch.tensor([0, 0, 50, 0, 1, 5, 0, V, 20, 3, 0, 7], dtype=torch.int32)
    pp = torch.tensor([1.0, 1.0, 1.0, 0.9, 1.0, 0.5, 0.3, 1.0, 0.95, 0.99, 0.6, 0.2])
    agree = 0
    total = 0
    lgd, td, kd, ppd = dev(lg), dev(temp), dev(k), dev(pp)     # keep the device tensors alive
    for trial in range(8):
        q = torch.empty(B, V).exponential_(generator=g)
        qd = dev(q)
        out = torch.empty(B, dtype=torch.int32, device="cuda")
        check_ok(kge, L.kge_sample(p(out), p(lgd), B, V, V, p(td), p(kd), p(ppd), p(qd),
                                   None, 0, None, 0, None), "sample")
        sync()
        kk = torch.where(k <= 0, torch.full_like(k, V), k).long()
        ref = osamp.sample(lg, temp, kk, pp, q)
        got = out.cpu().long()
        agree += (got == ref).sum().item()
        total += B
        # hard guarantees whatever the rounding at the top-p boundary
        masked = osamp.apply_top_k_top_p(lg.float() / torch.where(temp < 1e-5, torch.ones_like(temp), temp)[:, None], kk, pp)
        for b in range(B):
            if temp[b] < 1e-5:
                assert got[b] == lg[b].float().argmax()
            elif k[b] == 1:
                assert got[b] == lg[b].float().argmax()
    # fp32 cumsum order differs from a sort-based cumsum: allow a boundary token to flip rarely
    assert agree >= total - 2, f"{agree}/{total}"


def test_sample_philox_is_deterministic_and_in_support(kge, L):
    g = torch.Generator().manual_seed(2)
    B, V = 4, 2048
    lg = (torch.randn(B, V, generator=g) * 3).to(BF)
    temp = torch.ones(B)
    k = torch.full((B,), 10, dtype=torch.int32)
    pp = torch.ones(B)
    seeds = torch.tensor([1, 2, 3, 4], dtype=torch.int64)
    outs = []
    lgd, td, kd, ppd, sd = dev(lg), dev(temp), dev(k), dev(pp), dev(seeds)
    for step in (0, 0, 1):
        out = torch.empty(B, dtype=torch.int32, device="cuda")
        check_ok(kge, L.kge_sample(p(out), p(lgd), B, V, V, p(td), p(kd), p(ppd), None,
                                   p(sd), step, None, 0, None), "sample")
        sync()
        outs.append(out.cpu().tolist())
    assert outs[0] == outs[1]
    # kept set = every logit >= the 10th largest VALUE (ties survive, as in vLLM's `logits < kth` mask)
    kth = lg.float().topk(10, dim=-1).values[:, -1]
    for b in range(B):
        assert lg[b, outs[0][b]].float() >= kth[b] and lg[b, outs[2][b]].float() >= kth[b]


# ------------------------------------------------------------------------------ logprobs
@pytest.mark.parametrize("R,V,ld,top_n", [(3, 1000, 1000, 5), (5, 50272, 50304, 20), (2, 128256, 128256, 0), (4, 777, 784, 3), (1, 12, 16, 20)])
def test_logprobs_match_torch(kge, L, R, V, ld, top_n):
    """kge_logprobs against vLLM's recipe (sampler.py:300-349): log_softmax of fp32(logits), the sampled token's value,
    the top-N (ties by ascending id, which torch.topk leaves open) and the rank = #{x >= x_sampled}."""
    g = torch.Generator().manual_seed(R * 1000 + V)
    logits = (torch.randn(R, ld, generator=g) * 3).to(BF)           # bf16: plenty of exact ties
    logits[:, V:] = 100.0                                            # padding columns must be ignored
    sampled = torch.randint(0, V, (R,), generator=g, dtype=torch.int32)
    sampled[0] = int(logits[0, :V].float().argmax())
    ld_, sd = dev(logits), dev(sampled)
    vals = torch.zeros(R, top_n + 1, dtype=torch.float32, device="cuda")
    ids = torch.zeros(R, top_n + 1, dtype=torch.int32, device="cuda")
    rank = torch.zeros(R, dtype=torch.int32, device="cuda")
    check_ok(kge, L.kge_logprobs(p(vals), p(ids), p(rank), p(ld_), p(sd), R, V, ld, top_n, None), "logprobs")
    sync()
    x = logits[:, :V].float()
    lp = torch.log_softmax(x, -1)
    vals, ids, rank = vals.cpu(), ids.cpu(), rank.cpu()
    for r in range(R):
        t = int(sampled[r])
        assert ids[r, 0] == t and abs(vals[r, 0].item() - lp[r, t].item()) <= 2e-5 * max(1.0, abs(lp[r, t].item()))
        assert rank[r] == int((x[r] >= x[r, t]).sum())
        order = sorted(range(V), key=lambda j: (-x[r, j].item(), j))[:top_n] if V <= 2000 else None
        n_have = min(top_n, V)
        tv, _ = torch.topk(lp[r], n_have)
        assert torch.allclose(vals[r, 1:1 + n_have], tv, rtol=2e-5, atol=2e-5)
        got = ids[r, 1:1 + n_have].tolist()
        assert len(set(got)) == n_have and all(0 <= j < V for j in got)
        assert torch.equal(x[r, got], torch.topk(x[r], n_have).values)          # the picked ids carry the top values
        if order is not None:
            assert got == order[:n_have]
        if n_have < top_n:                                                        # fewer than N entries exist
            assert (ids[r, 1 + n_have:] == -1).all() and torch.isinf(vals[r, 1 + n_have:]).all()
    assert rank[0] == int((x[0] >= x[0].max()).sum())
