"""kge_paged_attn_decode_noalloc: the decode-attention entry point for hosts that cannot allocate.

Some users of libkge.so cannot give kge_paged_attn a workspace: their engine took all the device
memory it was allowed for the KV pool at start-up, or they call from a serving thread where the
device-wide stall of cudaMalloc is not acceptable.  For them the split-KV scratch is an 8 MiB array
built into the library.  This test drives that entry point through the C ABI and checks that

  * it returns what kge_paged_attn (caller-provided workspace) returns for the same inputs, on two
    calls in a row - the second call finds the first call's partials still in the shared array and
    must not be disturbed by them - and again for a different batch that reuses the same array;
  * once set up, a call leaves the amount of free device memory untouched;
  * a batch whose partials would not fit the built-in array is refused with a message that says so.

The file name sorts last on purpose, so a run that stops at the first failure has already run
everything else.  It has no dependency on the other tests and passes in a fresh process.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


def make_case(B, H, KV, hd, bs, seq_lens, seed):
    g = torch.Generator().manual_seed(seed)
    max_blocks = max((s + bs - 1) // bs for s in seq_lens)
    nblk = sum((s + bs - 1) // bs for s in seq_lens) + 2
    order = torch.randperm(nblk - 1, generator=g).add(1).tolist()
    bt = torch.zeros(B, max_blocks, dtype=torch.int32)
    for b, s in enumerate(seq_lens):
        for j in range((s + bs - 1) // bs):
            bt[b, j] = order.pop()
    t = {
        "kc": torch.randn(nblk, KV, bs, hd, generator=g).to(BF).cuda(),
        "vc": torch.randn(nblk, KV, bs, hd, generator=g).to(BF).cuda(),
        "q": torch.randn(B, (H + 2 * KV) * hd, generator=g).to(BF).cuda(),
        "bt": bt.cuda(),
        "sl": torch.tensor(seq_lens, dtype=torch.int32).cuda(),
        "qsl": torch.arange(B + 1, dtype=torch.int32).cuda(),
    }
    return t, max_blocks


def run_reference(kge, L, t, max_blocks, B, H, KV, hd, bs, max_seq):
    out = torch.full((B, H * hd), float("nan"), dtype=BF, device="cuda")
    nws = L.kge_attn_workspace_bytes(B, H, hd, max_seq)
    ws = torch.empty(max(nws, 16), dtype=torch.uint8, device="cuda")
    kge.check(L.kge_paged_attn(out.data_ptr(), t["q"].data_ptr(), (H + 2 * KV) * hd, t["kc"].data_ptr(),
                               t["vc"].data_ptr(), t["bt"].data_ptr(), max_blocks, t["sl"].data_ptr(),
                               t["qsl"].data_ptr(), B, B, 1, max_seq, H, KV, hd, bs, hd ** -0.5,
                               ws.data_ptr(), nws, None), "kge_paged_attn")
    torch.cuda.synchronize()
    return out


def run_noalloc(kge, L, t, max_blocks, B, H, KV, hd, bs, max_seq):
    out = torch.full((B, H * hd), float("nan"), dtype=BF, device="cuda")
    kge.check(L.kge_paged_attn_decode_noalloc(out.data_ptr(), t["q"].data_ptr(), (H + 2 * KV) * hd,
                                              t["kc"].data_ptr(), t["vc"].data_ptr(), t["bt"].data_ptr(),
                                              max_blocks, t["sl"].data_ptr(), t["qsl"].data_ptr(), B, max_seq,
                                              H, KV, hd, bs, hd ** -0.5, None), "kge_paged_attn_decode_noalloc")
    torch.cuda.synchronize()
    return out


def assert_same(got, ref):
    """The tolerance of tests/test_ops_gpu.py::test_paged_attention (vLLM's bf16 IR tolerance)."""
    g, r = got.float().cpu(), ref.float().cpu()
    assert torch.isfinite(g).all()
    err = (g - r).abs()
    tol = 1e-3 * r.abs().max() + 1.6e-2 * r.abs() + 4e-3
    assert (err <= tol).all(), f"max err {err.max().item():.4g}"


def test_noalloc_decode_matches_workspace_entry_point_twice(kge):
    kge.require_cuda()
    L = kge.lib()
    H, KV, hd, bs = 32, 8, 128, 16
    seq_a = [2176, 1500, 700]                 # few requests, long contexts -> split-KV, scratch in use
    ta, mba = make_case(3, H, KV, hd, bs, seq_a, seed=1)
    ref_a = run_reference(kge, L, ta, mba, 3, H, KV, hd, bs, max(seq_a))
    first = run_noalloc(kge, L, ta, mba, 3, H, KV, hd, bs, max(seq_a))      # includes the per-device set-up
    free_before = torch.cuda.mem_get_info()[0]
    second = run_noalloc(kge, L, ta, mba, 3, H, KV, hd, bs, max(seq_a))     # stale partials from `first` in the array
    free_after = torch.cuda.mem_get_info()[0]
    assert_same(first, ref_a)
    assert_same(second, ref_a)
    assert torch.equal(first, second)
    assert free_after == free_before, "the call changed the amount of free device memory"
    # a different batch (other head_dim, other split count) reuses the same array right after
    seq_b = [900, 33, 4000, 257, 1]
    tb, mbb = make_case(5, 16, 4, 64, bs, seq_b, seed=2)
    ref_b = run_reference(kge, L, tb, mbb, 5, 16, 4, 64, bs, max(seq_b))
    assert_same(run_noalloc(kge, L, tb, mbb, 5, 16, 4, 64, bs, max(seq_b)), ref_b)
    assert_same(run_noalloc(kge, L, ta, mba, 3, H, KV, hd, bs, max(seq_a)), ref_a)


def test_noalloc_rejects_a_batch_that_does_not_fit(kge):
    kge.require_cuda()
    L = kge.lib()
    cap = L.kge_paged_attn_decode_noalloc_max_bytes()
    assert cap == 8 << 20
    B, H, KV, hd = 64, 128, 8, 128            # 2 splits -> 64*128*2*130*4 B = 8.5 MB > 8 MiB
    x = torch.zeros(64, dtype=torch.int32, device="cuda")
    rc = L.kge_paged_attn_decode_noalloc(x.data_ptr(), x.data_ptr(), 8, x.data_ptr(), x.data_ptr(), x.data_ptr(), 1,
                                         x.data_ptr(), x.data_ptr(), B, 8192, H, KV, hd, 16, 0.1, None)
    assert rc == -1
    msg = L.kge_last_error().decode()
    assert "scratch" in msg and str(cap) in msg
    # unsupported shapes are refused too, never silently routed elsewhere
    assert L.kge_paged_attn_decode_noalloc(x.data_ptr(), x.data_ptr(), 8, x.data_ptr(), x.data_ptr(), x.data_ptr(), 1,
                                           x.data_ptr(), x.data_ptr(), 2, 100, 64, 2, 128, 16, 0.1, None) == -1
