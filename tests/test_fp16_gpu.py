"""float16 arithmetic (`--dtype float16`, ref:values-01-minimal-example8.yaml:30-31, ...example9.yaml:31-32): the `_f16`
twins of the ops against the oracle run in float16, and whole engines - the Qwen2.5 / Qwen3 families those two values
files deploy, plus Llama and OPT - against the float16 oracle.

Tolerances: vLLM's float16 defaults (vllm:ir/tolerances.py: atol 1e-3, rtol 1e-3; rms_norm override atol 1e-2) for the
ops, scaled by the tensor's magnitude where the op accumulates (GEMM, attention); greedy ids equal except at the oracle's
own float16 near-ties (adjacent float16 logits near |x| ~ 4 are 0.004 apart, so the margin is 8e-3 instead of bf16's 0.06)."""
import ctypes as C

import pytest
import torch

from oracle import model as om

pytestmark = pytest.mark.gpu
H16 = torch.float16
MARGIN16 = 8e-3


@pytest.fixture(scope="module")
def L(kge):
    kge.require_cuda()
    return kge.lib()


def p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


_KEEP = []


def dev(t):
    """Device copy that stays alive until the test ends: the C ABI gets raw pointers, so a temporary freed (and its
    memory reused) before the kernel runs would feed the kernel somebody else's bytes."""
    d = t.cuda()
    _KEEP.append(d)
    return d


@pytest.fixture(autouse=True)
def _release():
    yield
    torch.cuda.synchronize()
    _KEEP.clear()


def ok(kge, rc, what):
    assert rc == 0, f"{what}: {kge.lib().kge_last_error().decode()}"


def close(got, ref, atol, rtol, frac=0.0):
    err = (got.float() - ref.float()).abs()
    bad = (err > atol + rtol * ref.float().abs()).float().mean().item()
    assert bad <= frac, f"{bad:.2e} of the elements outside atol {atol} rtol {rtol}; max err {err.max().item():.4g}"


@pytest.mark.parametrize("T,d", [(1, 768), (7, 2048), (64, 4096), (33, 8192)])
def test_norms_f16(kge, L, T, d):
    g = torch.Generator().manual_seed(T + d)
    x = torch.randn(T, d, generator=g).to(H16)
    res = torch.randn(T, d, generator=g).to(H16)
    w = (1 + 0.1 * torch.randn(d, generator=g)).to(H16)
    b = (0.1 * torch.randn(d, generator=g)).to(H16)
    y = torch.empty(T, d, dtype=H16, device="cuda")
    ok(kge, L.kge_rmsnorm_f16(p(y), p(dev(x)), p(dev(w)), T, d, 1e-5, None), "rmsnorm_f16")
    torch.cuda.synchronize()
    close(y.cpu(), om.rms_norm(x, w, 1e-5), 1e-2, 1e-3)
    rd = dev(res.clone())
    ok(kge, L.kge_add_rmsnorm_f16(p(y), p(rd), p(dev(x)), p(dev(w)), T, d, 1e-5, None), "add_rmsnorm_f16")
    torch.cuda.synchronize()
    y_ref, r_ref = om.fused_add_rms_norm(x, res, w, 1e-5)
    close(y.cpu(), y_ref, 1e-2, 1e-3)
    assert torch.equal(rd.cpu().view(torch.int16), r_ref.view(torch.int16))          # the residual is one fp32 add, rounded once
    rd = dev(res.clone())
    ok(kge, L.kge_layernorm_f16(p(y), p(rd), p(dev(x)), p(dev(w)), p(dev(b)), T, d, 1e-5, None), "layernorm_f16")
    torch.cuda.synchronize()
    h = (res.float() + x.float()).to(H16)
    close(y.cpu(), om.layer_norm(h, w, b), 1e-2, 1e-3)
    assert torch.equal(rd.cpu().view(torch.int16), h.view(torch.int16))


@pytest.mark.parametrize("T,N,K,bias,act", [(1, 768, 768, True, 0), (64, 6144, 4096, False, 0), (64, 4096, 4096, False, 0), (20, 200, 72, True, 0),
                                            (300, 2048, 2048, False, 1), (2048, 1024, 4096, True, 0), (7, 50272, 768, False, 0)])
def test_gemm_f16(kge, L, T, N, K, bias, act):
    g = torch.Generator().manual_seed(T * 7 + N * 3 + K)
    x = torch.randn(T, K, generator=g).to(H16)
    w = (torch.randn(N, K, generator=g) * 0.05).to(H16)
    b = torch.randn(N, generator=g).to(H16) if bias else None
    y = torch.full((T, N), float("nan"), dtype=H16, device="cuda")
    nws = L.kge_gemm_workspace_bytes(T, N, K)
    ws = torch.zeros(max(nws, 16), dtype=torch.uint8, device="cuda")
    for wsp, nb in ((None, 0), (p(ws), nws)):
        y.fill_(float("nan"))
        ok(kge, L.kge_gemm_f16(p(y), p(dev(x)), p(dev(w)), p(dev(b)) if bias else None, T, N, K, act, wsp, nb, None), "gemm_f16")
        torch.cuda.synchronize()
        ref = x.float() @ w.float().t()
        if bias:
            ref = ref + b.float()
        if act:
            ref = ref.relu()
        got = y.cpu().float()
        assert torch.isfinite(got).all()
        scale = ref.abs().max().item()
        assert ((got - ref).abs() <= 1e-3 * scale + 2e-3 * ref.abs()).all(), (got - ref).abs().max().item()


def test_silu_gate_rope_and_sampling_f16(kge, L):
    g = torch.Generator().manual_seed(3)
    T, f, K = 64, 1024, 512
    x = torch.randn(T, K, generator=g).to(H16)
    w = (torch.randn(2 * f, K, generator=g) * 0.05).to(H16)
    y = torch.empty(T, f, dtype=H16, device="cuda")
    ok(kge, L.kge_gemm_silu_gate_f16(p(y), p(dev(x)), p(dev(w)), T, f, K, None), "gemm_silu_gate_f16")
    torch.cuda.synchronize()
    gu = (x.float() @ w.float().t()).to(H16)
    ref = om.silu_and_mul(gu)
    close(y.cpu(), ref, 2e-3 * ref.float().abs().max().item(), 4e-3)
    y2 = torch.empty(T, f, dtype=H16, device="cuda")
    ok(kge, L.kge_silu_mul_f16(p(y2), p(dev(gu)), T, f, None), "silu_mul_f16")
    torch.cuda.synchronize()
    close(y2.cpu(), ref, 1e-3, 2e-3)
    # RoPE + KV write: neox pairs, every op rounded to float16 (common.py:160-183)
    cfg = om.CONFIGS["llama-3.2-1b"]
    Hh, KV, hd, bs, Tt = 8, 2, 64, 16, 40
    qkv = torch.randn(Tt, (Hh + 2 * KV) * hd, generator=g).to(H16)
    pos = torch.arange(100, 100 + Tt, dtype=torch.int32)
    cs = om.rope_cos_sin_cache(cfg, 256, H16)
    slots = torch.arange(16, 16 + Tt, dtype=torch.int64)
    kc = torch.zeros(8, KV, bs, hd, dtype=H16, device="cuda")
    vc = torch.zeros(8, KV, bs, hd, dtype=H16, device="cuda")
    qd = dev(qkv.clone())
    ok(kge, L.kge_rope_kvwrite_f16(p(qd), p(dev(cs)), p(dev(pos)), p(dev(slots)), p(kc), p(vc), Tt, Hh, KV, hd, bs, None), "rope_kvwrite_f16")
    torch.cuda.synchronize()
    c, s_ = cs.index_select(0, pos.long()).chunk(2, dim=-1)
    q_ref = om.apply_rope_neox(qkv[:, :Hh * hd].reshape(Tt, Hh, hd), c, s_).reshape(Tt, Hh * hd)
    k_ref = om.apply_rope_neox(qkv[:, Hh * hd:(Hh + KV) * hd].reshape(Tt, KV, hd), c, s_)
    got = qd.cpu()
    close(got[:, :Hh * hd], q_ref, 1e-3, 2e-3, frac=5e-3)
    kcache = kc.cpu().permute(0, 2, 1, 3).reshape(-1, KV, hd)[16:16 + Tt]
    close(kcache, k_ref, 1e-3, 2e-3, frac=5e-3)
    vcache = vc.cpu().permute(0, 2, 1, 3).reshape(-1, KV, hd)[16:16 + Tt]
    assert torch.equal(vcache.reshape(Tt, -1).view(torch.int16), qkv[:, (Hh + KV) * hd:].contiguous().view(torch.int16))
    # argmax / top-k=1 sampling read float16 logits
    B, V = 5, 50272
    lg = torch.randn(B, V, generator=g).to(H16)
    out = torch.zeros(B, dtype=torch.int32, device="cuda")
    ok(kge, L.kge_argmax_f16(p(out), p(dev(lg)), B, V, V, None), "argmax_f16")
    torch.cuda.synchronize()
    assert out.cpu().tolist() == lg.float().argmax(-1).tolist()


@pytest.mark.parametrize("H,KV,hd,bs,q_lens,seq_lens", [
    (32, 8, 128, 64, [1] * 3, [2176, 1500, 700]),          # decode, TMA boxes of one 64-token block, split-KV + merge
    (32, 8, 128, 16, [1] * 4, [1, 17, 64, 300]),           # decode, 16-token blocks (cp.async gathers)
    (28, 4, 128, 16, [20, 1], [20, 300]),                  # Qwen2.5-7B group of 7, mixed
    (16, 8, 128, 64, [300, 64], [300, 64]),                # Qwen3 shape prefill
    (32, 8, 64, 16, [1, 30, 1, 7], [77, 30, 256, 135]),
    (32, 8, 128, 16, [2048], [2048]),
])
def test_paged_attention_f16(kge, L, H, KV, hd, bs, q_lens, seq_lens):
    g = torch.Generator().manual_seed(H + sum(seq_lens))
    B, T = len(q_lens), sum(q_lens)
    max_blocks = max((s + bs - 1) // bs for s in seq_lens)
    nblk = sum((s + bs - 1) // bs for s in seq_lens) + 3
    free = torch.randperm(nblk - 1, generator=g).add(1).tolist()
    kc = torch.randn(nblk, KV, bs, hd, generator=g).to(H16)
    vc = torch.randn(nblk, KV, bs, hd, generator=g).to(H16)
    bt = torch.zeros(B, max_blocks, dtype=torch.int32)
    for b, s in enumerate(seq_lens):
        for j in range((s + bs - 1) // bs):
            bt[b, j] = free.pop()
    qw = (H + 2 * KV) * hd
    qkv = torch.randn(T, qw, generator=g).to(H16)
    qsl = [0]
    for ql in q_lens:
        qsl.append(qsl[-1] + ql)
    out = torch.full((T, H * hd), float("nan"), dtype=H16, device="cuda")
    nws = L.kge_attn_workspace_bytes(B, H, hd, max(seq_lens))
    ws = torch.empty(max(nws, 16), dtype=torch.uint8, device="cuda")
    ok(kge, L.kge_paged_attn_f16(p(out), p(dev(qkv)), qw, p(dev(kc)), p(dev(vc)), p(dev(bt)), max_blocks, p(dev(torch.tensor(seq_lens, dtype=torch.int32))),
                                 p(dev(torch.tensor(qsl, dtype=torch.int32))), B, T, max(q_lens), max(seq_lens), H, KV, hd, bs, hd ** -0.5, p(ws), nws, None),
       "paged_attn_f16")
    torch.cuda.synchronize()
    got = out.cpu().float()
    assert torch.isfinite(got).all()
    for b in range(B):
        s, ql = seq_lens[b], q_lens[b]
        blocks = bt[b, :(s + bs - 1) // bs].long()
        k = kc[blocks].permute(0, 2, 1, 3).reshape(-1, KV, hd)[:s]
        v = vc[blocks].permute(0, 2, 1, 3).reshape(-1, KV, hd)[:s]
        q = qkv[qsl[b]:qsl[b + 1], :H * hd].reshape(ql, H, hd)
        ref = torch.cat([om.causal_attention(q[c0:c0 + 256], k[:s - ql + min(ql, c0 + 256)], v[:s - ql + min(ql, c0 + 256)], hd ** -0.5, s - ql + c0)
                         for c0 in range(0, ql, 256)], 0).float()
        err = (got[qsl[b]:qsl[b + 1]] - ref).abs()
        # P is rounded to float16 before the PV product (as a float16 flash kernel does): 2^-11 relative per term
        assert (err <= 1e-3 * ref.abs().max() + 4e-3 * ref.abs() + 1e-3).all(), f"request {b}: max err {err.max().item():.4g}"


def compare16(got, ref_ids, ref_logits):
    equal = near = 0
    for b, (g_, r) in enumerate(zip(got, ref_ids)):
        for s, (a, e) in enumerate(zip(g_, r)):
            if a == e:
                equal += 1
                continue
            top2 = ref_logits[s][b].topk(2).values
            assert (top2[0] - top2[1]).item() <= MARGIN16, f"request {b} step {s}: {a} vs {e}, oracle gap {(top2[0] - top2[1]).item():.4f}"
            assert abs(ref_logits[s][b][a].item() - top2[0].item()) <= MARGIN16
            near += 1
            break
    return equal, near


@pytest.mark.parametrize("name,stored,how", [
    ("tiny-qwen2", "float16", "auto"),          # a float16 checkpoint, --dtype auto: stays float16 (the ex8 model family)
    ("tiny-qwen3", "float16", "auto"),          # the ex9 model family
    ("tiny-qwen2", "bfloat16", "float16"),      # a bfloat16 checkpoint served with --dtype float16, as ex8 / ex9 do
    ("tiny-llama", "float16", "float16"),
    ("tiny-llama-hd128", "bfloat16", "float16"),
    ("tiny-opt", "float16", "auto"),
])
def test_float16_engine_matches_float16_oracle(kge, tmp_path, name, stored, how):
    cfg = om.CONFIGS[name]
    sd = H16 if stored == "float16" else torch.bfloat16
    w = om.make_weights(cfg, seed=0, dtype=sd)
    d = str(tmp_path / "m")
    om.save_model_dir(d, cfg, w, torch_dtype=stored)
    prompts = [om.make_prompts(cfg, 1, n, seed=n)[0] for n in (5, 16, 17, 33, 64, 100)]
    N = 20
    ref_ids, ref_lg = om.OracleModel(cfg, w, H16).generate_greedy(prompts, N, return_logits=True)   # weights converted as the loader does
    llm = kge.LLM(d, max_model_len=200, max_num_seqs=8, max_num_batched_tokens=128, num_gpu_blocks_override=128, dtype=how,
                  enable_prefix_caching=False)
    assert llm.dtype == "float16"
    outs = llm.generate(prompts, kge.SamplingParams(max_tokens=N, temperature=0.0, ignore_eos=True))
    # teacher-forced logits of the first prompt's prefill, float16 tolerance
    llm.close()
    got = [o.outputs[0].token_ids for o in outs]
    equal, near = compare16(got, ref_ids, ref_lg)
    assert near <= 2 and equal >= 0.75 * len(prompts) * N, (equal, near)
    # the same directory under --dtype bfloat16 is a different arithmetic: it must still run (and say what it runs in)
    llm = kge.LLM(d, max_model_len=200, max_num_seqs=8, max_num_batched_tokens=128, num_gpu_blocks_override=128, dtype="bfloat16")
    assert llm.dtype == "bfloat16"
    llm.generate(prompts[:1], kge.SamplingParams(max_tokens=2, temperature=0.0, ignore_eos=True))
    llm.close()


def test_float16_logits_tolerance(kge, tmp_path):
    """Prefill and teacher-forced decode logits of a float16 engine within vLLM's float16 tolerance (scaled by the logits'
    magnitude) of the float16 oracle: max-abs <= 4e-3 * max|logit| + 2e-3 * |ref|."""
    cfg = om.CONFIGS["small-llama"]
    w = om.make_weights(cfg, seed=0, dtype=H16)
    d = str(tmp_path / "m")
    om.save_model_dir(d, cfg, w, torch_dtype="float16")
    prompts = om.make_prompts(cfg, 2, 40, seed=9)
    ref_ids, ref_lg = om.OracleModel(cfg, w, H16).generate_greedy(prompts, 6, return_logits=True)
    llm = kge.LLM(d, max_model_len=128, max_num_seqs=4, max_num_batched_tokens=128, num_gpu_blocks_override=32)
    bs, bt = 16, [[1, 2, 3], [4, 5, 6]]
    ids = prompts[0] + prompts[1]
    slots = [bt[b][t // bs] * bs + t % bs for b in range(2) for t in range(40)]
    out, lg = llm.raw_step(ids, list(range(40)) * 2, [0, 40, 80], [40, 40], bt, 3, slots, [39, 79], want_logits=True)
    for s in range(6):
        ref, got = ref_lg[s], torch.from_numpy(lg)
        scale = ref.abs().max().item()
        err = (got - ref).abs()
        assert (err <= 4e-3 * scale + 2e-3 * ref.abs()).all(), f"step {s}: max-abs {err.max().item():.4g} (scale {scale:.3g})"
        if s == 5:
            break
        t = 40 + s
        out, lg = llm.raw_step([ref_ids[0][s], ref_ids[1][s]], [t, t], [0, 1, 2], [t + 1, t + 1], bt, 3,
                               [bt[b][t // bs] * bs + t % bs for b in range(2)], [0, 1], want_logits=True)
    llm.close()
