"""End-to-end parity of the engine (C++ scheduler + paged KV + sm_100a kernels, through the C ABI)
against the CPU oracle on identical seeded safetensors and token-id prompts.

Bar (BASELINE.json north_star): greedy ids bit-exact for the first N tokens while the oracle's own
top-1/top-2 margin exceeds bf16 noise; logits within bf16 tolerance (max-abs stated per test).
A mismatch is only accepted at a step where the oracle's top-2 margin is below `MARGIN` (two fp32
GEMM summation orders legitimately round differently there); after such a step the sequences
diverge by construction and the comparison for that request stops.
"""
import pytest
import torch

from oracle import model as om

pytestmark = pytest.mark.gpu
BF = torch.bfloat16
MARGIN = 0.06      # logits are bf16: adjacent values near |x|~4 are 0.03 apart


@pytest.fixture(scope="module")
def model_dirs(tmp_path_factory):
    """name -> (cfg, weights, directory) written once per session."""
    cache = {}

    def get(name, seed=0):
        key = (name, seed)
        if key not in cache:
            cfg = om.CONFIGS[name]
            w = om.make_weights(cfg, seed=seed)
            d = tmp_path_factory.mktemp(f"{name}-{seed}")
            om.save_model_dir(str(d), cfg, w)
            cache[key] = (cfg, w, str(d))
        return cache[key]
    return get


def compare_greedy(got, ref_ids, ref_logits, margin_of=None):
    """Returns (#tokens compared equal, #requests that diverged at a near-tie).  margin_of(row of oracle logits) gives the
    largest oracle gap at which a different greedy choice is still consistent with the logits tolerance (default MARGIN)."""
    equal = near_tie = 0
    for b, (g, r) in enumerate(zip(got, ref_ids)):
        assert len(g) == len(r)
        for s, (a, e) in enumerate(zip(g, r)):
            if a == e:
                equal += 1
                continue
            lim = margin_of(ref_logits[s][b]) if margin_of else MARGIN
            top2 = ref_logits[s][b].topk(2).values
            margin = (top2[0] - top2[1]).item()
            assert margin <= lim, f"request {b} step {s}: got {a}, oracle {e}, oracle margin {margin:.4f}"
            assert abs(ref_logits[s][b][a].item() - top2[0].item()) <= lim
            near_tie += 1
            break
    return equal, near_tie


def logits_bound_margin(row):
    """Two logits, each within 3e-2 * max|logit| + 1.6e-2 * |logit| of the oracle's, can swap order when the oracle's gap
    between them is at most the sum of the two bounds."""
    top = row.topk(1).values.item()
    return 2.0 * (3e-2 * row.abs().max().item() + 1.6e-2 * abs(top))


@pytest.mark.parametrize("name", ["tiny-opt", "tiny-llama", "tiny-llama-hd128", "tiny-qwen2", "tiny-qwen3"])
def test_greedy_generation_matches_oracle(kge, model_dirs, name):
    cfg, w, d = model_dirs(name)
    prompts = [om.make_prompts(cfg, 1, n, seed=n)[0] for n in (5, 16, 17, 33, 64, 100)]
    N = 24
    ref_ids, ref_lg = om.OracleModel(cfg, w, BF).generate_greedy(prompts, N, return_logits=True)
    llm = kge.LLM(d, max_model_len=200, max_num_seqs=8, max_num_batched_tokens=128, num_gpu_blocks_override=128)
    outs = llm.generate(prompts, kge.SamplingParams(max_tokens=N, temperature=0.0, ignore_eos=True))
    assert llm.kernel_launches > 0
    llm.close()
    got = [o.outputs[0].token_ids for o in outs]
    assert all(o.outputs[0].finish_reason == "length" for o in outs)
    equal, near = compare_greedy(got, ref_ids, ref_lg)
    assert near <= 2 and equal >= 0.75 * len(prompts) * N, (equal, near)   # every divergence is an asserted near-tie


@pytest.mark.parametrize("name", ["small-opt", "small-llama"])
def test_real_width_models_logits_and_ids(kge, model_dirs, name):
    """opt-125m / llama-3.2-1b widths and vocabularies (2 layers): prefill logits within bf16
    tolerance of the oracle (max-abs <= 3e-2 * max|logit| + 1.6e-2 * |ref|), KL <= 1e-3, and the
    first 8 greedy tokens equal."""
    cfg, w, d = model_dirs(name)
    prompts = om.make_prompts(cfg, 2, 40, seed=9)
    m = om.OracleModel(cfg, w, BF)
    ref_ids, ref_lg = m.generate_greedy(prompts, 8, return_logits=True)
    llm = kge.LLM(d, max_model_len=128, max_num_seqs=4, max_num_batched_tokens=128, num_gpu_blocks_override=32)
    # raw step: both prompts prefilled in one batch, logits read back
    bs = 16
    ids = prompts[0] + prompts[1]
    pos = list(range(40)) * 2
    bt = [[1, 2, 3], [4, 5, 6]]
    slots = [bt[b][t // bs] * bs + t % bs for b in range(2) for t in range(40)]
    out, lg = llm.raw_step(ids, pos, [0, 40, 80], [40, 40], bt, 3, slots, [39, 79], want_logits=True)
    ref = ref_lg[0]
    got = torch.from_numpy(lg)
    scale = ref.abs().max().item()
    err = (got - ref).abs()
    assert (err <= 3e-2 * scale + 1.6e-2 * ref.abs()).all(), f"max-abs {err.max().item():.4g} vs scale {scale:.3g}"
    kl = torch.nn.functional.kl_div(torch.log_softmax(got, -1), torch.log_softmax(ref, -1), log_target=True,
                                    reduction="batchmean").item()
    assert kl <= 1e-3, kl
    assert out.tolist() == got.argmax(-1).tolist()          # the engine's ids are the argmax of its own logits
    for b in range(2):
        t2 = ref[b].topk(2).values
        assert out[b] == ref_ids[b][0] or (t2[0] - t2[1]).item() <= MARGIN, (b, out[b], ref_ids[b][0])
    # teacher-forced decode: feed the ORACLE's tokens so a near-tie cannot derail the comparison, and
    # check every decode step's logits (decode attention + split-K GEMMs at real width)
    for s in range(1, 8):
        ids = [ref_ids[0][s - 1], ref_ids[1][s - 1]]
        t = 40 + s - 1
        out, lg = llm.raw_step(ids, [t, t], [0, 1, 2], [t + 1, t + 1], bt, 3,
                               [bt[b][t // bs] * bs + t % bs for b in range(2)], [0, 1], want_logits=True)
        ref, got = ref_lg[s], torch.from_numpy(lg)
        err = (got - ref).abs()
        assert (err <= 3e-2 * scale + 1.6e-2 * ref.abs()).all(), f"decode step {s}: max-abs {err.max().item():.4g}"
        for b in range(2):
            t2 = ref[b].topk(2).values
            assert out[b] == ref_ids[b][s] or (t2[0] - t2[1]).item() <= MARGIN, (s, b)
    llm.close()
    llm = kge.LLM(d, max_model_len=128, max_num_seqs=4, max_num_batched_tokens=128, num_gpu_blocks_override=32)
    outs = llm.generate(prompts, kge.SamplingParams(max_tokens=8, temperature=0.0, ignore_eos=True))
    llm.close()
    # free-running: with a 50k-128k vocabulary and random weights the oracle's own bf16 top-2 gap is
    # below MARGIN on a large share of steps, so each request may leave the oracle's path at such a
    # step (compare_greedy asserts that every divergence IS such a near-tie)
    equal, near = compare_greedy([o.outputs[0].token_ids for o in outs], ref_ids, ref_lg)
    assert near <= 2, (equal, near)


# ----------------------------------------------------------------- BASELINE.json configurations at full size
# Row n1 of the scope table: parity on the models the bench is quoted on, not on toy shapes.
#   C1  facebook/opt-125m, all 12 layers, 1 x 128-in/128-out    (ref:values-01-minimal-example.yaml:7)
#   C2  Llama-3.2-1B, all 16 layers, 4 x 128-in/32-out
#   C3  Llama-3.1-8B width (d 4096, 32/8 heads of 128, ffn 14336, vocab 128256), 3 layers, one 2048-token prompt
#       prefilled in ONE step (cta_group::2 GEMMs + tcgen05 attention over 32 key tiles) + 8 decode steps
import dataclasses
import json
import os

BASELINE_CASES = {
    # name: (oracle config, requests, prompt length, teacher-forced decode steps, free-running tokens)
    "c1-opt-125m": (om.CONFIGS["opt-125m"], 1, 128, 127, 128),
    "c2-llama-3.2-1b": (om.CONFIGS["llama-3.2-1b"], 4, 128, 31, 32),
    "c3-llama-3.1-8b-width-3-layers": (dataclasses.replace(om.CONFIGS["llama-3.1-8b"], num_layers=3), 1, 2048, 8, 9),
}


def _record(name, **kw):
    """Measured parity figures, kept next to the run's other outputs when that directory exists."""
    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "parity_r02.jsonl"), "a") as fh:
            fh.write(json.dumps(dict(case=name, **kw)) + "\n")


@pytest.mark.parametrize("case", sorted(BASELINE_CASES))
def test_baseline_config_parity(kge, tmp_path, case):
    """Bar (north_star): logits within bf16 tolerance of the oracle - |err| <= 3e-2 * max|logit| + 1.6e-2 * |ref| for all but
    1e-4 of a step's logits and twice that for every one, KL(engine || oracle) <= 1e-3 per row - at the prefill and at EVERY decode step (the oracle's tokens are fed back, so a
    near-tie cannot derail the comparison); greedy ids equal to the oracle's wherever the oracle's own bf16 top-2 gap
    exceeds MARGIN; and, free-running through the scheduler, every divergence is such a near-tie."""
    cfg, n, pin, n_dec, n_free = BASELINE_CASES[case]
    w = om.make_weights(cfg, seed=0)
    d = str(tmp_path / "model")
    om.save_model_dir(d, cfg, w)
    prompts = om.make_prompts(cfg, n, pin, seed=11)
    m = om.OracleModel(cfg, w, BF, max_pos=pin + n_free + 8)
    ref_ids, ref_lg = m.generate_greedy(prompts, n_free, return_logits=True)
    del m, w
    bs = 16
    nb = (pin + n_free + bs - 1) // bs
    llm = kge.LLM(d, max_model_len=pin + n_free + 8, max_num_seqs=max(n, 4), max_num_batched_tokens=max(256, n * pin),
                  num_gpu_blocks_override=n * nb + 2, enable_prefix_caching=False)
    bt = [[1 + b * nb + j for j in range(nb)] for b in range(n)]
    slot = lambda b, t: bt[b][t // bs] * bs + t % bs      # noqa: E731
    worst_abs = worst_kl = worst_out = 0.0
    tie_steps = 0

    def check(step, out, lg):
        nonlocal worst_abs, worst_kl, tie_steps, worst_out
        ref, got = ref_lg[step], torch.from_numpy(lg)
        scale = ref.abs().max().item()
        err = (got - ref).abs()
        # n x 50k-128k logits per step through 12-16 layers: the tail of bf16 rounding noise between two correct
        # implementations reaches past the two-layer bound on a few logits in a hundred thousand (measured: <= 1.1x),
        # so the bound holds for all but 1e-4 of them and twice the bound for every one
        tol = 3e-2 * scale + 1.6e-2 * ref.abs()
        outside = (err > tol).float().mean().item()
        assert outside <= 1e-4 and (err <= 2 * tol).all(), \
            f"{case} step {step}: {outside:.2e} of the logits outside the bound, max-abs {err.max().item():.4g} (scale {scale:.3g})"
        worst_out = max(worst_out, outside)
        kl = torch.nn.functional.kl_div(torch.log_softmax(ref, -1), torch.log_softmax(got, -1), log_target=True,
                                        reduction="none").sum(-1).max().item()
        assert kl <= 1e-3, f"{case} step {step}: KL {kl:.3g}"
        worst_abs, worst_kl = max(worst_abs, err.max().item() / scale), max(worst_kl, kl)
        assert out.tolist() == got.argmax(-1).tolist()
        for b in range(n):
            if out[b] != ref_ids[b][step]:
                # both tokens' logits are within the bound above, so the greedy choice may legitimately differ exactly when
                # the oracle's own gap between them is no larger than the two bounds together
                a_, e_ = int(out[b]), ref_ids[b][step]
                gap = (ref[b][e_] - ref[b][a_]).item()
                assert gap <= (tol[b][a_] + tol[b][e_]).item(), f"{case} step {step} request {b}: {a_} vs {e_}, oracle gap {gap:.4f}"
                tie_steps += 1

    qsl = [b * pin for b in range(n + 1)]
    out, lg = llm.raw_step([t for p in prompts for t in p], list(range(pin)) * n, qsl, [pin] * n, bt, nb,
                           [slot(b, t) for b in range(n) for t in range(pin)], [q - 1 for q in qsl[1:]], want_logits=True)
    check(0, out, lg)
    for s in range(1, n_dec + 1):
        t = pin + s - 1
        out, lg = llm.raw_step([ref_ids[b][s - 1] for b in range(n)], [t] * n, list(range(n + 1)), [t + 1] * n, bt, nb,
                               [slot(b, t) for b in range(n)], list(range(n)), want_logits=True)
        check(s, out, lg)
    llm.close()
    # free-running through the scheduler (chunked prefill, paged KV, graphs): the first tokens are the oracle's
    llm = kge.LLM(d, max_model_len=pin + n_free + 8, max_num_seqs=max(n, 4), max_num_batched_tokens=max(256, n * pin),
                  num_gpu_blocks_override=n * nb + 2, enable_prefix_caching=False)
    outs = llm.generate(prompts, kge.SamplingParams(max_tokens=n_free, temperature=0.0, ignore_eos=True))
    llm.close()
    equal, near = compare_greedy([o.outputs[0].token_ids for o in outs], ref_ids, ref_lg, logits_bound_margin)   # asserts each divergence is one
    _record(case, max_abs_over_scale=worst_abs, max_fraction_outside_bound=worst_out, max_kl=worst_kl, teacher_forced_steps=n_dec + 1, near_tie_steps=tie_steps,
            free_running_equal=equal, free_running_total=n * n_free, free_running_near_tie_divergences=near)


def test_chunked_prefill_and_preemption_do_not_change_tokens(kge, model_dirs):
    """Scheduling must not change values: a roomy engine, one with a 32-token step budget (prompts
    split across steps and mixed with decodes) and one whose KV pool forces preemption + recompute
    all emit the same greedy ids."""
    cfg, w, d = model_dirs("tiny-llama")
    prompts = [om.make_prompts(cfg, 1, n, seed=100 + n)[0] for n in (70, 9, 45, 33, 81, 20)]
    sp = kge.SamplingParams(max_tokens=20, temperature=0.0, ignore_eos=True)
    results = []
    stats = []
    for kw in (dict(max_num_batched_tokens=512, num_gpu_blocks_override=128),
               dict(max_num_batched_tokens=32, num_gpu_blocks_override=128),
               dict(max_num_batched_tokens=64, num_gpu_blocks_override=22)):
        llm = kge.LLM(d, max_model_len=128, max_num_seqs=8, **kw)
        outs = llm.generate(prompts, sp)
        stats.append(llm.stats())
        llm.close()
        results.append([o.outputs[0].token_ids for o in outs])
    assert stats[2]["num_preemptions"] > 0, "the small pool was meant to preempt"
    ref_ids, ref_lg = om.OracleModel(cfg, w, BF).generate_greedy(prompts, 20, return_logits=True)
    for r in results:
        equal, near = compare_greedy(r, ref_ids, ref_lg)     # asserts every mismatch is a near-tie
        assert near <= 2 and equal >= 0.75 * 6 * 20, (equal, near)
    # between engine configurations the arithmetic per token is the same kernels on the same values,
    # except the split of a prompt into chunks (general vs decode attention kernel): near-tie only
    for r in results[1:]:
        diff = sum(a != b for x, y in zip(results[0], r) for a, b in zip(x, y))
        assert diff <= 20, diff


def test_replay_of_recorded_steps_reproduces_the_job(kge, model_dirs):
    """bench.py's `value` leg: the steps the scheduler made for a job (chunked prompts mixed with decodes, a preemption,
    requests of different lengths finishing at different steps) are recorded and replayed device-resident - same batches,
    same kernels, sampled ids handed over on the device - and must give exactly the recorded job's tokens, every time."""
    cfg, w, d = model_dirs("tiny-llama")
    prompts = [om.make_prompts(cfg, 1, n, seed=300 + n)[0] for n in (70, 9, 45, 33, 81, 20)]
    sps = [kge.SamplingParams(max_tokens=m, temperature=0.0, ignore_eos=True) for m in (20, 7, 15, 20, 3, 11)]
    for kw in (dict(max_num_batched_tokens=512, num_gpu_blocks_override=128), dict(max_num_batched_tokens=32, num_gpu_blocks_override=128),
               dict(max_num_batched_tokens=64, num_gpu_blocks_override=22)):
        llm = kge.LLM(d, max_model_len=128, max_num_seqs=8, enable_prefix_caching=False, **kw)
        llm.record_steps(True)
        outs = llm.generate(prompts, sps)
        llm.record_steps(False)
        want = {o.request_id: o.outputs[0].token_ids for o in outs}
        for _ in range(2):
            got, steps, ms, dms = llm.replay_recorded(256)
            assert got == want and steps > 0 and ms > 0 and 0 <= dms <= ms
        # a random job replays too: the per-row noise seeds travel with the recorded metadata
        rsp = kge.SamplingParams(max_tokens=9, temperature=1.2, top_p=0.9, seed=11, ignore_eos=True)
        llm.record_steps(True)
        outs = llm.generate(prompts[:3], rsp)
        llm.record_steps(False)
        got, _, _, _ = llm.replay_recorded(64)
        assert got == {o.request_id: o.outputs[0].token_ids for o in outs}
        llm.close()


def test_prefix_caching_reuses_blocks_and_keeps_tokens(kge, model_dirs):
    """--enable-prefix-caching: requests that share a prompt prefix (and a second job with the same prompts) take the
    shared full blocks from the pool instead of recomputing them; the tokens are those of the oracle (near-ties
    aside) and of the engine without caching."""
    cfg, w, d = model_dirs("tiny-llama")
    shared = om.make_prompts(cfg, 1, 48, seed=7)[0]                       # three full blocks of 16
    prompts = [shared + om.make_prompts(cfg, 1, n, seed=200 + n)[0] for n in (5, 17, 30)] + [shared[:40], list(shared)]
    sp = kge.SamplingParams(max_tokens=12, temperature=0.0, ignore_eos=True)
    ref_ids, ref_lg = om.OracleModel(cfg, w, BF).generate_greedy(prompts, 12, return_logits=True)
    plain = kge.LLM(d, max_model_len=128, max_num_seqs=8, max_num_batched_tokens=64, num_gpu_blocks_override=64,
                    enable_prefix_caching=False)
    base = [o.outputs[0].token_ids for o in plain.generate(prompts, sp)]
    assert plain.prefix_cache_stats() == (0, 0)
    plain.close()
    llm = kge.LLM(d, max_model_len=128, max_num_seqs=8, max_num_batched_tokens=64, num_gpu_blocks_override=64,
                  enable_prefix_caching=True)
    first = [o.outputs[0].token_ids for o in llm.generate(prompts, sp)]
    q1, h1 = llm.prefix_cache_stats()
    # a 64-token step admits the first prompt alone; the others arrive after its blocks are registered and hit them
    assert q1 == sum(len(p) for p in prompts) and h1 >= 48 + 48 + 32 + 32, (q1, h1)
    second = [o.outputs[0].token_ids for o in llm.generate(prompts, sp)]   # everything but each prompt's last block is cached now
    q2, h2 = llm.prefix_cache_stats()
    assert h2 - h1 >= sum((len(p) - 1) // 16 * 16 for p in prompts), (h1, h2)
    assert llm.stats()["num_running"] == 0
    llm.close()
    for got in (first, second):
        equal, near = compare_greedy(got, ref_ids, ref_lg)      # asserts every mismatch is a near-tie of the oracle
        assert near <= 2 and equal >= 0.75 * len(prompts) * 12, (equal, near)
        assert sum(a != b for x, y in zip(base, got) for a, b in zip(x, y)) <= 24


def test_stop_rules_eos_and_stop_ids(kge, model_dirs):
    cfg, w, d = model_dirs("tiny-llama")
    prompts = om.make_prompts(cfg, 1, 12, seed=4)
    ref = om.OracleModel(cfg, w, BF).generate_greedy(prompts, 12)[0]
    llm = kge.LLM(d, max_model_len=64, max_num_seqs=4, max_num_batched_tokens=64, num_gpu_blocks_override=16)
    stop_tok = ref[5]
    first = ref.index(stop_tok)
    out = llm.generate(prompts, kge.SamplingParams(max_tokens=12, temperature=0.0, ignore_eos=True,
                                                   stop_token_ids=[stop_tok]))[0].outputs[0]
    assert out.finish_reason == "stop" and out.token_ids == ref[:first + 1]
    out = llm.generate(prompts, kge.SamplingParams(max_tokens=12, min_tokens=first + 2, temperature=0.0,
                                                   ignore_eos=True, stop_token_ids=[stop_tok]))[0].outputs[0]
    assert len(out.token_ids) >= first + 2
    out = llm.generate(prompts, kge.SamplingParams(max_tokens=3, temperature=0.0, ignore_eos=True))[0].outputs[0]
    assert out.finish_reason == "length" and out.token_ids == ref[:3]
    # max_model_len cap: prompt 12 + outputs stops at 64 total tokens
    out = llm.generate(prompts, kge.SamplingParams(max_tokens=500, temperature=0.0, ignore_eos=True))[0].outputs[0]
    assert out.finish_reason == "length" and len(out.token_ids) == 64 - 12
    llm.close()


def test_min_tokens_masks_stop_tokens_in_the_logits(kge, model_dirs):
    """vLLM's MinTokensLogitsProcessor: until a request has min_tokens outputs, its EOS and stop ids are not merely ignored
    as stop conditions - their logits are -inf, so they cannot be sampled at all.  Take a token greedy decoding first emits
    at some step k, make it a stop id with min_tokens = 8: it must be absent from the first 8 outputs (the runner-up is taken
    instead), and the request stops at its first occurrence afterwards."""
    cfg, w, d = model_dirs("tiny-llama")
    prompts = om.make_prompts(cfg, 2, 12, seed=4)
    llm = kge.LLM(d, max_model_len=96, max_num_seqs=4, max_num_batched_tokens=64, num_gpu_blocks_override=32, enable_prefix_caching=False)
    # each request alone, as the masked runs below are: the same batches through the same kernels, so any difference is the mask's
    free = [llm.generate([pr], kge.SamplingParams(max_tokens=40, temperature=0.0, ignore_eos=True))[0].outputs[0].token_ids for pr in prompts]
    for b in range(2):
        # the first token greedy decoding emits at a step >= 2 that it has not emitted before (a tiny random model repeats itself)
        k = next(i for i in range(2, 8) if free[b][i] not in free[b][:i])
        banned = free[b][k]
        out = llm.generate([prompts[b]], kge.SamplingParams(max_tokens=40, min_tokens=8, temperature=0.0, ignore_eos=True,
                                                            stop_token_ids=[banned]))[0].outputs[0]
        assert out.token_ids[:k] == free[b][:k]                      # untouched until the banned token would have come
        assert banned not in out.token_ids[:8], (banned, out.token_ids)
        assert out.token_ids[k] != banned
        if banned in out.token_ids:
            assert out.finish_reason == "stop" and out.token_ids.index(banned) == len(out.token_ids) - 1 >= 8
        # a random-sampling request obeys the mask too
        out = llm.generate([prompts[b]], kge.SamplingParams(max_tokens=12, min_tokens=12, temperature=1.0, top_k=2, seed=3, ignore_eos=True,
                                                            stop_token_ids=[banned, free[b][0]]))[0].outputs[0]
        assert banned not in out.token_ids and free[b][0] not in out.token_ids and len(out.token_ids) == 12
    llm.close()


def test_random_sampling_respects_top_k(kge, model_dirs):
    cfg, w, d = model_dirs("tiny-llama")
    prompts = om.make_prompts(cfg, 4, 10, seed=6)
    llm = kge.LLM(d, max_model_len=64, max_num_seqs=4, max_num_batched_tokens=64, num_gpu_blocks_override=32)
    a = llm.generate(prompts, kge.SamplingParams(max_tokens=6, temperature=1.0, top_k=1, ignore_eos=True))
    g = llm.generate(prompts, kge.SamplingParams(max_tokens=6, temperature=0.0, ignore_eos=True))
    assert [o.outputs[0].token_ids for o in a] == [o.outputs[0].token_ids for o in g]   # top_k=1 == greedy
    r1 = llm.generate(prompts, kge.SamplingParams(max_tokens=6, temperature=1.5, seed=5, ignore_eos=True))
    assert all(len(o.outputs[0].token_ids) == 6 for o in r1)
    llm.close()


def test_seeded_sampling_is_independent_of_other_traffic(kge, model_dirs):
    """A request with its own seed draws its noise from (seed, index of the token being sampled) alone - as vLLM's
    per-request generator does (vllm:v1/sample/ops/topk_topp_sampler.py:180-184) - so the same seeded request gives the
    same tokens whether it runs alone on a fresh engine, after other jobs, or batched with other requests.  seed=0 is a
    seed like any other."""
    cfg, w, d = model_dirs("tiny-llama")
    prompts = om.make_prompts(cfg, 5, 14, seed=21)
    mk = lambda: kge.LLM(d, max_model_len=96, max_num_seqs=8, max_num_batched_tokens=64, num_gpu_blocks_override=64,      # noqa: E731
                         enable_prefix_caching=False)
    for seed in (0, 12345):
        sp = kge.SamplingParams(max_tokens=16, temperature=1.3, top_p=0.95, seed=seed, ignore_eos=True)
        llm = mk()
        alone = llm.generate([prompts[0]], sp)[0].outputs[0].token_ids
        # other traffic first (advances the engine's step counter), then the same request inside a batch of unseeded ones
        llm.generate(prompts[1:3], kge.SamplingParams(max_tokens=9, temperature=1.0, ignore_eos=True))
        other = kge.SamplingParams(max_tokens=16, temperature=0.8, ignore_eos=True)
        mixed = llm.generate([prompts[3], prompts[0], prompts[4]], [other, sp, other])[1].outputs[0].token_ids
        again = llm.generate([prompts[0]], sp)[0].outputs[0].token_ids
        llm.close()
        assert alone == mixed == again, (seed, alone, mixed, again)
    llm = mk()
    a = llm.generate([prompts[0]], kge.SamplingParams(max_tokens=16, temperature=1.3, seed=0, ignore_eos=True))[0].outputs[0].token_ids
    b = llm.generate([prompts[0]], kge.SamplingParams(max_tokens=16, temperature=1.3, seed=1, ignore_eos=True))[0].outputs[0].token_ids
    llm.close()
    assert a != b        # different seeds, different streams (16 tokens at temperature 1.3 over a 1024-token vocabulary)


def test_engine_rejects_what_it_cannot_serve(kge, model_dirs, tmp_path):
    import json
    cfg, w, d = model_dirs("tiny-llama")
    with pytest.raises(kge.KgeError, match="max_model_len"):
        kge.LLM(d, max_model_len=10 ** 6)
    llm = kge.LLM(d, max_model_len=64, max_num_seqs=4, max_num_batched_tokens=64, num_gpu_blocks_override=8)
    with pytest.raises(kge.KgeError, match="longer than max_model_len"):
        llm.add_request(list(range(3, 70)), kge.SamplingParams())
    with pytest.raises(kge.KgeError, match="out of vocabulary"):
        llm.add_request([5, cfg.vocab_size + 3], kge.SamplingParams())
    llm.close()
    bad = dict(cfg.to_hf(), model_type="mamba")
    (tmp_path / "config.json").write_text(json.dumps(bad))
    with pytest.raises(kge.KgeError, match="unsupported model_type"):
        kge.LLM(str(tmp_path), load_format="dummy")
    # a KV pool that cannot hold one request of max_model_len tokens: refused at start-up, as vLLM does
    with pytest.raises(kge.KgeError, match="larger than the maximum number of tokens that can be stored in KV cache"):
        kge.LLM(d, max_model_len=128, max_num_seqs=4, max_num_batched_tokens=64, num_gpu_blocks_override=5)
    # a checkpoint whose tensors do not have the shapes config.json implies must not be copied into buffers sized from
    # the config: larger vocabulary rows, a narrower MLP, a bias of another length
    import shutil
    for name, shape in (("model.embed_tokens.weight", (cfg.vocab_size + 64, cfg.hidden_size)),
                        ("model.layers.1.mlp.down_proj.weight", (cfg.hidden_size, cfg.ffn // 2)),
                        ("model.layers.0.input_layernorm.weight", (cfg.hidden_size * 2,))):
        w2 = dict(w)
        w2[name] = torch.zeros(*shape, dtype=BF)
        dd = tmp_path / ("bad-" + name.replace(".", "_"))
        om.save_model_dir(str(dd), cfg, w2)
        with pytest.raises(kge.KgeError, match="does not match config.json"):
            kge.LLM(str(dd), max_model_len=64, max_num_seqs=4, max_num_batched_tokens=64, num_gpu_blocks_override=8)
        shutil.rmtree(dd)
