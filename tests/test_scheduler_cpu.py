"""C++ scheduler / block manager (through the C ABI) against the Python restatement of vLLM's
policy (oracle/scheduler.py) on randomised workloads: identical per-step schedules, sampled-token
bookkeeping, preemptions and finish reasons.  CPU only."""
import ctypes as C
import random

import pytest

from oracle.scheduler import OracleScheduler, Req, SchedConfig


def run_pair(kge, cfg: SchedConfig, reqs, eos, seed):
    L = kge.lib()
    s = L.kge_sched_create(cfg.max_num_seqs, cfg.max_num_batched_tokens, cfg.max_model_len, cfg.block_size,
                           cfg.num_blocks, int(cfg.enable_chunked_prefill), eos)
    o = OracleScheduler(cfg)
    for r in reqs:
        o.add(Req(r["rid"], list(r["prompt"]), r["max_tokens"], r["ignore_eos"], eos, tuple(r["stop"]), r["min_tokens"]))
        sp = kge.SamplingParams(max_tokens=r["max_tokens"], min_tokens=r["min_tokens"], temperature=0.0,
                                ignore_eos=r["ignore_eos"], stop_token_ids=list(r["stop"])).to_c()
        arr = (C.c_int32 * len(r["prompt"]))(*r["prompt"])
        assert L.kge_sched_add(s, r["rid"], arr, len(r["prompt"]), C.byref(sp)) == 0
    rng = random.Random(seed)
    ids = (C.c_int64 * 4096)()
    nt = (C.c_int32 * 4096)()
    outs = (kge.StepOutput * 4096)()
    steps = 0
    finished = {}
    while o.running or o.waiting:
        n = L.kge_sched_schedule(s, ids, nt, 4096)
        got = [(ids[i], nt[i]) for i in range(n)]
        exp_d = o.schedule()
        assert got == list(exp_d.items()), f"step {steps}"
        assert got, "no progress"
        by = {r.rid: r for r in o.running}
        need = [rid for rid, _ in got if by[rid].num_computed == by[rid].num_tokens]
        toks = {rid: rng.choice([eos, 7, 11, 13, 17, 19]) for rid in need}
        done = o.update(exp_d, toks)
        tarr = (C.c_int32 * max(1, len(need)))(*[toks[r] for r in need])
        m = L.kge_sched_update(s, None, tarr, len(need), outs, 4096)
        assert m == n
        got_fin = {outs[i].request_id: outs[i].finished for i in range(m) if outs[i].finished}
        exp_fin = {r.rid: (1 if r.status == "stop" else 2) for r in done}
        assert got_fin == exp_fin, f"step {steps}"
        for i in range(m):
            exp_tok = toks.get(outs[i].request_id, -1)
            assert outs[i].token == exp_tok
        finished.update(got_fin)
        run, wait, free = C.c_int32(), C.c_int32(), C.c_int32()
        L.kge_sched_counts(s, C.byref(run), C.byref(wait), C.byref(free))
        assert (run.value, wait.value, free.value) == (len(o.running), len(o.waiting), o.free_blocks)
        steps += 1
        assert steps < 20000
    L.kge_sched_destroy(s)
    return steps, finished, sum(r.num_preemptions for r in o.finished)


def make_reqs(rng, n, max_prompt, max_out):
    return [dict(rid=i, prompt=[rng.randrange(20, 900) for _ in range(rng.randrange(1, max_prompt))],
                 max_tokens=rng.randrange(1, max_out), ignore_eos=rng.random() < 0.5,
                 stop=[13] if rng.random() < 0.3 else [], min_tokens=rng.choice([0, 0, 3]))
            for i in range(n)]


@pytest.mark.parametrize("seed", range(6))
def test_random_workloads_match_oracle(kge, seed):
    rng = random.Random(seed)
    cfg = SchedConfig(max_num_seqs=rng.choice([2, 4, 16]), max_num_batched_tokens=rng.choice([16, 48, 256]),
                      max_model_len=160, block_size=rng.choice([4, 16]), num_blocks=rng.choice([24, 40, 400]),
                      enable_chunked_prefill=True)
    reqs = make_reqs(rng, 30, min(80, (cfg.num_blocks - 1) * cfg.block_size // 2), 40)
    steps, fin, _ = run_pair(kge, cfg, reqs, eos=2, seed=seed)
    assert len(fin) == 30 and steps > 0


def test_preemption_path_is_exercised(kge):
    """A pool too small for all running requests forces tail preemption + recompute."""
    cfg = SchedConfig(max_num_seqs=8, max_num_batched_tokens=64, max_model_len=200, block_size=4, num_blocks=30)
    reqs = [dict(rid=i, prompt=list(range(30, 50)), max_tokens=40, ignore_eos=True, stop=[], min_tokens=0) for i in range(6)]
    steps, fin, preempt = run_pair(kge, cfg, reqs, eos=2, seed=0)
    assert len(fin) == 6 and all(v == 2 for v in fin.values())
    assert preempt > 0


def test_rejects_bad_requests(kge):
    L = kge.lib()
    s = L.kge_sched_create(4, 64, 32, 16, 8, 1, 2)
    sp = kge.SamplingParams(max_tokens=4).to_c()
    arr = (C.c_int32 * 40)(*range(40))
    assert L.kge_sched_add(s, 1, arr, 0, C.byref(sp)) == -1          # empty prompt
    assert L.kge_sched_add(s, 1, arr, 40, C.byref(sp)) == -1         # longer than max_model_len
    assert L.kge_sched_add(s, 1, arr, 8, C.byref(sp)) == 0
    assert L.kge_sched_add(s, 1, arr, 8, C.byref(sp)) == -1          # duplicate id
    assert b"duplicate" in L.kge_last_error()
    L.kge_sched_destroy(s)
