"""Automatic prefix caching: the Python restatement (oracle/prefix_cache.py) and the product's C++ scheduler behind the C
ABI replayed against traces of vLLM's REAL v1 Scheduler with enable_prefix_caching=True
(tests/golden/sched_prefix_vllm.json, made by tests/golden/make_sched_prefix_golden.py from the installed vLLM 0.22.0):
shared prefixes, duplicates, whole-block prompts, late arrivals, eviction of cached blocks, preemption.  Per step: same
requests, same token counts, same order; every admission takes the same number of tokens from the cache; the same
requests finish for the same reason; the pool has the same number of free blocks.  CPU only."""
import ctypes as C
import json
import os

import pytest

from oracle.prefix_cache import PrefixCachingScheduler
from oracle.scheduler import Req, SchedConfig

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sched_prefix_vllm.json")
CASES = json.load(open(GOLD))
IDS = lambda c: f"seed{c['seed']}"  # noqa: E731


def test_fixture_exercises_the_cache():
    hits = [n for c in CASES for st in c["steps"] for _, n in st["cached"] if n > 0]
    assert len(hits) > 100 and max(hits) >= 64
    assert sum(1 for c in CASES for st in c["steps"] if st["preempted"]) >= 10
    assert any(c["config"]["reserve_full_isl"] for c in CASES) and any(not c["config"]["reserve_full_isl"] for c in CASES)
    # pools small enough that cached blocks must be evicted: some step ends with (almost) nothing free
    assert sum(1 for c in CASES if min(st["free_blocks"] for st in c["steps"]) <= 2) >= 3


@pytest.mark.parametrize("case", CASES, ids=IDS)
def test_oracle_prefix_caching_reproduces_vllm(case):
    k = case["config"]
    o = PrefixCachingScheduler(SchedConfig(max_num_seqs=k["max_num_seqs"], max_num_batched_tokens=k["max_num_batched_tokens"],
                                           max_model_len=k["max_model_len"], block_size=k["block_size"], num_blocks=k["num_blocks"],
                                           enable_chunked_prefill=True, reserve_full_isl=k["reserve_full_isl"]))
    reqs = {r["rid"]: r for r in case["requests"]}
    for i, st in enumerate(case["steps"]):
        for rid in st["arrivals"]:
            r = reqs[rid]
            o.add(Req(r["rid"], list(r["prompt"]), r["max_tokens"], r["ignore_eos"], case["eos"], tuple(r["stop"]), r["min_tokens"]))
        sched = o.schedule()
        assert [[a, b] for a, b in sched.items()] == st["sched"], f"step {i}"
        assert [[rid, o.cached_tokens[rid]] for rid, _ in st["cached"]] == st["cached"], f"step {i}"
        done = o.update(sched, {a: b for a, b in st["tokens"]})
        assert {r.rid: ("stop" if r.status == "stop" else "length") for r in done} == {a: b for a, b in st["finished"]}, f"step {i}"
        assert o.free_blocks == st["free_blocks"], f"step {i}"
    assert not o.running and not o.waiting


@pytest.mark.parametrize("case", CASES, ids=IDS)
def test_cpp_prefix_caching_reproduces_vllm(kge, case):
    L = kge.lib()
    k = case["config"]
    s = L.kge_sched_create(k["max_num_seqs"], k["max_num_batched_tokens"], k["max_model_len"], k["block_size"], k["num_blocks"], 1,
                           case["eos"])
    L.kge_sched_set_reserve_full_isl(s, int(k["reserve_full_isl"]))
    L.kge_sched_set_prefix_caching(s, 1)
    reqs = {r["rid"]: r for r in case["requests"]}
    ids = (C.c_int64 * 4096)()
    nt = (C.c_int32 * 4096)()
    outs = (kge.StepOutput * 4096)()
    run, wait, free = C.c_int32(), C.c_int32(), C.c_int32()
    for i, st in enumerate(case["steps"]):
        for rid in st["arrivals"]:
            r = reqs[rid]
            sp = kge.SamplingParams(max_tokens=r["max_tokens"], min_tokens=r["min_tokens"], temperature=0.0, ignore_eos=r["ignore_eos"],
                                    stop_token_ids=list(r["stop"])).to_c()
            arr = (C.c_int32 * len(r["prompt"]))(*r["prompt"])
            assert L.kge_sched_add(s, r["rid"], arr, len(r["prompt"]), C.byref(sp)) == 0
        n = L.kge_sched_schedule(s, ids, nt, 4096)
        assert [[ids[j], nt[j]] for j in range(n)] == st["sched"], f"step {i}"
        assert [[rid, L.kge_sched_cached_tokens(s, rid)] for rid, _ in st["cached"]] == st["cached"], f"step {i}"
        toks = [v for _, v in st["tokens"]]
        tarr = (C.c_int32 * max(1, len(toks)))(*toks)
        m = L.kge_sched_update(s, None, tarr, len(toks), outs, 4096)
        got = {outs[j].request_id: ("stop" if outs[j].finished == 1 else "length") for j in range(m) if outs[j].finished}
        assert got == {a: b for a, b in st["finished"]}, f"step {i}"
        L.kge_sched_counts(s, C.byref(run), C.byref(wait), C.byref(free))
        assert free.value == st["free_blocks"], f"step {i}"
    assert (run.value, wait.value) == (0, 0) and free.value == k["num_blocks"] - 1
    L.kge_sched_destroy(s)
