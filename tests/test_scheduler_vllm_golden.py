"""Both schedulers - the Python restatement (oracle/scheduler.py) and the product's C++ scheduler behind the C
ABI - replayed against traces of vLLM's REAL v1 Scheduler (tests/golden/sched_vllm.json, made by
tests/golden/make_sched_golden.py from the installed vLLM 0.22.0): per step, the same requests get the same
number of tokens in the same order, the same requests are preempted (implied by the schedules) and every request
finishes at the same step with the same reason - under both settings of --scheduler-reserve-full-isl (vLLM's
default is on).  CPU only."""
import ctypes as C
import json
import os

import pytest

from oracle.scheduler import OracleScheduler, Req, SchedConfig

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sched_vllm.json")
CASES = json.load(open(GOLD))


def test_fixture_covers_both_admission_modes_and_preemption():
    on = [c for c in CASES if c["config"]["reserve_full_isl"]]
    off = [c for c in CASES if not c["config"]["reserve_full_isl"]]
    assert on and off
    assert any(st["preempted"] for c in on for st in c["steps"]) and any(st["preempted"] for c in off for st in c["steps"])
    # the admission gate changes at least one trace, otherwise the switch is untested
    assert any(a["steps"] != b["steps"] for a, b in zip(on, off) if a["seed"] == b["seed"])


def sched_config(c):
    k = c["config"]
    return SchedConfig(max_num_seqs=k["max_num_seqs"], max_num_batched_tokens=k["max_num_batched_tokens"],
                       max_model_len=k["max_model_len"], block_size=k["block_size"], num_blocks=k["num_blocks"],
                       enable_chunked_prefill=True, reserve_full_isl=k["reserve_full_isl"])


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"seed{c['seed']}-{'isl' if c['config']['reserve_full_isl'] else 'noisl'}")
def test_oracle_scheduler_reproduces_vllm(case):
    o = OracleScheduler(sched_config(case))
    for r in case["requests"]:
        o.add(Req(r["rid"], list(r["prompt"]), r["max_tokens"], r["ignore_eos"], case["eos"], tuple(r["stop"]), r["min_tokens"]))
    for i, st in enumerate(case["steps"]):
        sched = o.schedule()
        assert [[k, v] for k, v in sched.items()] == st["sched"], f"step {i}"
        done = o.update(sched, {k: v for k, v in st["tokens"]})
        got = {r.rid: ("stop" if r.status == "stop" else "length") for r in done}
        assert got == {k: v for k, v in st["finished"]}, f"step {i}"
    assert not o.running and not o.waiting


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"seed{c['seed']}-{'isl' if c['config']['reserve_full_isl'] else 'noisl'}")
def test_cpp_scheduler_reproduces_vllm(kge, case):
    L = kge.lib()
    k = case["config"]
    s = L.kge_sched_create(k["max_num_seqs"], k["max_num_batched_tokens"], k["max_model_len"], k["block_size"], k["num_blocks"], 1,
                           case["eos"])
    L.kge_sched_set_reserve_full_isl(s, int(k["reserve_full_isl"]))
    for r in case["requests"]:
        sp = kge.SamplingParams(max_tokens=r["max_tokens"], min_tokens=r["min_tokens"], temperature=0.0, ignore_eos=r["ignore_eos"],
                                stop_token_ids=list(r["stop"])).to_c()
        arr = (C.c_int32 * len(r["prompt"]))(*r["prompt"])
        assert L.kge_sched_add(s, r["rid"], arr, len(r["prompt"]), C.byref(sp)) == 0
    ids = (C.c_int64 * 4096)()
    nt = (C.c_int32 * 4096)()
    outs = (kge.StepOutput * 4096)()
    for i, st in enumerate(case["steps"]):
        n = L.kge_sched_schedule(s, ids, nt, 4096)
        assert [[ids[j], nt[j]] for j in range(n)] == st["sched"], f"step {i}"
        toks = [v for _, v in st["tokens"]]          # recorded in schedule order
        tarr = (C.c_int32 * max(1, len(toks)))(*toks)
        m = L.kge_sched_update(s, None, tarr, len(toks), outs, 4096)
        got = {outs[j].request_id: ("stop" if outs[j].finished == 1 else "length") for j in range(m) if outs[j].finished}
        assert got == {k2: v for k2, v in st["finished"]}, f"step {i}"
    run, wait, free = C.c_int32(), C.c_int32(), C.c_int32()
    L.kge_sched_counts(s, C.byref(run), C.byref(wait), C.byref(free))
    assert (run.value, wait.value, free.value) == (0, 0, k["num_blocks"] - 1)
    L.kge_sched_destroy(s)
