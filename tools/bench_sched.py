"""Host cost of one scheduling step of the C++ scheduler (no GPU): kge_sched_schedule + kge_sched_update with the C3
shape (64 running requests, 2048-token prompts, block 16), with and without prefix caching.  Prints JSON lines."""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kge_loader import load  # noqa: E402

kge = load()
L = kge.lib()


def run(prefix_caching, B=64, pin=2048, pout=256):
    s = L.kge_sched_create(B, 8192, pin + pout + 16, 16, B * ((pin + pout + 15) // 16) + 2, 1, -1)
    L.kge_sched_set_prefix_caching(s, int(prefix_caching))
    sp = kge.SamplingParams(max_tokens=pout, temperature=0.0, ignore_eos=True).to_c()
    for i in range(B):
        arr = (C.c_int32 * pin)(*[(i * 7919 + t * 31) % 50000 + 10 for t in range(pin)])
        assert L.kge_sched_add(s, i, arr, pin, C.byref(sp)) == 0
    ids, nt, outs = (C.c_int64 * 4096)(), (C.c_int32 * 4096)(), (kge.StepOutput * 4096)()
    toks = (C.c_int32 * 4096)(*([7] * 4096))
    steps, t_sched, t_upd = 0, 0.0, 0.0
    decode_steps, t_decode = 0, 0.0
    run_, wait_, free_ = C.c_int32(), C.c_int32(), C.c_int32()
    while True:
        L.kge_sched_counts(s, C.byref(run_), C.byref(wait_), C.byref(free_))
        if run_.value == 0 and wait_.value == 0:
            break
        t0 = time.perf_counter()
        n = L.kge_sched_schedule(s, ids, nt, 4096)
        t1 = time.perf_counter()
        m = sum(1 for j in range(n) if nt[j] == 1) if n < 4096 else 0
        L.kge_sched_update(s, None, toks, n, outs, 4096)
        t2 = time.perf_counter()
        steps += 1
        t_sched += t1 - t0
        t_upd += t2 - t1
        if m == n == B:
            decode_steps += 1
            t_decode += (t1 - t0) + (t2 - t1)
    L.kge_sched_destroy(s)
    return {"prefix_caching": bool(prefix_caching), "steps": steps, "schedule_us_per_step": round(1e6 * t_sched / steps, 2),
            "update_us_per_step": round(1e6 * t_upd / steps, 2),
            "full_decode_step_us": round(1e6 * t_decode / max(1, decode_steps), 2), "full_decode_steps": decode_steps}


if __name__ == "__main__":
    for pc in (False, True):
        print(json.dumps(run(pc)), flush=True)
