"""Scheduler / block-manager oracle.  TEST INFRASTRUCTURE (see oracle/__init__.py).

Restates the FCFS policy of vllm:v1/core/sched/scheduler.py:329-830 (RUNNING first under a
token budget, tail preemption with recompute when blocks run out, then WAITING admission up
to max_num_seqs), :1283 update_from_output, the stop rules of
vllm:v1/core/sched/utils.py:94-130, and the block accounting of
vllm:v1/core/kv_cache_manager.py:236-430 / block_pool.py (block 0 is the reserved null
block, so num_blocks-1 are usable).  No prefix caching, no speculative tokens.
Only WHO runs HOW MANY tokens each step is compared with the product; block ids are free.
Checked step by step against traces of the real vLLM Scheduler (tests/golden/sched_vllm.json,
tests/test_scheduler_vllm_golden.py).
"""
from __future__ import annotations

from collections import deque
from dataclasses import dataclass, field


@dataclass
class SchedConfig:
    max_num_seqs: int = 1024            # vllm:engine/arg_utils.py:2311-2331 (>=70 GiB GPUs)
    max_num_batched_tokens: int = 8192
    max_model_len: int = 4096
    block_size: int = 16
    num_blocks: int = 1024              # including the null block 0
    enable_chunked_prefill: bool = True
    # --scheduler-reserve-full-isl (vLLM default on): admit a waiting request only if free blocks
    # cover its whole known sequence (kv_cache_manager.py:346-360, full_sequence_must_fit)
    reserve_full_isl: bool = True


@dataclass
class Req:
    rid: int
    prompt: list
    max_tokens: int
    ignore_eos: bool = False
    eos: int = -1
    stop_ids: tuple = ()
    min_tokens: int = 0
    output: list = field(default_factory=list)
    num_computed: int = 0
    num_blocks: int = 0
    status: str = "waiting"            # waiting|running|preempted|stop|length
    num_preemptions: int = 0

    @property
    def num_tokens(self) -> int:
        return len(self.prompt) + len(self.output)


class OracleScheduler:
    def __init__(self, cfg: SchedConfig):
        self.cfg = cfg
        self.free_blocks = cfg.num_blocks - 1
        self.waiting: deque[Req] = deque()
        self.running: list[Req] = []
        self.finished: list[Req] = []

    def add(self, r: Req) -> None:
        self.waiting.append(r)

    def _alloc(self, r: Req, n_new: int) -> bool:
        bs = self.cfg.block_size
        need = -(-(r.num_computed + n_new) // bs) - r.num_blocks
        if need > self.free_blocks:
            return False
        self.free_blocks -= max(need, 0)
        r.num_blocks += max(need, 0)
        return True

    def _free(self, r: Req) -> None:
        self.free_blocks += r.num_blocks
        r.num_blocks = 0

    def _admit(self, r: Req, budget: int):
        """WAITING -> RUNNING (scheduler.py:560-760).  Returns the tokens scheduled this step, or None
        when the request (and so everything behind it) has to wait."""
        c = self.cfg
        n = r.num_tokens - r.num_computed
        if not c.enable_chunked_prefill and n > budget:
            return None
        n = min(n, budget)
        if c.reserve_full_isl:
            full = min(r.num_tokens, c.max_model_len)
            if -(-full // c.block_size) - r.num_blocks > self.free_blocks:
                return None
        if not self._alloc(r, n):
            return None
        return n

    def schedule(self) -> dict[int, int]:
        """Returns {rid: num_scheduled_tokens} in scheduling order and advances
        num_computed (scheduler.py _update_after_schedule)."""
        c = self.cfg
        budget = c.max_num_batched_tokens
        sched: dict[int, int] = {}
        preempted = False
        i = 0
        while i < len(self.running) and budget > 0:
            r = self.running[i]
            n = min(r.num_tokens - r.num_computed, budget, c.max_model_len - 1 - r.num_computed)
            if n <= 0:
                i += 1
                continue
            ok = True
            while not self._alloc(r, n):
                victim = self.running.pop()
                self._free(victim)
                victim.num_computed = 0
                victim.status = "preempted"
                victim.num_preemptions += 1
                self.waiting.appendleft(victim)
                preempted = True
                if victim.rid in sched:     # cannot happen with tail preemption; kept for clarity
                    budget += sched.pop(victim.rid)
                if victim is r:
                    ok = False
                    break
            if not ok:
                break
            sched[r.rid] = n
            budget -= n
            i += 1
        if not preempted:
            while self.waiting and budget > 0:
                if len(self.running) == c.max_num_seqs:
                    break
                r = self.waiting[0]
                n = self._admit(r, budget)
                if n is None:
                    break
                self.waiting.popleft()
                r.status = "running"
                self.running.append(r)
                sched[r.rid] = n
                budget -= n
        by_id = {r.rid: r for r in self.running}
        for rid, n in sched.items():
            by_id[rid].num_computed += n
        return sched

    def update(self, sched: dict[int, int], sampled: dict[int, int]) -> list[Req]:
        """sampled[rid] = token sampled for requests whose scheduled chunk reached their last
        known token.  Returns requests finished this step."""
        done = []
        for r in list(self.running):
            if r.rid not in sched or r.num_computed < r.num_tokens:
                continue
            tok = sampled[r.rid]
            r.output.append(tok)
            fin = None
            if len(r.output) >= r.min_tokens:      # utils.py:100-101: nothing stops earlier
                if not r.ignore_eos and tok == r.eos:
                    fin = "stop"
                elif tok in r.stop_ids:
                    fin = "stop"
                elif (r.num_tokens >= self.cfg.max_model_len
                      or len(r.output) >= r.max_tokens):
                    fin = "length"
            if fin:
                r.status = fin
                self._free(r)
                self.running.remove(r)
                self.finished.append(r)
                done.append(r)
        return done
