"""Scheduler oracle with automatic prefix caching.  TEST INFRASTRUCTURE (see oracle/__init__.py).

Adds to oracle/scheduler.py what vLLM does when `--enable-prefix-caching` is on (its V1 default):
  block pool        vllm:v1/core/block_pool.py:184-209 (get_cached_block: first block cached under a hash),
                    :211-290 (cache_full_blocks), :333-363 (get_new_blocks pops the FRONT of the free queue and evicts
                    the block's hash), :402-417 (touch), :419-433 (free_blocks: decrement all, then append the ones
                    that reached zero, in the given order)
  per-request view  vllm:v1/core/single_type_kv_cache_manager.py:88-181 (blocks to allocate = missing blocks + hit blocks
                    that are eviction candidates), :183-240 (hit blocks are touched, then become the request's first
                    blocks), :277-323 (cache the blocks that became full), :339-353 (free in REVERSE order, so a
                    request's tail is evicted before its head), :483-528 (longest cache hit = leading run of cached
                    full blocks)
  cache lookup      vllm:v1/core/kv_cache_manager.py:194-235 (at most num_tokens - 1 tokens may hit: the last token is
                    always recomputed for its logits), :346-360 (full-sequence admission gate counts hits)
  scheduler         vllm:v1/core/sched/scheduler.py:590-596 (lookup only for requests with nothing computed: new or
                    preempted), :807-812 (num_computed_tokens starts at the hit length)
Block hashes are a chain over (parent, the block's tokens); here they are interned tuples, so equal hash <=> equal
prefix exactly.  Block ids matter now (which cached block is evicted decides later hits), so the pool hands out real ids
in vLLM's order.  Checked step by step against the real Scheduler (tests/golden/sched_prefix_vllm.json).
"""
from __future__ import annotations

from collections import OrderedDict

from .scheduler import OracleScheduler, Req, SchedConfig


class BlockPool:
    def __init__(self, num_blocks: int):
        self.free: OrderedDict[int, None] = OrderedDict((b, None) for b in range(1, num_blocks))   # block 0 = null block
        self.ref = [0] * num_blocks
        self.hash_of: list = [None] * num_blocks
        self.cached: dict = {}                       # hash -> OrderedDict{block id: None}, insertion order

    def num_free(self) -> int:
        return len(self.free)

    def get_cached(self, h):
        d = self.cached.get(h)
        return next(iter(d)) if d else None

    def touch(self, blocks) -> None:
        for b in blocks:
            if self.ref[b] == 0:
                del self.free[b]
            self.ref[b] += 1

    def get_new(self, n: int) -> list[int]:
        out = []
        for _ in range(n):
            b, _ = self.free.popitem(last=False)
            h = self.hash_of[b]
            if h is not None:                        # evict: the block stops being a cache entry
                d = self.cached.get(h)
                if d is not None and b in d:
                    del d[b]
                    if not d:
                        del self.cached[h]
                self.hash_of[b] = None
            self.ref[b] = 1
            out.append(b)
        return out

    def free_blocks(self, ordered) -> None:
        ordered = list(ordered)
        for b in ordered:
            self.ref[b] -= 1
        for b in ordered:
            if self.ref[b] == 0:
                self.free[b] = None

    def cache(self, b: int, h) -> None:
        self.hash_of[b] = h
        self.cached.setdefault(h, OrderedDict())[b] = None


class PrefixCachingScheduler(OracleScheduler):
    def __init__(self, cfg: SchedConfig):
        super().__init__(cfg)
        self.pool = BlockPool(cfg.num_blocks)
        self.blocks: dict[int, list[int]] = {}       # rid -> block ids
        self.num_cached: dict[int, int] = {}         # rid -> leading blocks that carry a hash
        self._intern: dict = {}
        self.cached_tokens: dict[int, int] = {}      # rid -> tokens served from the cache at its last admission

    # -- hashes
    def _hashes(self, r: Req, nblocks: int) -> list:
        bs = self.cfg.block_size
        toks = r.prompt + r.output
        out, parent = [], 0
        for i in range(nblocks):
            key = (parent, tuple(toks[i * bs:(i + 1) * bs]))
            parent = self._intern.setdefault(key, len(self._intern) + 1)
            out.append(parent)
        return out

    def _cache_full_blocks(self, r: Req, num_tokens: int) -> None:
        nfull = num_tokens // self.cfg.block_size
        done = self.num_cached.get(r.rid, 0)
        if done >= nfull:
            return
        hs = self._hashes(r, nfull)
        for i in range(done, nfull):
            self.pool.cache(self.blocks[r.rid][i], hs[i])
        self.num_cached[r.rid] = nfull

    # -- hooks of OracleScheduler
    def _alloc(self, r: Req, n_new: int) -> bool:          # RUNNING request grows
        bs = self.cfg.block_size
        have = self.blocks.setdefault(r.rid, [])
        need = max(-(-(r.num_computed + n_new) // bs) - len(have), 0)
        if need > self.pool.num_free():
            return False
        have.extend(self.pool.get_new(need))
        self._cache_full_blocks(r, min(r.num_computed + n_new, r.num_tokens))
        return True

    def _free(self, r: Req) -> None:
        self.pool.free_blocks(reversed(self.blocks.pop(r.rid, [])))
        self.num_cached.pop(r.rid, None)

    def _admit(self, r: Req, budget: int):
        c, bs = self.cfg, self.cfg.block_size
        hits = []
        if r.num_computed == 0:
            for h in self._hashes(r, (r.num_tokens - 1) // bs):
                b = self.pool.get_cached(h)
                if b is None:
                    break
                hits.append(b)
        nc = len(hits) * bs
        n = r.num_tokens - nc
        if not c.enable_chunked_prefill and n > budget:
            return None
        n = min(n, budget)
        evictable = sum(self.pool.ref[b] == 0 for b in hits)
        if c.reserve_full_isl:
            full = -(-min(r.num_tokens, c.max_model_len) // bs)
            if max(full - len(hits), 0) + evictable > self.pool.num_free():
                return None
        need = max(-(-(nc + n) // bs) - len(hits), 0)
        if need + evictable > self.pool.num_free():
            return None
        self.pool.touch(hits)
        self.blocks[r.rid] = list(hits)
        self.num_cached[r.rid] = len(hits)
        self.blocks[r.rid].extend(self.pool.get_new(need))
        r.num_computed = nc
        self.cached_tokens[r.rid] = nc
        self._cache_full_blocks(r, min(nc + n, r.num_tokens))
        return n

    @property
    def free_blocks(self) -> int:
        return self.pool.num_free()

    @free_blocks.setter
    def free_blocks(self, v) -> None:       # the base class initialises a counter; the pool is the truth here
        pass
