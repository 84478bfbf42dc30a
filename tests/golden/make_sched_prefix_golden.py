"""Generate tests/golden/sched_prefix_vllm.json: per-step schedules of vLLM's REAL v1 Scheduler with automatic prefix
caching ON (its default), on workloads built to hit the cache: shared prefixes of several lengths, exact duplicates,
prompts that are whole multiples of the block size, requests arriving while others run, pools small enough to evict
cached blocks and to preempt.  Same conventions as make_sched_golden.py; each step also records how many tokens every
newly admitted request took from the cache.  Runs only in the build container.

  python tests/golden/make_sched_prefix_golden.py
"""
import json
import os
import random
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import model as om  # noqa: E402


def make_scheduler(cfg_dir, conf):
    from vllm.config import CacheConfig, DeviceConfig, ModelConfig, SchedulerConfig, VllmConfig
    from vllm.v1.core.sched.scheduler import Scheduler
    from vllm.v1.kv_cache_interface import FullAttentionSpec, KVCacheConfig, KVCacheGroupSpec
    from vllm.v1.structured_output import StructuredOutputManager
    mc = ModelConfig(model=cfg_dir, tokenizer=cfg_dir, skip_tokenizer_init=True, dtype="bfloat16", max_model_len=conf["max_model_len"])
    sc = SchedulerConfig(max_model_len=conf["max_model_len"], is_encoder_decoder=False, max_num_batched_tokens=conf["max_num_batched_tokens"],
                         max_num_seqs=conf["max_num_seqs"], enable_chunked_prefill=True,
                         scheduler_reserve_full_isl=conf["reserve_full_isl"], async_scheduling=False)
    cc = CacheConfig(block_size=conf["block_size"], gpu_memory_utilization=0.9, enable_prefix_caching=True)
    cc.num_gpu_blocks = conf["num_blocks"]
    vc = VllmConfig(model_config=mc, cache_config=cc, scheduler_config=sc, device_config=DeviceConfig(device="cpu"))
    spec = FullAttentionSpec(conf["block_size"], num_kv_heads=2, head_size=64, dtype=torch.bfloat16)
    kvc = KVCacheConfig(num_blocks=conf["num_blocks"], kv_cache_tensors=[], kv_cache_groups=[KVCacheGroupSpec(["layer0"], spec)])
    return Scheduler(vc, kvc, StructuredOutputManager(vc), block_size=conf["block_size"])


def run_case(cfg_dir, seed, eos=2):
    from vllm.sampling_params import SamplingParams
    from vllm.utils.hashing import sha256
    from vllm.v1.core.kv_cache_utils import get_request_block_hasher, init_none_hash
    from vllm.v1.outputs import ModelRunnerOutput
    from vllm.v1.request import Request
    rng = random.Random(1000 + seed)
    bs = rng.choice([4, 16])
    conf = dict(max_num_seqs=rng.choice([3, 8, 16]), max_num_batched_tokens=rng.choice([24, 64, 256]), max_model_len=200,
                block_size=bs, num_blocks=rng.choice([30, 48, 80, 400]) if bs == 4 else rng.choice([12, 20, 40, 200]),
                reserve_full_isl=bool(seed % 2))
    init_none_hash(sha256)
    hasher = get_request_block_hasher(bs, sha256)
    s = make_scheduler(cfg_dir, conf)
    cap = (conf["num_blocks"] - 1) * bs
    prefixes = [[rng.randrange(20, 900) for _ in range(rng.choice([bs, 2 * bs, 3 * bs + 1, 5 * bs, 37]))] for _ in range(3)]
    reqs = []
    for i in range(18):
        kind = rng.random()
        if kind < 0.2 and reqs:
            prompt = list(rng.choice(reqs)["prompt"])                               # exact duplicate
        elif kind < 0.75:
            prompt = list(rng.choice(prefixes)) + [rng.randrange(20, 900) for _ in range(rng.randrange(0, 30))]
        else:
            prompt = [rng.randrange(20, 900) for _ in range(rng.randrange(1, 60))]
        if rng.random() < 0.25:
            prompt = prompt[:max(bs, len(prompt) // bs * bs)]                        # whole blocks: the last one is recomputed
        prompt = prompt[:max(1, min(len(prompt), cap // 2, conf["max_model_len"] - 50))]
        reqs.append(dict(rid=i, prompt=prompt, max_tokens=rng.randrange(1, 30), ignore_eos=rng.random() < 0.6,
                         stop=[13] if rng.random() < 0.2 else [], min_tokens=0, arrive=0 if i < 6 else rng.randrange(1, 60)))
    pending = sorted(reqs, key=lambda r: (r["arrive"], r["rid"]))
    steps = []
    tok_rng = random.Random(seed + 5000)
    step_no = 0
    while pending or s.has_unfinished_requests():
        arrivals = []
        while pending and (pending[0]["arrive"] <= step_no or not s.has_unfinished_requests()):
            r = pending.pop(0)
            sp = SamplingParams(max_tokens=r["max_tokens"], min_tokens=r["min_tokens"], temperature=0.0, ignore_eos=r["ignore_eos"],
                                stop_token_ids=list(r["stop"]))
            sp.update_from_generation_config({}, eos)
            s.add_request(Request(str(r["rid"]), list(r["prompt"]), sp, None, arrival_time=float(step_no), block_hasher=hasher))
            arrivals.append(r["rid"])
        out = s.schedule()
        sched = [[int(k), int(v)] for k, v in out.num_scheduled_tokens.items()]
        assert sched, "no progress"
        req_ids = list(out.num_scheduled_tokens)
        cached = [[int(n.req_id), int(n.num_computed_tokens)] for n in out.scheduled_new_reqs]
        cr = out.scheduled_cached_reqs
        for rid, resumed_ids, nct in zip(cr.req_ids, [rid in cr.resumed_req_ids for rid in cr.req_ids], cr.num_computed_tokens):
            if resumed_ids:
                cached.append([int(rid), int(nct)])
        sampled, toks = [], {}
        for rid in req_ids:
            req = s.requests[rid]
            if req.num_computed_tokens == req.num_tokens:
                t = tok_rng.choice([eos, 7, 11, 13, 17, 19])
                sampled.append([t])
                toks[int(rid)] = t
            else:
                sampled.append([])
        mro = ModelRunnerOutput(req_ids=req_ids, req_id_to_index={r: i for i, r in enumerate(req_ids)}, sampled_token_ids=sampled)
        eco = s.update_from_output(out, mro)
        fin = {}
        for outs in eco.values():
            for o in outs.outputs:
                if o.finish_reason is not None:
                    fin[int(o.request_id)] = str(o.finish_reason)
        steps.append(dict(arrivals=arrivals, sched=sched, cached=cached, tokens=[[k, v] for k, v in toks.items()],
                          finished=[[k, v] for k, v in fin.items()], preempted=sorted(int(x) for x in (out.preempted_req_ids or [])),
                          free_blocks=int(s.kv_cache_manager.block_pool.get_num_free_blocks())))
        step_no += 1
        assert step_no < 5000
    return dict(seed=seed, eos=eos, config=conf, requests=reqs, steps=steps)


def main():
    cfg_dir = "/tmp/kge_sched_cfg"
    os.makedirs(cfg_dir, exist_ok=True)
    json.dump(om.CONFIGS["tiny-llama"].to_hf(), open(os.path.join(cfg_dir, "config.json"), "w"))
    cases = [run_case(cfg_dir, seed) for seed in range(16)]
    with open(os.path.join(HERE, "sched_prefix_vllm.json"), "w") as f:
        json.dump(cases, f, separators=(",", ":"))
    hits = sum(1 for c in cases for st in c["steps"] for _, n in st["cached"] if n > 0)
    print("sched_prefix_vllm.json:", len(cases), "cases,", sum(len(c["steps"]) for c in cases), "steps,", hits, "admissions with cache hits,",
          sum(1 for c in cases for st in c["steps"] if st["preempted"]), "steps with preemption")


if __name__ == "__main__":
    main()
