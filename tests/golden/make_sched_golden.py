"""Generate tests/golden/sched_vllm.json: per-step schedules of vLLM's REAL v1 Scheduler
(vllm/v1/core/sched/scheduler.py of the installed 0.22.0, constructed on the CPU with a hand-built
KVCacheConfig) on randomised workloads with scripted "sampled" tokens.  Runs only in the build
container.  The fixtures pin rows a4/a5 of SURVEY.md section 8: oracle/scheduler.py and the C++
scheduler must reproduce, step by step, who runs how many tokens, who is preempted and how every
request finishes.

  python tests/golden/make_sched_golden.py
"""
import json
import os
import random
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import model as om  # noqa: E402


def make_scheduler(cfg_dir, max_num_seqs, budget, max_model_len, block_size, num_blocks, reserve_full_isl):
    from vllm.config import CacheConfig, DeviceConfig, ModelConfig, SchedulerConfig, VllmConfig
    from vllm.v1.core.sched.scheduler import Scheduler
    from vllm.v1.kv_cache_interface import FullAttentionSpec, KVCacheConfig, KVCacheGroupSpec
    from vllm.v1.structured_output import StructuredOutputManager
    mc = ModelConfig(model=cfg_dir, tokenizer=cfg_dir, skip_tokenizer_init=True, dtype="bfloat16", max_model_len=max_model_len)
    sc = SchedulerConfig(max_model_len=max_model_len, is_encoder_decoder=False, max_num_batched_tokens=budget,
                         max_num_seqs=max_num_seqs, enable_chunked_prefill=True, scheduler_reserve_full_isl=reserve_full_isl,
                         async_scheduling=False)
    cc = CacheConfig(block_size=block_size, gpu_memory_utilization=0.9, enable_prefix_caching=False)
    cc.num_gpu_blocks = num_blocks
    vc = VllmConfig(model_config=mc, cache_config=cc, scheduler_config=sc, device_config=DeviceConfig(device="cpu"))
    spec = FullAttentionSpec(block_size, num_kv_heads=2, head_size=64, dtype=torch.bfloat16)
    kvc = KVCacheConfig(num_blocks=num_blocks, kv_cache_tensors=[], kv_cache_groups=[KVCacheGroupSpec(["layer0"], spec)])
    return Scheduler(vc, kvc, StructuredOutputManager(vc), block_size=block_size)


def run_case(cfg_dir, seed, reserve_full_isl, eos=2):
    from vllm.sampling_params import SamplingParams
    from vllm.v1.outputs import ModelRunnerOutput
    from vllm.v1.request import Request
    rng = random.Random(seed)
    conf = dict(max_num_seqs=rng.choice([2, 4, 16]), max_num_batched_tokens=rng.choice([16, 48, 256]), max_model_len=160,
                block_size=rng.choice([4, 16]), num_blocks=rng.choice([24, 40, 400]))
    if seed % 4 == 3:       # force the preemption path
        conf.update(max_num_seqs=8, max_num_batched_tokens=64, block_size=4, num_blocks=30)
    s = make_scheduler(cfg_dir, conf["max_num_seqs"], conf["max_num_batched_tokens"], conf["max_model_len"],
                       conf["block_size"], conf["num_blocks"], reserve_full_isl)
    conf["reserve_full_isl"] = reserve_full_isl
    cap = (conf["num_blocks"] - 1) * conf["block_size"]
    reqs = []
    for i in range(14):
        n = rng.randrange(1, min(80, cap // 2))
        r = dict(rid=i, prompt=[rng.randrange(20, 900) for _ in range(n)], max_tokens=rng.randrange(1, 40),
                 ignore_eos=rng.random() < 0.5, stop=[13] if rng.random() < 0.3 else [], min_tokens=rng.choice([0, 0, 3]))
        if seed % 4 == 3:
            r.update(prompt=list(range(30, 50)), max_tokens=40, ignore_eos=True, stop=[], min_tokens=0)
        reqs.append(r)
        sp = SamplingParams(max_tokens=r["max_tokens"], min_tokens=r["min_tokens"], temperature=0.0, ignore_eos=r["ignore_eos"],
                            stop_token_ids=list(r["stop"]))
        sp.update_from_generation_config({}, eos)       # sets eos_token_id the way the engine does
        s.add_request(Request(str(i), list(r["prompt"]), sp, None, arrival_time=float(i)))
    steps = []
    tok_rng = random.Random(seed + 1000)
    finished = {}
    while s.has_unfinished_requests():
        out = s.schedule()
        sched = [[int(k), int(v)] for k, v in out.num_scheduled_tokens.items()]
        assert sched, "no progress"
        req_ids = [k for k in out.num_scheduled_tokens]
        sampled, toks = [], {}
        for rid in req_ids:
            req = s.requests[rid]
            if req.num_computed_tokens == req.num_tokens:        # scheduler advanced it already: last known token reached
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
        finished.update(fin)
        steps.append(dict(sched=sched, tokens=[[k, v] for k, v in toks.items()], finished=[[k, v] for k, v in fin.items()],
                          preempted=sorted(int(x) for x in (out.preempted_req_ids or []))))
        assert len(steps) < 5000
    return dict(seed=seed, eos=eos, config=conf, requests=reqs, steps=steps)


def main():
    cfg_dir = "/tmp/kge_sched_cfg"
    os.makedirs(cfg_dir, exist_ok=True)
    json.dump(om.CONFIGS["tiny-llama"].to_hf(), open(os.path.join(cfg_dir, "config.json"), "w"))
    cases = [run_case(cfg_dir, seed, isl) for isl in (True, False) for seed in range(8)]
    with open(os.path.join(HERE, "sched_vllm.json"), "w") as f:
        json.dump(cases, f, separators=(",", ":"))
    print("sched_vllm.json:", len(cases), "cases,", sum(len(c["steps"]) for c in cases), "steps,",
          sum(1 for c in cases for st in c["steps"] if st["preempted"]), "steps with preemption")


if __name__ == "__main__":
    main()
