#!/usr/bin/env python
"""Headline benchmark: output tokens/s at fixed concurrency (BASELINE.json `metric`).

  python bench.py --gpus N --steps K --warmup W [--workload c3|c2|c1] [--impl ours|reference]

A "step" is one whole job of the workload: every request prefilled (2k-token prompts, 8192 tokens
per engine step as the continuous-batching scheduler cuts them) and decoded greedily to its output length.
  value  device-resident replay of exactly the steps the scheduler made for that job: all step metadata is in HBM
         before the timed region, sampled ids feed the next step on the device, CUDA events on the engine's stream
         (kge_engine_record_steps / kge_engine_replay_recorded).  Same steps, so the same tokens as `e2e`.
  e2e    the same job through the public API (LLM.generate -> C++ continuous-batching scheduler):
         host token lists in, host token lists out; every engine step copies its metadata from pinned
         host memory and reads the sampled ids back, both inside the timed region.
At N > 1 every rank is an independent replica (request-level data parallelism, BASELINE config 5:
no data-path collective), timing is barrier-bracketed and the max over ranks; scaling is weak.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (oracle config key, concurrency B, prompt_len, out_len, requests prefilled per step)
    "c3": ("llama-3.1-8b", 64, 2048, 256, 4),
    "c2": ("llama-3.2-1b", 32, 128, 128, 32),
    "c1": ("opt-125m", 1, 128, 128, 1),
    # tensor parallel over all --gpus ranks (BASELINE config 4): one model, NCCL all-reduce, strong scaling
    "c4": ("llama-3.1-70b", 64, 2048, 256, 4),
}
CPU_SAMPLE = {"c3": (4, 128, 8), "c2": (8, 128, 16), "c1": (1, 128, 128), "c4": (1, 64, 2)}   # (requests, in, out)


def model_cfg(name):
    """HF config.json of the named model (published hyper-parameters)."""
    l3 = lambda f: {"factor": f, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                    "original_max_position_embeddings": 8192, "rope_type": "llama3"}
    if name == "opt-125m":
        return {"model_type": "opt", "vocab_size": 50272, "hidden_size": 768, "num_hidden_layers": 12, "ffn_dim": 3072,
                "num_attention_heads": 12, "max_position_embeddings": 2048, "do_layer_norm_before": True,
                "word_embed_proj_dim": 768, "activation_function": "relu", "bos_token_id": 2, "eos_token_id": 2}
    base = {"model_type": "llama", "vocab_size": 128256, "rms_norm_eps": 1e-5, "rope_theta": 500000.0,
            "max_position_embeddings": 131072, "bos_token_id": 128000, "eos_token_id": 128001}
    if name == "llama-3.2-1b":
        return dict(base, hidden_size=2048, num_hidden_layers=16, intermediate_size=8192, num_attention_heads=32,
                    num_key_value_heads=8, head_dim=64, tie_word_embeddings=True, rope_scaling=l3(32.0))
    if name == "llama-3.1-8b":
        return dict(base, hidden_size=4096, num_hidden_layers=32, intermediate_size=14336, num_attention_heads=32,
                    num_key_value_heads=8, head_dim=128, tie_word_embeddings=False, rope_scaling=l3(8.0))
    if name == "llama-3.1-70b":
        return dict(base, hidden_size=8192, num_hidden_layers=80, intermediate_size=28672, num_attention_heads=64,
                    num_key_value_heads=8, head_dim=128, tie_word_embeddings=False, rope_scaling=l3(8.0))
    raise ValueError(name)


def make_prompts(vocab, n, length, seed=0):
    import torch
    g = torch.Generator().manual_seed(seed)
    return torch.randint(1000, vocab - 1000, (n, length), generator=g).tolist()


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def max_over_ranks(x, world, device="cuda"):
    """Timing of a multi-rank run = the slowest rank's (all_reduce MAX); identity at world 1."""
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def job_throughput(replicas, B, pout, seconds):
    """Whole-job output tokens/s: every replica decodes B*pout tokens in the (max-over-ranks) time."""
    return replicas * B * pout / seconds


def rank_prompts(vocab, B, pin, rank, tp):
    """Replicas serve different requests (seed = rank); the ranks of one TP group serve the same ones."""
    return make_prompts(vocab, B, pin, seed=0 if tp > 1 else rank)


def latency_probe(llm, prompts, sp, clock=time.perf_counter):
    """p50 time-to-first-token and inter-token latency of one job through the public API, with vLLM's bench definitions
    (vllm:benchmarks/lib/endpoint_request_func.py:231-238, serve.py:578,590): all requests are sent at t0; a request's
    TTFT is the time of its first token, its ITLs are the gaps between its following tokens; medians over requests /
    over all gaps.  Tokens of one engine step share that step's completion time."""
    times = {}
    t0 = clock()

    def on_step(outs):
        t = clock()
        for rid, tok, _f, _n in outs:
            if tok >= 0:
                times.setdefault(rid, []).append(t)
    llm.generate(prompts, sp, on_step=on_step)
    ttft = sorted(ts[0] - t0 for ts in times.values() if ts)
    itl = sorted(b - a for ts in times.values() for a, b in zip(ts, ts[1:]))
    med = lambda v: v[len(v) // 2] if len(v) % 2 else 0.5 * (v[len(v) // 2 - 1] + v[len(v) // 2])   # noqa: E731
    return {"p50_ttft_ms": 1e3 * med(ttft) if ttft else None, "p50_itl_ms": 1e3 * med(itl) if itl else None,
            "requests": len(ttft), "definition": "all requests sent at t0 through LLM.generate; tokens of a step share its completion time"}


NEAR_TIE = 0.0625      # two bf16 steps of a logit in [4, 8): what two fp32 summation orders can legitimately flip


def classify_divergences(kge, llm, prompts, pout, toks_e2e, toks_res):
    """The two legs run the same prompts on the same weights but batch them differently (whole prompts per step vs the
    scheduler's chunked / mixed steps), so rows go through different kernels and fp32 summation orders and a greedy
    choice between two almost equal logits can flip; after the first flip a request follows another path by
    construction.  This replays the e2e job once more with per-token log-probabilities, requires it to reproduce the
    timed job's tokens exactly (the engine is deterministic), and for every request's FIRST token that differs between
    the legs records the engine's own top-1/top-2 log-probability gap at that step and whether the other leg's token
    is that runner-up.  not_near_tie counts first divergences that this does not explain."""
    sp = kge.SamplingParams(max_tokens=pout, temperature=0.0, ignore_eos=True, logprobs=2)
    outs = llm.generate(prompts, sp)
    toks_lp = [o.outputs[0].token_ids for o in outs]
    gaps, bad, first_steps = [], 0, []
    for b, (x, y) in enumerate(zip(toks_e2e, toks_res)):
        s = next((i for i, (a, c) in enumerate(zip(x, y)) if a != c), None)
        if s is None:
            continue
        first_steps.append(s)
        lp = outs[b].outputs[0].logprobs[s]                    # {token: Logprob}, sampled first, then the top-2
        top = sorted(((v.logprob, t) for t, v in lp.items()), reverse=True)
        gap = top[0][0] - top[1][0] if len(top) > 1 else float("inf")
        runner_up = {t for _, t in top[:2]}
        gaps.append(gap)
        if not (gap <= NEAR_TIE and y[s] in runner_up):
            bad += 1
    gaps.sort()
    return {"deterministic_replay": toks_lp == toks_e2e, "requests": len(toks_e2e), "first_divergences": len(gaps),
            "not_near_tie": bad, "near_tie_threshold": NEAR_TIE, "max_gap": gaps[-1] if gaps else None,
            "median_gap": gaps[len(gaps) // 2] if gaps else None,
            "median_first_divergent_step": sorted(first_steps)[len(first_steps) // 2] if first_steps else None}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d.get("bf16_tflops_sustained", d["bf16_tflops"]), "measured"
    return 6650.0, 1400.0, "fallback"


_CPU_MODELS = {}


def cpu_reference(workload, threads=None):
    """The CPU oracle (PyTorch restatement of vLLM's arithmetic) timed on a bounded sample of the
    workload on the host cores.  Returns (output tok/s, dict).  The model is built once per process
    (outside the timed region)."""
    import torch
    from oracle import model as om
    # every core this process may run on, whatever OMP_NUM_THREADS says (torchrun exports OMP_NUM_THREADS=1, which made
    # the reference arm of a multi-rank launch run on one core)
    torch.set_num_threads(threads or len(os.sched_getaffinity(0)))
    cores = torch.get_num_threads()
    name = WORKLOADS[workload][0]
    n, pin, pout = CPU_SAMPLE[workload]
    cfg = om.CONFIGS[name]
    if workload not in _CPU_MODELS:
        _CPU_MODELS[workload] = om.OracleModel(cfg, om.make_timing_weights(cfg), torch.bfloat16, max_pos=pin + pout + 8)
    m = _CPU_MODELS[workload]
    prompts = make_prompts(cfg.vocab_size, n, pin)
    t0 = time.perf_counter()
    out = m.generate_greedy(prompts, pout)
    dt = time.perf_counter() - t0
    assert all(len(o) == pout for o in out)
    v = n * pout / dt
    return v, {"value": v, "unit": "output tokens/s", "cores": cores, "kind": "port",
               "sample": f"{name} bf16 CPU oracle (layer-aliased timing weights), {n} requests x {pin}-in/{pout}-out, "
                         f"{dt:.1f} s wall"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    for _ in range(args.warmup_ref):
        cpu_reference(args.workload)
    info = None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        v, info = cpu_reference(args.workload)
        vals.append(v)
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    v = sum(vals) / len(vals)
    name, B, pin, pout, _ = WORKLOADS[args.workload]
    info["value"] = v
    emit({
        "impl": "reference", "metric": "output tokens/sec at fixed concurrency", "value": v, "unit": "output tokens/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup_ref, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {name}, {B} concurrent requests, {pin}-in/{pout}-out, greedy "
                               f"(CPU arm runs the bounded sample named in cpu_baseline.sample)"},
        "cpu_baseline": info,
        "e2e": {"value": v, "unit": "output tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def connect_tp(kge, llm, how, rank):
    """Join the ranks of one tensor-parallel engine: peer-memory exchange (default) or NCCL (--tp-exchange nccl, what the
    reference's --disable-custom-all-reduce selects)."""
    import torch
    import torch.distributed as dist
    if how == "nccl":
        import ctypes as C
        buf = C.create_string_buffer(128)
        if rank == 0:
            kge.check(kge.lib().kge_nccl_unique_id(buf), "kge_nccl_unique_id")
        t = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).cuda()
        dist.broadcast(t, 0)
        llm.init_tp(bytes(t.cpu().numpy().tobytes()))
    else:
        llm.tp_connect_torch()


# decode-only HBM ceilings of Llama-3.1-70B on the C3 load, output tok/s
C4_DECODE_CEILING = {2: 4540.0, 4: 9080.0, 8: 18160.0}


def tp_subrecord(kge, args, world, rank, local, mx, barrier):
    """BASELINE config 4 on the same ranks: ONE Llama-3.1-70B engine sharded over all `world` GPUs (tensor parallel,
    strong scaling), same 64 x 2048-in/256-out load, device-resident replay like the headline `value`, plus one job
    through the public API and the per-category device times of an instrumented job."""
    import tempfile
    name, B, pin, pout, pbatch = WORKLOADS["c4"]
    hf = model_cfg(name)
    prompts = rank_prompts(hf["vocab_size"], B, pin, rank, world)
    bs = args.block_size
    nblk = B * ((pin + pout + bs - 1) // bs) + 2
    with tempfile.TemporaryDirectory() as d:
        json.dump(hf, open(os.path.join(d, "config.json"), "w"))
        llm = kge.LLM(d, max_model_len=pin + pout + 16, max_num_seqs=B, max_num_batched_tokens=pbatch * pin,
                      num_gpu_blocks_override=nblk, load_format="dummy", seed=1234, device=local, tensor_parallel_size=world,
                      tp_rank=rank, block_size=bs, enable_prefix_caching=False)
    connect_tp(kge, llm, args.tp_exchange, rank)
    log(f"tp{world} engine ready: {llm.weight_bytes / 1e9:.1f} GB weights/rank")
    llm.run_resident(prompts, pout, pbatch)
    barrier()
    ms = dms = 0.0
    steps = max(1, min(2, args.steps))
    for _ in range(steps):
        _, m, dm = llm.run_resident(prompts, pout, pbatch)
        ms += m
        dms += dm
    barrier()
    ms, dms = mx(ms / steps), mx(dms / steps)
    sp = kge.SamplingParams(max_tokens=pout, temperature=0.0, ignore_eos=True)
    barrier()
    t0 = time.perf_counter()
    llm.generate(prompts, sp)
    barrier()
    e2e_s = mx(time.perf_counter() - t0)
    llm.set_timing(True)
    llm.generate(prompts, sp)
    stats = llm.kernel_stats()
    llm.set_timing(False)
    barrier()
    llm.close()
    total = sum(s["ms"] for s in stats.values())
    ar = stats.get("all_reduce", {"ms": 0.0, "launches": 0})
    value = B * pout / (ms * 1e-3)
    dec_step = dms / (pout - 1)
    ceil = C4_DECODE_CEILING.get(world)
    return {"workload": f"c4: {name}, {B} concurrent requests, {pin}-in/{pout}-out, greedy, tensor parallel over {world} GPUs",
            "parallelism": f"tp{world}", "exchange": ("peer memory: fused all-reduce + residual + norm kernel, vocab-sharded LM head + logits all-gather"
                                                     if args.tp_exchange == "peer" else "NCCL all-reduce / all-gather"),
            "scaling": "strong", "value": value, "unit": "output tokens/s", "ms_per_job": ms, "steps": steps,
            "decode_ms_per_step": dec_step, "decode_tokens_per_s": B / (dec_step * 1e-3),
            "decode_hbm_ceiling_tokens_per_s": ceil, "decode_frac_of_hbm_ceiling": (B / (dec_step * 1e-3) / ceil) if ceil else None,
            "prefill_ms_per_8192_token_step": (ms - dms) / (B / pbatch),
            "e2e": {"value": B * pout / e2e_s, "unit": "output tokens/s", "ms_per_job": e2e_s * 1e3},
            "exchange_ms_of_instrumented_job": ar["ms"], "exchange_launches": ar["launches"],
            "exchange_share_of_kernel_time": ar["ms"] / total if total else None,
            "kernels_ms": {k: round(s["ms"], 1) for k, s in stats.items()}}


def log(msg):
    sys.stderr.write(f"[bench {time.strftime('%H:%M:%S')} rank {os.environ.get('RANK', '0')}] {msg}\n")
    sys.stderr.flush()


def run_ours(args):
    import torch
    import torch.distributed as dist
    from kge_loader import load
    kge = load()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    kge.require_cuda()
    name, B, pin, pout, pbatch = WORKLOADS[args.workload]
    hf = model_cfg(name)
    vocab = hf["vocab_size"]
    tp = world if args.workload == "c4" else 1      # c4: ONE model sharded over all ranks; else replicas
    replicas = 1 if tp > 1 else world
    prompts = rank_prompts(vocab, B, pin, rank, tp)
    bs = args.block_size
    nblk = B * ((pin + pout + bs - 1) // bs) + 2
    with tempfile.TemporaryDirectory() as d:
        json.dump(hf, open(os.path.join(d, "config.json"), "w"))
        llm = kge.LLM(d, max_model_len=pin + pout + 16, max_num_seqs=max(B, 8),
                      max_num_batched_tokens=max(pbatch * pin, 256), num_gpu_blocks_override=nblk,
                      load_format="dummy", seed=1234, device=local, tensor_parallel_size=tp, tp_rank=rank if tp > 1 else 0,
                      block_size=bs, enable_prefix_caching=False)      # the same prompts are replayed every step: a cache hit would skip the prefill
    if tp > 1:
        connect_tp(kge, llm, args.tp_exchange, rank)
    sp = kge.SamplingParams(max_tokens=pout, temperature=0.0, ignore_eos=True)
    log(f"engine ready: {llm.weight_bytes / 1e9:.1f} GB weights, {llm.num_gpu_blocks} KV blocks, tp={tp}")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def mx(x):
        return max_over_ranks(x, world)

    # ---- value: device-resident replay of the job's own steps.  One job goes through the public API with recording on;
    # the replay then runs exactly the steps the scheduler made (same batches, block tables and kernels) with all metadata
    # in HBM and sampled ids handed from step to step on the device, CUDA events on the engine's stream.
    llm.record_steps(True)
    outs_rec = llm.generate(prompts, sp)
    llm.record_steps(False)
    rec_ids = [o.request_id for o in outs_rec]
    toks_rec = [o.outputs[0].token_ids for o in outs_rec]
    cap = B * pout + 64
    for i in range(args.warmup):
        _, n_steps, ms_w, _ = llm.replay_recorded(cap)
        log(f"replay warm-up {i}: {ms_w:.0f} ms, {n_steps} steps")
    barrier()
    clocks = ClockSampler(local)
    clocks.start()
    l0 = llm.kernel_launches
    ms_total = dec_total = 0.0
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        rep, n_steps, ms, dms = llm.replay_recorded(cap)
        ms_total += ms
        dec_total += dms
    barrier()
    wall_resident = time.perf_counter() - t_wall0
    launches = llm.kernel_launches - l0
    clk = clocks.summary()
    toks_res = [rep[r] for r in rec_ids]
    ms_step = mx(ms_total / args.steps)
    value = job_throughput(replicas, B, pout, ms_step * 1e-3)
    # fixed-shape companion (whole prompts, pbatch per step, then full-batch decode): the decode step at full batch
    llm.run_resident(prompts, pout, pbatch)
    _, ms_fix, dms_fix = llm.run_resident(prompts, pout, pbatch)
    dec_ms_step = mx(dms_fix)
    fixed_shape = {"what": f"{B // pbatch} prefill steps of {pbatch} whole prompts, then {pout - 1} decode steps of all {B} requests (kge_engine_run_resident)",
                   "ms_per_job": mx(ms_fix), "value": job_throughput(replicas, B, pout, mx(ms_fix) * 1e-3), "decode_ms_per_step": dec_ms_step / max(1, pout - 1),
                   "prefill_ms_per_step": (mx(ms_fix) - dec_ms_step) / (B / pbatch)}

    # ---- e2e: public API with host buffers, scheduler-driven, per-step H2D/D2H inside the timing
    io0 = llm.io_bytes()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        outs = llm.generate(prompts, sp)
    barrier()
    e2e_s = mx((time.perf_counter() - t0) / args.steps)
    io1 = llm.io_bytes()
    e2e = job_throughput(replicas, B, pout, e2e_s)
    log(f"value {value:.0f} tok/s, e2e {e2e:.0f} tok/s")
    toks_e2e = [o.outputs[0].token_ids for o in outs]
    same = sum(a == b for x, y in zip(toks_e2e, toks_res) for a, b in zip(x, y))
    # both legs run the same steps, so their tokens must be the same; if they ever differ, say where and how closely
    divergence = classify_divergences(kge, llm, prompts, pout, toks_e2e, toks_res) if same != B * pout else \
        {"requests": B, "first_divergences": 0, "not_near_tie": 0, "recorded_job_equal_to_timed_job": toks_rec == toks_e2e}
    log(f"legs: {same}/{B * pout} tokens equal; {divergence}")

    # ---- latency (extra job through the public API; BASELINE.json's metric also names p50 TTFT and ITL)
    try:
        latency = latency_probe(llm, prompts, sp)
    except Exception as ex:      # never let the side measurement take the bench line down
        latency = {"error": f"{type(ex).__name__}: {ex}"}

    # ---- per-kernel breakdown (instrumented extra job; never part of a headline number)
    llm.set_timing(True)
    llm.generate(prompts, sp)
    stats = llm.kernel_stats()
    llm.set_timing(False)
    hbm, tf, src = peaks()
    kernels = {}
    for k, s in stats.items():
        ms_k = s["ms"]
        kernels[k] = {"launches": s["launches"], "ms": round(ms_k, 3),
                      "GB/s": round(s["bytes"] / ms_k / 1e6, 1) if ms_k else None,
                      "TFLOP/s": round(s["flops"] / ms_k / 1e9, 1) if ms_k and s["flops"] else None}
    dom = max(stats, key=lambda k: stats[k]["ms"])
    tensor_bound = dom in ("gemm_prefill", "attn_general")
    s = stats[dom]
    per_launch_ms = s["ms"] / s["launches"]
    if tensor_bound:
        ach = s["flops"] / s["ms"] / 1e9
        roof = {"kernel": dom, "bound": "tensor", "achieved": ach, "peak": tf, "unit": "TFLOP/s", "frac": ach / tf}
    else:
        ach = s["bytes"] / s["ms"] / 1e6
        roof = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm}
    traffic, traffic_of = None, None
    try:        # DRAM bytes of one launch of this kernel from the committed `ncu --set full` capture (never measured live)
        t = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_traffic.json")))[dom]
        traffic, traffic_of = t["traffic"], f"{t['launch']}; algorithmic bytes of that launch {t['algorithmic']}; {t['source']}"
    except (OSError, KeyError, ValueError):
        pass
    roof.update({"traffic": traffic, "traffic_of": traffic_of, "avg_launch_ms": per_launch_ms, "launches": s["launches"],
                 "peak_source": f"MEASURED_PEAKS.json ({src}; sustained bf16 for a kernel inside a long step)"})
    line = {
        "metric": "output tokens/sec at fixed concurrency", "value": value, "unit": "output tokens/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong" if tp > 1 else "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic token-id prompts U[1000,V-1000), hash-random weights (--load-format dummy)",
        "config": {"workload": f"{args.workload}: {name}, {B} concurrent requests, {pin}-in/{pout}-out, greedy, "
                               f"paged KV block {bs}, {pbatch} prompts per prefill step",
                   "parallelism": (f"tp{tp} ({'peer-memory fused all-reduce + residual + norm' if args.tp_exchange == 'peer' else 'NCCL bf16 all-reduce'} after o_proj and down_proj)" if tp > 1
                                   else f"dp{world} (independent replicas, no collective)"),
                   "l2": "every step streams weights + KV >> 126 MB L2 (inputs larger than L2, no flush needed)",
                   "prefix_caching": "off: every job recomputes every prompt (the same prompts are replayed each step)"},
        "steps_per_job": n_steps, "trailing_decode_ms": mx(dec_total / args.steps), "fixed_shape": fixed_shape,
        "decode_ms_per_token_step": fixed_shape["decode_ms_per_step"],
        "e2e": {"value": e2e, "unit": "output tokens/s", "h2d_bytes_per_step": (io1[0] - io0[0]) // args.steps,
                "d2h_bytes_per_step": (io1[1] - io0[1]) // args.steps, "ms_per_step": e2e_s * 1e3,
                "timer": "host wall clock between device synchronisations; includes the C++ scheduler",
                "tokens_equal_to_resident_leg": f"{same}/{B * pout}", "leg_divergence": divergence},
        "latency": latency, "gpu_launches": launches, "clocks": clk, "roofline": roof, "kernels": kernels,
        "resident_wall_s": wall_resident,
    }
    llm.close()
    if world > 1 and tp == 1 and not args.no_tp:
        # BASELINE config 4 on the same ranks.  Guarded: whatever happens in this leg, the headline line is printed -
        # a watchdog prints it without the sub-record and ends the process if the leg does not finish in time.
        import threading

        def give_up():
            if rank == 0:
                line["tp"] = {"error": f"tensor-parallel leg did not finish within {args.tp_timeout} s"}
                emit(line)
            os._exit(0)
        dog = threading.Timer(args.tp_timeout, give_up)
        dog.daemon = True
        dog.start()
        try:
            line["tp"] = tp_subrecord(kge, args, world, rank, local, mx, barrier)
        except Exception as ex:
            line["tp"] = {"error": f"{type(ex).__name__}: {ex}"[:500]}
        dog.cancel()
    if world == 1 and not args.no_http and args.workload != "c4":
        # the same job as it is deployed: SSE clients -> kge-router -> `vllm serve` (one engine process on this GPU);
        # a side measurement, never allowed to take the bench line down
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import http_bench
            line["http"] = http_bench.measure(1, args.workload, jobs=1, warm_jobs=1, log=log, startup_timeout=240)
            line["http"]["vs_e2e"] = line["http"]["value"] / e2e
        except Exception as ex:
            line["http"] = {"error": f"{type(ex).__name__}: {ex}"[:400]}
    if rank == 0:
        if world == 1 and not args.no_cpu:
            _, line["cpu_baseline"] = cpu_reference(args.workload)
        emit(line)
    if world > 1:
        try:
            dist.destroy_process_group()
        except Exception:
            pass


_REAL_STDOUT = None


def protect_stdout():
    """stdout carries exactly ONE JSON line: anything libraries print there (NCCL's version banner,
    torchrun notices) is sent to stderr instead; emit() writes to the saved descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--block-size", type=int, default=64, help="tokens per KV block (the server's default, 64)")
    ap.add_argument("--no-http", action="store_true", help="at N = 1, skip the HTTP (router + server) sub-record")
    ap.add_argument("--no-tp", action="store_true", help="at N > 1, skip the tensor-parallel (config 4) sub-record")
    ap.add_argument("--tp-exchange", default="peer", choices=["peer", "nccl"], help="TP data path: peer memory (default) or NCCL")
    ap.add_argument("--tp-timeout", type=float, default=420.0, help="seconds the tensor-parallel leg may take at N > 1")
    args = ap.parse_args()
    args.warmup_ref = args.warmup        # same warm-up count as the GPU arm
    protect_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
