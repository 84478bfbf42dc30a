"""Third opinion (oracle plan item 3): the installed vLLM 0.22.0 running its own CUDA
kernels on a B200, on the oracle's seeded tiny models (identical safetensors) and token-id prompts, greedy.
NOT the pinned version (v0.8.x) and not vLLM's CPU backend - but it is the dependency the reference deploys,
executing the same model graph, so its tokens and logprobs pin the oracle and the product end to end.

Runs ONLY on a GPU box (gpurun); needs vllm.  Writes gpurun_out/vllm_cuda.json, which is then committed as
tests/golden/vllm_cuda.json.  Nothing reads vllm at test time.

  python tests/golden/make_vllm_cuda_golden.py            # parity fixtures, one subprocess per model
  python tests/golden/make_vllm_cuda_golden.py --perf     # + C3 throughput of vLLM on this GPU (beat-this row)
"""
import argparse
import json
import os
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "..", "..")
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "gpurun_out")
MODELS = ["tiny-llama-hd128", "tiny-qwen3", "tiny-opt", "tiny-qwen2", "tiny-llama"]
DEFAULT_MODELS = MODELS[:3]
PROMPT_LENS = (5, 17, 33, 64)
NEW_TOKENS = 16
TOPN = 5


def env_setup():
    os.environ.setdefault("VLLM_ENABLE_V1_MULTIPROCESSING", "0")
    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    os.environ.setdefault("TRANSFORMERS_OFFLINE", "1")
    os.environ.setdefault("VLLM_NO_USAGE_STATS", "1")
    os.environ.setdefault("VLLM_LOGGING_LEVEL", "WARNING")


def one_model(name, backend):
    env_setup()
    import torch
    from oracle import model as om
    from vllm import LLM, SamplingParams
    cfg = om.CONFIGS[name]
    d = f"/tmp/kge_vllm_{name}"
    om.save_model_dir(d, cfg, om.make_weights(cfg, seed=0))
    prompts = [om.make_prompts(cfg, 1, n, seed=n)[0] for n in PROMPT_LENS]
    kw = dict(model=d, skip_tokenizer_init=True, dtype="bfloat16", enforce_eager=True, max_model_len=256, max_num_seqs=8,
              gpu_memory_utilization=0.3, seed=0, disable_log_stats=True, enable_prefix_caching=False)
    if backend:
        kw["attention_config"] = {"backend": backend}
    llm = LLM(**kw)
    sp = SamplingParams(temperature=0.0, max_tokens=NEW_TOKENS, ignore_eos=True, detokenize=False, logprobs=TOPN)
    outs = llm.generate([{"prompt_token_ids": p} for p in prompts], sp, use_tqdm=False)
    res = []
    for p, o in zip(prompts, outs):
        c = o.outputs[0]
        steps = []
        for lp in c.logprobs:                       # dict token_id -> Logprob, per generated token
            items = sorted(lp.items(), key=lambda kv: kv[1].rank)
            steps.append([[int(t), float(v.logprob)] for t, v in items])
        res.append(dict(prompt=p, ids=[int(t) for t in c.token_ids], top=steps))
    import vllm
    try:
        be = llm.llm_engine.engine_core.engine_core.model_executor.driver_worker.worker.model_runner.attn_groups[0][0].backend
        be = getattr(be, "__name__", type(be).__name__)
    except Exception:
        be = backend or "default"
    json.dump(dict(model=name, vllm=vllm.__version__, torch=torch.__version__, gpu=torch.cuda.get_device_name(0),
                   backend=str(be), requested_backend=backend or "", new_tokens=NEW_TOKENS, requests=res),
              open(os.path.join(OUT, f"vllm_cuda_{name}.json"), "w"))
    print("ok", name, be, flush=True)


def perf():
    """C3 of BASELINE.json on vLLM itself: llama-3.1-8b shapes, dummy weights, 64 x 2048-in / 256-out, greedy."""
    env_setup()
    import torch
    from oracle import model as om
    from vllm import LLM, SamplingParams
    cfg = om.CONFIGS["llama-3.1-8b"]
    d = "/tmp/kge_vllm_c3"
    os.makedirs(d, exist_ok=True)
    json.dump(cfg.to_hf(), open(os.path.join(d, "config.json"), "w"))
    t0 = time.time()
    llm = LLM(model=d, skip_tokenizer_init=True, dtype="bfloat16", load_format="dummy", max_model_len=4096, max_num_seqs=64,
              max_num_batched_tokens=8192, gpu_memory_utilization=0.9, seed=0, disable_log_stats=True,
              enable_prefix_caching=False, compilation_config={"cudagraph_capture_sizes": [1, 2, 4, 8, 16, 32, 64]})
    t_init = time.time() - t0
    prompts = [{"prompt_token_ids": p} for p in om.make_prompts(cfg, 64, 2048, seed=0)]
    llm.generate(prompts[:8], SamplingParams(temperature=0.0, max_tokens=8, ignore_eos=True, detokenize=False), use_tqdm=False)
    sp = SamplingParams(temperature=0.0, max_tokens=256, ignore_eos=True, detokenize=False)
    runs = []
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.time()
        outs = llm.generate(prompts, sp, use_tqdm=False)
        torch.cuda.synchronize()
        dt = time.time() - t0
        n = sum(len(o.outputs[0].token_ids) for o in outs)
        runs.append(dict(seconds=dt, output_tokens=n, tok_s=n / dt))
        print("run", runs[-1], flush=True)
    # where vLLM's time goes: the same prompts with one output token (prefill only) and with 65 (prefill + 64 decode steps
    # of the last-admitted request), so prefill and decode can be told apart without a profiler
    split = {}
    for n_out in (1, 65):
        spn = SamplingParams(temperature=0.0, max_tokens=n_out, ignore_eos=True, detokenize=False)
        best = 1e9
        for _ in range(2):
            torch.cuda.synchronize()
            t0 = time.time()
            llm.generate(prompts, spn, use_tqdm=False)
            torch.cuda.synchronize()
            best = min(best, time.time() - t0)
        split[f"seconds_out{n_out}"] = best
        print("split", n_out, best, flush=True)
    import vllm
    json.dump(dict(workload="c3: llama-3.1-8b dummy weights, 64 x 2048-in/256-out, greedy, offline LLM.generate (host prompts)", split=split,
                   vllm=vllm.__version__, gpu=torch.cuda.get_device_name(0), init_seconds=t_init, runs=runs,
                   best_tok_s=max(r["tok_s"] for r in runs)), open(os.path.join(OUT, "vllm_cuda_c3.json"), "w"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--one")
    ap.add_argument("--backend", default="")
    ap.add_argument("--perf", action="store_true")
    ap.add_argument("--perf-only", action="store_true")
    ap.add_argument("--models", default=",".join(DEFAULT_MODELS))
    ap.add_argument("--timeout", type=int, default=150)
    a = ap.parse_args()
    os.makedirs(OUT, exist_ok=True)
    if a.one:
        return one_model(a.one, a.backend)
    if a.perf_only:
        return perf()
    for i, name in enumerate(a.models.split(",")):
        ok = False
        for backend in ("", "TRITON_ATTN"):          # default selection first; Triton attention if that cannot start
            log = open(os.path.join(OUT, f"vllm_cuda_{name}.log"), "a")
            try:
                rc = subprocess.run([sys.executable, os.path.abspath(__file__), "--one", name, "--backend", backend],
                                    stdout=log, stderr=subprocess.STDOUT, timeout=a.timeout).returncode
            except subprocess.TimeoutExpired:
                rc = -9
            print(name, backend or "default", "rc", rc, flush=True)
            f = os.path.join(OUT, f"vllm_cuda_{name}.json")
            if rc == 0 and os.path.exists(f):
                ok = True
                break
        if i == 0 and not ok:
            print("vLLM did not start on this box with either attention backend; see gpurun_out/vllm_cuda_*.log")
            sys.exit(3)
    merged = [json.load(open(os.path.join(OUT, f"vllm_cuda_{m}.json"))) for m in MODELS
              if os.path.exists(os.path.join(OUT, f"vllm_cuda_{m}.json"))]
    json.dump(merged, open(os.path.join(OUT, "vllm_cuda.json"), "w"), separators=(",", ":"))
    print("vllm_cuda.json:", [m["model"] for m in merged])
    if a.perf:
        log = open(os.path.join(OUT, "vllm_cuda_c3.log"), "w")
        try:
            rc = subprocess.run([sys.executable, os.path.abspath(__file__), "--perf-only"], stdout=log, stderr=subprocess.STDOUT,
                                timeout=330).returncode
        except subprocess.TimeoutExpired:
            rc = -9
        print("perf rc", rc)


if __name__ == "__main__":
    main()
