"""Real-mode (CUDA graphs + programmatic dependent launch) decode step time of C3 (llama-3.1-8b shapes, 64
requests, ctx 2048..2048+N) under different launch-shape knobs.  One engine per setting (captured graphs keep
their launch shapes).  python tools/probe_decode.py '{"gemm_ws":1}' '{"gemm_even_grid":0}' ..."""
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from kge_loader import load  # noqa: E402

kge = load()
L = kge.lib()
BS = int(os.environ.get("PROBE_BLOCK", "64"))
DEFAULTS = {"gemm_wide": 1, "attn_persistent": 1, "gemm_l2_hints": 0, "smem_carveout": 1, "gemm_ws": 0, "attn_tma": 1, "pdl": 1, "gemm_even_grid": 1, "attn_decode_splits": 0, "attn_split_balance": 0, "attn_cluster_merge": 0}


def main():
    kge.require_cuda()
    settings = [json.loads(a) for a in sys.argv[1:]] or [{}]
    name, B, pin, pout, pbatch = bench.WORKLOADS["c3"]
    pout = int(os.environ.get("PROBE_OUT", "49"))
    hf = bench.model_cfg(name)
    prompts = bench.rank_prompts(hf["vocab_size"], B, pin, 0, 1)
    for st in settings:
        kn = dict(DEFAULTS, **st)
        for k, v in kn.items():
            kge.check(L.kge_set_tunable(k.encode(), int(v)), "kge_set_tunable")
        with tempfile.TemporaryDirectory() as d:
            json.dump(hf, open(os.path.join(d, "config.json"), "w"))
            llm = kge.LLM(d, max_model_len=pin + 256 + 16, max_num_seqs=B, max_num_batched_tokens=pbatch * pin,
                          num_gpu_blocks_override=B * ((pin + 256 + BS - 1) // BS) + 2, load_format="dummy", seed=1234, block_size=BS)
        llm.run_resident(prompts, pout, pbatch)
        res = []
        for _ in range(2):
            _, ms, dms = llm.run_resident(prompts, pout, pbatch)
            res.append((ms, dms))
        llm.close()
        ms, dms = min(res, key=lambda r: r[1])
        print(json.dumps({"knobs": st, "decode_ms_per_step": round(dms / (pout - 1), 4), "prefill_ms_per_step": round((ms - dms) / (B / pbatch), 3)}),
              flush=True)


if __name__ == "__main__":
    main()
