"""The job profiled by tools/ncu_capture.sh: one device-resident C3 pass with a short decode tail
(64 x 2048-token prompts in 16 prefill steps, then NCU_OUT-1 decode steps).  Prints nothing that is a bench value."""
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from kge_loader import load  # noqa: E402

kge = load()
kge.require_cuda()
name, B, pin, _, pbatch = bench.WORKLOADS["c3"]
pout = int(os.environ.get("NCU_OUT", "9"))
hf = bench.model_cfg(name)
prompts = bench.rank_prompts(hf["vocab_size"], B, pin, 0, 1)
with tempfile.TemporaryDirectory() as d:
    json.dump(hf, open(os.path.join(d, "config.json"), "w"))
    llm = kge.LLM(d, max_model_len=pin + 256 + 16, max_num_seqs=B, max_num_batched_tokens=pbatch * pin,
                  num_gpu_blocks_override=B * ((pin + 256 + 63) // 64) + 2, load_format="dummy", seed=1234, block_size=64)
_, ms, dms = llm.run_resident(prompts, pout, pbatch)
print(json.dumps({"ncu_job_ms": ms, "decode_ms": dms, "launches": llm.kernel_launches}))
llm.close()
