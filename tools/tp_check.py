"""Tensor-parallel parity check: one process per rank (RANK / WORLD_SIZE env), each with its own engine on
GPU `rank`, NCCL unique id passed through a file.  Rank 0 compares the greedy ids with the CPU oracle and
prints one JSON line.  Used by tests/test_tp_gpu.py and by hand:
  for r in 0 1; do RANK=$r WORLD_SIZE=2 python tools/tp_check.py tiny-llama /tmp/idfile & done; wait"""
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from kge_loader import load  # noqa: E402
from oracle import model as om  # noqa: E402


def main():
    name, idfile = sys.argv[1], sys.argv[2]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    kge = load()
    L = kge.lib()
    torch.cuda.set_device(rank)
    cfg = om.CONFIGS[name]
    w = om.make_weights(cfg, seed=0)
    prompts = [om.make_prompts(cfg, 1, n, seed=n)[0] for n in (9, 40, 70)]
    N = 12
    with tempfile.TemporaryDirectory() as d:
        om.save_model_dir(d, cfg, w)
        llm = kge.LLM(d, max_model_len=128, max_num_seqs=8, max_num_batched_tokens=96, num_gpu_blocks_override=64,
                      tensor_parallel_size=world, tp_rank=rank, device=rank)
    if rank == 0:
        import ctypes as C
        buf = C.create_string_buffer(128)
        kge.check(L.kge_nccl_unique_id(buf), "unique id")
        with open(idfile + ".tmp", "wb") as f:
            f.write(buf.raw)
        os.replace(idfile + ".tmp", idfile)
        uid = buf.raw
    else:
        t0 = time.time()
        while not os.path.exists(idfile):
            assert time.time() - t0 < 120, "rank 0 never wrote the NCCL id"
            time.sleep(0.05)
        uid = open(idfile, "rb").read()
    llm.init_tp(uid)
    outs = llm.generate(prompts, kge.SamplingParams(max_tokens=N, temperature=0.0, ignore_eos=True))
    got = [o.outputs[0].token_ids for o in outs]
    llm.close()
    if rank == 0:
        ref_ids, ref_lg = om.OracleModel(cfg, w, torch.bfloat16).generate_greedy(prompts, N, return_logits=True)
        equal = near = bad = 0
        for b, (g, r) in enumerate(zip(got, ref_ids)):
            for s, (a, e) in enumerate(zip(g, r)):
                if a == e:
                    equal += 1
                    continue
                t2 = ref_lg[s][b].topk(2).values
                if (t2[0] - t2[1]).item() <= 0.06 and abs(ref_lg[s][b][a].item() - t2[0].item()) <= 0.06:
                    near += 1          # bf16 near-tie: TP sums bf16-rounded partials, so low bits differ by design
                else:
                    bad += 1
                break
        print(json.dumps({"model": name, "tp": world, "equal": equal, "near_tie_divergences": near, "bad": bad,
                          "total": len(prompts) * N}))


if __name__ == "__main__":
    main()
