import sys, torch, tempfile
sys.path.insert(0, '.')
from kge_loader import load
from oracle import model as om
kge = load()
BF = torch.bfloat16
cfg = om.CONFIGS["small-llama"]
w = om.make_weights(cfg, seed=0)
prompts = om.make_prompts(cfg, 2, 40, seed=9)
m = om.OracleModel(cfg, w, BF)
ref_ids, ref_lg = m.generate_greedy(prompts, 2, return_logits=True)
with tempfile.TemporaryDirectory() as d:
    om.save_model_dir(d, cfg, w)
    llm = kge.LLM(d, max_model_len=128, max_num_seqs=4, max_num_batched_tokens=128, num_gpu_blocks_override=32)
    bs = 16
    ids = prompts[0] + prompts[1]
    pos = list(range(40)) * 2
    bt = [[1, 2, 3], [4, 5, 6]]
    slots = [bt[b][t // bs] * bs + t % bs for b in range(2) for t in range(40)]
    out, lg = llm.raw_step(ids, pos, [0, 40, 80], [40, 40], bt, 3, slots, [39, 79], want_logits=True)
    got = torch.from_numpy(lg); ref = ref_lg[0]
    print("out", out.tolist(), "got.argmax", got.argmax(-1).tolist(), "ref.argmax", ref.argmax(-1).tolist())
    print("scale", ref.abs().max().item(), "maxerr", (got-ref).abs().max().item(), "meanerr", (got-ref).abs().mean().item())
    for b in range(2):
        print(b, "ref top3", ref[b].topk(3), "got top3", got[b].topk(3))
        print("   got at ref argmax", got[b][ref[b].argmax()].item(), "ref at got argmax", ref[b][got[b].argmax()].item())
    # single-request prefill for comparison
    out1, lg1 = llm.raw_step(prompts[0], list(range(40)), [0, 40], [40], [[7, 8, 9]], 3,
                             [[7,8,9][t // bs] * bs + t % bs for t in range(40)], [39], want_logits=True)
    g1 = torch.from_numpy(lg1)
    print("single: out", out1.tolist(), "maxerr vs ref", (g1[0]-ref[0]).abs().max().item(), "vs batched", (g1[0]-got[0]).abs().max().item())
    llm.close()
