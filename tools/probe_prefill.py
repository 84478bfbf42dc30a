"""In-engine per-category kernel times of single C3 steps (llama-3.1-8b shapes, dummy weights), through
kge_step with timing on: a pure prefill step (4 x 2048), a chunk-misaligned prefill step as the vLLM policy
produces them, and a decode step (64 requests at ctx 2176).  Prints one JSON line per step shape.
  python tools/probe_prefill.py [--layers N]"""
import argparse
import json
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from kge_loader import load  # noqa: E402

kge = load()
BS = 16


def batch(reqs):
    """reqs: list of (prefix_len, q_len).  Blocks are handed out contiguously per request."""
    ids, pos, qsl, sl, rows, slots = [], [], [0], [], [], []
    nb_max = max((p + q + BS - 1) // BS for p, q in reqs)
    bt = np.zeros((len(reqs), nb_max), dtype=np.int32)
    nxt = 1
    rng = np.random.default_rng(0)
    for i, (p, q) in enumerate(reqs):
        nb = (p + q + BS - 1) // BS
        bt[i, :nb] = np.arange(nxt, nxt + nb)
        nxt += nb
        ids += rng.integers(1000, 100000, q).tolist()
        pos += list(range(p, p + q))
        slots += [int(bt[i, t // BS]) * BS + t % BS for t in range(p, p + q)]
        qsl.append(qsl[-1] + q)
        sl.append(p + q)
        rows.append(qsl[-1] - 1)
    return ids, pos, qsl, sl, bt, nb_max, slots, rows, nxt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--iters", type=int, default=3)
    a = ap.parse_args()
    kge.require_cuda()
    hf = bench.model_cfg("llama-3.1-8b")
    hf["num_hidden_layers"] = a.layers
    shapes = {
        "prefill 4x2048": [(0, 2048)] * 4,
        "prefill misaligned (4@2044, 3x2048, 2037)": [(2044, 4), (0, 2048), (0, 2048), (0, 2048), (0, 2037)],
        "mixed 7 decodes + misaligned prefill": [(2048 + i, 1) for i in range(7)] + [(2044, 4), (0, 2048), (0, 2048), (0, 2048), (0, 2030)],
        "decode 64 @2176": [(2175, 1)] * 64,
    }
    with tempfile.TemporaryDirectory() as d:
        json.dump(hf, open(os.path.join(d, "config.json"), "w"))
        llm = kge.LLM(d, max_model_len=2400, max_num_seqs=64, max_num_batched_tokens=8192, num_gpu_blocks_override=64 * 150 + 2,
                      load_format="dummy", seed=1234)
    for name, reqs in shapes.items():
        b = batch(reqs)
        args = b[:8]
        for _ in range(2):
            llm.raw_step(*args)
        llm.set_timing(True)
        for _ in range(a.iters):
            llm.raw_step(*args)
        st = llm.kernel_stats()
        llm.set_timing(False)
        row = {"step": name, "layers": a.layers, "T": len(b[0])}
        tot = 0.0
        for k, s in st.items():
            row[k] = {"n": s["launches"] // a.iters, "ms": round(s["ms"] / a.iters, 3),
                      "us_per_launch": round(1e3 * s["ms"] / s["launches"], 1),
                      "TF": round(s["flops"] / s["ms"] / 1e9, 1) if s["flops"] else None,
                      "GB/s": round(s["bytes"] / s["ms"] / 1e6, 1)}
            tot += s["ms"] / a.iters
        row["sum_ms"] = round(tot, 3)
        print(json.dumps(row), flush=True)
    llm.close()


if __name__ == "__main__":
    main()
