"""The N > 1 host logic of bench.py on CPU: two gloo ranks (127.0.0.1) agree on the max-over-ranks
time, serve different prompts as replicas (same prompts inside a TP group) and aggregate whole-job
throughput; rank 0 alone runs the `--impl reference` arm."""
import json
import os
import subprocess
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(rank, world, port, q):
    import bench
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = 1.0 + rank                                   # rank 1 is the slow one
    t = bench.max_over_ranks(mine, world, device="cpu")
    pr = bench.rank_prompts(5000, 2, 8, rank, tp=1)
    pt = bench.rank_prompts(5000, 2, 8, rank, tp=2)
    q.put((rank, t, bench.job_throughput(world, 64, 256, t), pr, pt))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_reduction_and_sharding():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, t0, v0, pr0, pt0), (r1, t1, v1, pr1, pt1) = res
    assert t0 == t1 == 2.0                               # slowest rank decides
    assert v0 == v1 == 2 * 64 * 256 / 2.0                # whole-job aggregate over both replicas
    assert pr0 != pr1                                    # replicas serve different requests
    assert pt0 == pt1                                    # a TP group serves the same requests


def test_world_one_is_identity():
    import bench
    assert bench.max_over_ranks(3.5, 1) == 3.5


def test_reference_arm_rank_gt0_prints_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       env=env, capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_reference_arm_line_shape():
    """c1 sample (opt-125m, 1 x 128-in/128-out) is small enough for the CPU suite."""
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c1",
                        "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-1500:]
    line = json.loads(p.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "output tokens/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "output tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_latency_probe_uses_vllm_bench_definitions():
    """TTFT = first token - send time; ITL = gaps between a request's following tokens; medians."""
    import bench

    class FakeLLM:
        def generate(self, prompts, sp, on_step=None):
            for step in range(1, 7):                      # request 2's prompt needs two steps
                on_step([(rid, 5, 0, 1) if step >= (1 if rid < 2 else 2) else (rid, -1, 0, 8) for rid in range(3)])

    ticks = iter([0.0] + [0.010 * i + (0.005 if i == 4 else 0.0) for i in range(1, 8)])
    r = bench.latency_probe(FakeLLM(), None, None, clock=lambda: next(ticks))
    assert r["requests"] == 3 and abs(r["p50_ttft_ms"] - 10.0) < 1e-9           # first tokens at 10, 10, 20 ms
    assert abs(r["p50_itl_ms"] - 10.0) < 1e-6                                   # gaps: mostly 10 ms, one 15 and one 5
