"""Config 5 measured the way it is deployed (ref:old_README.md:1175, ref:values-01-minimal-example2.yaml:10): N engine
processes (`bin/vllm serve`, one GPU each) behind the request router, driven over HTTP with streaming completions.

Definitions are vLLM's bench client's (vllm:benchmarks/serve.py:573,578,590; benchmarks/lib/endpoint_request_func.py:
231-238): every request is sent at t0 (fixed concurrency = all of them in flight), output throughput = generated tokens /
wall time until the last stream closes, TTFT = first SSE chunk - send, ITL = gaps between a request's chunks, medians.
Prompts are token-id lists (no tokenizer offline), max_tokens fixed, ignore_eos, temperature 0, stream true.

  python tools/http_bench.py --gpus 1 --workload c3 [--jobs 2] [--out FILE]        # prints one JSON document
Used by bench.py at N = 1 (the `http` sub-record) and by hand for N replicas."""
import argparse
import asyncio
import json
import os
import socket
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = os.path.join(ROOT, "kubernetes-gpu-cluster_b200")


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def wait_http(port, path, timeout):
    import http.client
    t0 = time.time()
    while time.time() - t0 < timeout:
        try:
            c = http.client.HTTPConnection("127.0.0.1", port, timeout=2)
            c.request("GET", path)
            if c.getresponse().status == 200:
                return True
        except OSError:
            pass
        time.sleep(0.2)
    return False


async def one_request(port, body: bytes, rec):
    reader, writer = await asyncio.open_connection("127.0.0.1", port)
    req = (b"POST /v1/completions HTTP/1.1\r\nHost: bench\r\nContent-Type: application/json\r\nConnection: close\r\n"
           b"Content-Length: " + str(len(body)).encode() + b"\r\n\r\n" + body)
    sent = time.perf_counter()
    writer.write(req)
    await writer.drain()
    times = []
    status = None
    while True:
        line = await reader.readline()
        if not line:
            break
        if status is None and line.startswith(b"HTTP/1.1 "):
            status = int(line.split()[1])
        if line.startswith(b"data: "):
            if line.startswith(b"data: [DONE]"):
                break
            if b'"token_ids":[' in line or b'"text"' in line:
                times.append(time.perf_counter())
            elif b'"error"' in line:
                rec["errors"].append(line[:200].decode(errors="replace"))
    writer.close()
    rec["status"].append(status)
    rec["sent"].append(sent)
    rec["chunks"].append(times)


async def run_job(port, bodies):
    rec = {"status": [], "sent": [], "chunks": [], "errors": []}
    t0 = time.perf_counter()
    await asyncio.gather(*(one_request(port, b, rec) for b in bodies))
    return rec, time.perf_counter() - t0


def median(v):
    v = sorted(v)
    return None if not v else (v[len(v) // 2] if len(v) % 2 else 0.5 * (v[len(v) // 2 - 1] + v[len(v) // 2]))


def measure(gpus, workload, jobs, warm_jobs, log=lambda m: None, startup_timeout=300, env_extra=None, shape=None):
    """shape = (model name, streams per engine, prompt length, output length, prompts per prefill step) overrides the named
    workload's; env_extra is added to the engines' environment (the CPU test points LD_LIBRARY_PATH at the mock engine)."""
    import bench
    name, B, pin, pout, pbatch = shape or bench.WORKLOADS[workload]
    hf = bench.model_cfg(name)
    served = "bench-model"
    procs = []
    tmp = tempfile.TemporaryDirectory()
    try:
        json.dump(hf, open(os.path.join(tmp.name, "config.json"), "w"))
        ports = [free_port() for _ in range(gpus)]
        nblk = B * ((pin + pout + 63) // 64) + 2        # the server's default KV block is 64 tokens
        for g, port in enumerate(ports):
            env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(g), **(env_extra or {}))
            cmd = [os.path.join(PKG, "bin", "vllm"), "serve", tmp.name, "--host", "127.0.0.1", "--port", str(port), "--served-model-name", served,
                   "--load-format", "dummy", "--seed", "1234", "--max-model-len", str(pin + pout + 16), "--max-num-seqs", str(max(B, 8)),
                   "--max-num-batched-tokens", str(max(pbatch * pin, 256)), "--num-gpu-blocks-override", str(nblk), "--no-enable-prefix-caching"]
            procs.append(subprocess.Popen(cmd, env=env, stdout=subprocess.DEVNULL, stderr=open(os.path.join(tmp.name, f"serve{g}.log"), "w")))
        for port in ports:
            if not wait_http(port, "/health", startup_timeout):
                raise RuntimeError("engine did not come up: " + open(os.path.join(tmp.name, "serve0.log")).read()[-500:])
        rport = free_port()
        procs.append(subprocess.Popen([os.path.join(PKG, "bin", "vllm-router"), "--host", "127.0.0.1", "--port", str(rport), "--service-discovery", "static",
                                       "--static-backends", ",".join(f"http://127.0.0.1:{p}" for p in ports), "--routing-logic", "roundrobin"],
                                      stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL))
        if not wait_http(rport, "/health", 30):
            raise RuntimeError("router did not come up")
        log(f"{gpus} engine(s) + router up")
        bodies = []
        for g in range(gpus):
            for p in bench.make_prompts(hf["vocab_size"], B, pin, seed=g):
                bodies.append(json.dumps({"model": served, "prompt": p, "max_tokens": pout, "temperature": 0, "ignore_eos": True,
                                          "stream": True}).encode())
        for _ in range(warm_jobs):
            asyncio.run(run_job(rport, bodies))
        tput, ttft, itl, bad = [], [], [], 0
        for _ in range(jobs):
            rec, wall = asyncio.run(run_job(rport, bodies))
            ntok = sum(len(c) for c in rec["chunks"])
            bad += sum(1 for st, c in zip(rec["status"], rec["chunks"]) if st != 200 or len(c) != pout) + len(rec["errors"])
            tput.append(ntok / wall)
            for sent, c in zip(rec["sent"], rec["chunks"]):
                if c:
                    ttft.append(c[0] - sent)
                    itl.extend(y - x for x, y in zip(c, c[1:]))
        return {"path": f"asyncio SSE client -> kge-router (round robin) -> {gpus} x `vllm serve` (kge-serve), one GPU each",
                "workload": f"{workload}: {name}, {B} concurrent streams per engine, {pin}-in/{pout}-out, greedy, ignore_eos",
                "gpus": gpus, "requests_per_job": len(bodies), "jobs": jobs, "warm_jobs": warm_jobs,
                "value": sum(tput) / len(tput), "unit": "output tokens/s", "per_job": [round(t, 1) for t in tput],
                "p50_ttft_ms": 1e3 * median(ttft) if ttft else None, "p50_itl_ms": 1e3 * median(itl) if itl else None,
                "bad_requests": bad,
                "definitions": "vLLM bench client: tokens / wall of the job; TTFT = first chunk - send; ITL = gaps between a request's chunks; medians"}
    finally:
        for p in procs:
            p.terminate()
        for p in procs:
            try:
                p.wait(timeout=20)
            except subprocess.TimeoutExpired:
                p.kill()
        tmp.cleanup()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--jobs", type=int, default=2)
    ap.add_argument("--warm-jobs", type=int, default=1)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    doc = measure(a.gpus, a.workload, a.jobs, a.warm_jobs, log=lambda m: sys.stderr.write(m + "\n"))
    s = json.dumps(doc)
    print(s)
    if a.out:
        open(a.out, "w").write(s + "\n")
