"""tools/http_bench.py - config 5 as deployed: SSE clients -> kge-router -> N `vllm serve` processes - exercised on the
CPU: the unmodified server binaries run on the mock engine library (tests/cpp/mock_kge.cc), two "replicas" behind the
router, and the client's accounting (tokens, TTFT, ITL, failures) is checked."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "kubernetes-gpu-cluster_b200")
sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.fixture(scope="module")
def mock_lib(tmp_path_factory):
    d = tmp_path_factory.mktemp("mocklib")
    cmd = ["g++", "-O1", "-std=c++17", "-shared", "-fPIC", f"-I{os.path.join(ROOT, 'include')}", f"-I{os.path.join(PKG, 'csrc')}",
           os.path.join(ROOT, "tests", "cpp", "mock_kge.cc"), os.path.join(PKG, "csrc", "scheduler.cc"), "-o", str(d / "libkge.so")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    return str(d)


def test_two_replicas_behind_the_router(mock_lib):
    import http_bench
    doc = http_bench.measure(2, "c2", jobs=2, warm_jobs=1, env_extra={"LD_LIBRARY_PATH": mock_lib},
                             shape=("llama-3.2-1b", 8, 40, 12, 8), startup_timeout=60)
    assert doc["bad_requests"] == 0, doc
    assert doc["requests_per_job"] == 16 and doc["gpus"] == 2
    assert doc["value"] > 0 and len(doc["per_job"]) == 2
    # every stream delivered its 12 tokens one chunk at a time: 11 gaps per request
    assert doc["p50_ttft_ms"] is not None and doc["p50_ttft_ms"] >= 0
    assert doc["p50_itl_ms"] is not None and doc["p50_itl_ms"] >= 0
