"""Host-side C++ of the product under AddressSanitizer + UndefinedBehaviorSanitizer: the scheduler / block pool replaying
every recorded vLLM trace (and an abort storm), and the tokenizer / NFC / stop-string code on hostile input.  The sources
are compiled here straight from the product tree (they have no CUDA in them).  CPU only."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "kubernetes-gpu-cluster_b200")
pytestmark = pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
SAN = ["-fsanitize=address,undefined,float-cast-overflow", "-fno-sanitize-recover=all", "-fno-omit-frame-pointer", "-g", "-O1", "-std=c++17"]
ENV = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1")


def build(tmp_path, name, sources, includes):
    exe = str(tmp_path / name)
    cmd = ["g++", *SAN, *[f"-I{i}" for i in includes], *sources, "-o", exe]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if p.returncode and "sanitize" in p.stderr and "cannot find" in p.stderr:
        pytest.skip("sanitizer runtime not installed")
    assert p.returncode == 0, p.stderr[-3000:]
    return exe


def test_scheduler_and_block_pool_under_sanitizers(tmp_path):
    exe = build(tmp_path, "sched_replay", [os.path.join(ROOT, "tests", "cpp", "sched_replay.cc"), os.path.join(PKG, "csrc", "scheduler.cc")],
                [os.path.join(PKG, "csrc")])
    for fixture, pc in (("sched_vllm.json", 0), ("sched_prefix_vllm.json", 1)):
        p = subprocess.run([exe, os.path.join(ROOT, "tests", "golden", fixture), str(pc)], capture_output=True, text=True, timeout=300, env=ENV)
        assert p.returncode == 0 and p.stdout.startswith("ok"), (fixture, p.stdout[-500:], p.stderr[-3000:])


def test_tokenizer_nfc_stop_strings_under_sanitizers(tmp_path):
    pytest.importorskip("tokenizers")
    from test_tokenizer_cpu import build as build_tok
    d = tmp_path / "tok"
    d.mkdir()
    build_tok("qwen", str(d), nfc=True)
    exe = build(tmp_path, "text_fuzz", [os.path.join(ROOT, "tests", "cpp", "text_fuzz.cc")], [os.path.join(PKG, "host"), os.path.join(PKG, "csrc")])
    p = subprocess.run([exe, str(d)], capture_output=True, text=True, timeout=300, env=ENV)
    assert p.returncode == 0 and p.stdout.startswith("ok"), (p.stdout[-500:], p.stderr[-3000:])
