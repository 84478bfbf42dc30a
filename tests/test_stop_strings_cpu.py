"""The server's per-request text stream (host/stop_strings.h: incremental detokenisation, stop strings, hold-back of
a possible stop-string prefix, min_tokens, include_stop_str_in_output, stop-token termination) against fixtures
recorded from vLLM's real FastIncrementalDetokenizer (tests/golden/stop_vllm.json, made by
tests/golden/make_stop_golden.py).  Driven through `vllm tokenize DIR`.  CPU only."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "kubernetes-gpu-cluster_b200", "bin", "vllm")
GOLD = os.path.join(ROOT, "tests", "golden", "stop_vllm.json")
pytestmark = pytest.mark.skipif(not os.path.exists(BIN), reason="server binary not built")


def test_text_stream_matches_vllm_detokenizer(tmp_path):
    g = json.load(open(GOLD))
    (tmp_path / "tokenizer.json").write_text(g["tokenizer_json"])
    cases = g["cases"]
    assert sum(c["matched"] >= 0 for c in cases) > 100 and sum(c["matched"] < 0 for c in cases) > 100
    inp = "".join(json.dumps({k: c[k] for k in ("ids", "stop", "include_stop_str_in_output", "min_tokens", "stop_terminated")}) + "\n"
                  for c in cases)
    p = subprocess.run([BIN, "tokenize", str(tmp_path)], input=inp, capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    outs = [json.loads(line) for line in p.stdout.splitlines()]
    assert len(outs) == len(cases)
    for c, o in zip(cases, outs):
        key = (c["stop"], c["include_stop_str_in_output"], c["min_tokens"], c["stop_terminated"])
        assert o["matched"] == c["matched"], key
        assert o["tokens_used"] == c["tokens_used"], key
        assert o["text"] == c["text"], key
        # what a streaming client has received in total, and that nothing beyond the final text ever went out
        assert "".join(o["deltas"]) == "".join(c["deltas"]) == c["text"], key
        # piece boundaries may differ while a multi-byte character is incomplete, never by more than the hold-back
        got, ref = "", ""
        hold = max(len(s) for s in c["stop"]) - 1
        for a, b in zip(o["deltas"], c["deltas"]):
            got += a
            ref += b
            assert c["text"].startswith(got) and abs(len(got) - len(ref)) <= max(hold, 1) + 1, key
