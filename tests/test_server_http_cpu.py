"""The HTTP server (the unmodified bin/vllm = kge-serve) on a machine without a GPU: tests/cpp/mock_kge.cc - test
infrastructure, a stand-in libkge.so that runs the REAL scheduler and "samples" with a fixed formula - is built into a
temporary directory and put first on LD_LIBRARY_PATH.  What is under test is everything above the C ABI: routes, request
parsing and validation, the OpenAI response shapes, SSE framing, tokenizer / chat rendering, stop strings and stop_reason,
echo, logprobs shapes, generation_config defaults, chat max_tokens default, n > 1, aborts on disconnect, /metrics.
(The same routes run against the real engine in tests/test_server_gpu.py.)  CPU only."""
import json
import os
import shutil
import socket
import subprocess
import time
import urllib.error
import urllib.request

import pytest

from oracle import model as om

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "kubernetes-gpu-cluster_b200")
BIN = os.path.join(PKG, "bin", "vllm")
pytestmark = pytest.mark.skipif(not os.path.exists(BIN) or shutil.which("g++") is None, reason="needs the server binary and g++")
VOCAB = om.CONFIGS["tiny-llama"].vocab_size


def mock_next(last, pos):
    return (7 * last + 3 * pos + 11) % VOCAB


def mock_generate(prompt_ids, n):
    seq = list(prompt_ids)
    for _ in range(n):
        seq.append(mock_next(seq[-1], len(seq)))
    return seq[len(prompt_ids):]


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.fixture(scope="module")
def mock_lib(tmp_path_factory):
    d = tmp_path_factory.mktemp("mocklib")
    cmd = ["g++", "-O1", "-std=c++17", "-shared", "-fPIC", f"-I{os.path.join(ROOT, 'include')}", f"-I{os.path.join(PKG, 'csrc')}",
           os.path.join(ROOT, "tests", "cpp", "mock_kge.cc"), os.path.join(PKG, "csrc", "scheduler.cc"), "-o", str(d / "libkge.so")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    return str(d)


def start(mock_lib, model_dir, log, *extra, env_extra=None, max_model_len=256):
    port = free_port()
    env = dict(os.environ, LD_LIBRARY_PATH=mock_lib, MOCK_KGE_LOG=str(log), **(env_extra or {}))
    mml = ["--max-model-len", str(max_model_len)] if max_model_len else []
    proc = subprocess.Popen([BIN, "serve", str(model_dir), "--host", "127.0.0.1", "--port", str(port), *mml,
                             "--served-model-name", "m", *extra], stderr=subprocess.PIPE, env=env)
    for _ in range(200):
        try:
            with urllib.request.urlopen(f"http://127.0.0.1:{port}/health", timeout=1) as r:
                if r.status == 200:
                    return proc, port
        except Exception:
            if proc.poll() is not None:
                break
            time.sleep(0.03)
    proc.kill()
    raise RuntimeError("server did not start: " + proc.stderr.read().decode()[-2000:])


def stop(proc):
    proc.terminate()
    try:
        proc.wait(timeout=10)
    except subprocess.TimeoutExpired:
        proc.kill()


def http(port, path, body=None, timeout=30):
    req = urllib.request.Request(f"http://127.0.0.1:{port}{path}", data=None if body is None else json.dumps(body).encode(),
                                 headers={"Content-Type": "application/json"})
    try:
        with urllib.request.urlopen(req, timeout=timeout) as r:
            return r.status, r.read().decode()
    except urllib.error.HTTPError as e:
        return e.code, e.read().decode()


def sse(port, path, body):
    req = urllib.request.Request(f"http://127.0.0.1:{port}{path}", data=json.dumps(body).encode(), headers={"Content-Type": "application/json"})
    with urllib.request.urlopen(req, timeout=30) as r:
        assert r.headers["Content-Type"].startswith("text/event-stream")
        raw = r.read().decode()
    events = [e[6:] for e in raw.split("\n\n") if e.startswith("data: ")]
    assert events[-1] == "[DONE]"
    return [json.loads(e) for e in events[:-1]]


@pytest.fixture(scope="module")
def srv(mock_lib, tmp_path_factory):
    """Server on a model directory WITH a Llama-3-style tokenizer; yields (port, tokenizer, log path)."""
    from test_tokenizer_cpu import build
    d = tmp_path_factory.mktemp("model")
    (d / "config.json").write_text(json.dumps(om.CONFIGS["tiny-llama"].to_hf()))
    tok = build("llama3", str(d))
    log = d / "requests.log"
    proc, port = start(mock_lib, d, log)
    yield port, tok, log
    stop(proc)


def last_logged(log):
    return json.loads(open(log).read().strip().splitlines()[-1])


def test_routes_and_token_id_completions(srv):
    port, tok, log = srv
    st, body = http(port, "/v1/models")
    m = json.loads(body)["data"][0]
    assert st == 200 and m["id"] == "m" and m["object"] == "model" and m["owned_by"] == "vllm" and m["max_model_len"] == 256
    assert http(port, "/health")[0] == 200 and http(port, "/ping")[0] == 200
    st, body = http(port, "/v1/completions", {"model": "m", "prompt": [5, 6, 7], "max_tokens": 5, "temperature": 0})
    r = json.loads(body)
    assert st == 200 and r["object"] == "text_completion" and r["id"].startswith("cmpl-") and r["model"] == "m"
    c = r["choices"][0]
    assert c["token_ids"] == mock_generate([5, 6, 7], 5) and c["finish_reason"] == "length" and c["stop_reason"] is None and c["logprobs"] is None
    assert c["text"] == tok.decode(c["token_ids"], skip_special_tokens=True)
    assert r["usage"] == {"prompt_tokens": 3, "total_tokens": 8, "completion_tokens": 5}
    lg = last_logged(log)
    assert (lg["max_tokens"], lg["temperature"], lg["top_k"], lg["top_p"], lg["ignore_eos"]) == (5, 0, 0, 1, 0)
    # defaults: 16 tokens, temperature 1.0 (vllm:sampling_params.py)
    st, body = http(port, "/v1/completions", {"prompt": [5, 6, 7]})
    assert len(json.loads(body)["choices"][0]["token_ids"]) == 16 and last_logged(log)["temperature"] == 1
    # batch of prompts and n > 1: choices in order, indices 0..
    st, body = http(port, "/v1/completions", {"prompt": [[5, 6], [9]], "n": 2, "max_tokens": 2, "temperature": 0, "seed": 7})
    ch = json.loads(body)["choices"]
    assert [c["index"] for c in ch] == [0, 1, 2, 3]
    assert [c["token_ids"] for c in ch] == [mock_generate([5, 6], 2)] * 2 + [mock_generate([9], 2)] * 2
    st, m = http(port, "/metrics")
    for name in ("vllm:num_requests_running", "vllm:num_requests_waiting", "vllm:kv_cache_usage_perc", "vllm:prompt_tokens_total",
                 "vllm:generation_tokens_total", "vllm:request_success_total", "vllm:prefix_cache_queries_total", "vllm:prefix_cache_hits_total"):
        assert name in m, name


def test_errors(srv):
    port, tok, log = srv
    for body, code, needle in [({"prompt": [5, 10 ** 7]}, 400, "out of vocabulary"), ({"prompt": [5, 1e30]}, 400, "out of vocabulary"),
                               ({"prompt": []}, 400, "prompt"), ({"prompt": [5], "max_tokens": 0}, 400, "max_tokens"),
                               ({"prompt": [5], "top_p": 0}, 400, "top_p"), ({"prompt": [5], "temperature": -1}, 400, "temperature"),
                               ({"prompt": [5], "n": 1000}, 400, "n must"), ({"prompt": [5], "logprobs": 50}, 400, "logprobs"),
                               ({"prompt": [5], "stop": 3}, 400, "stop"), ({"prompt": list(range(3, 300)), "max_tokens": 4}, 400, "max_model_len"),
                               ({"model": "other", "prompt": [5]}, 404, "does not exist"), ({"max_tokens": 3}, 400, "prompt")]:
        st, out = http(port, "/v1/completions", body)
        e = json.loads(out)["error"]
        assert st == code and e["code"] == code and needle in e["message"] and e["type"] in ("BadRequestError", "NotFoundError"), (body, out)
    req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/completions", data=b"{not json", headers={"Content-Type": "application/json"})
    with pytest.raises(urllib.error.HTTPError) as ei:
        urllib.request.urlopen(req, timeout=10)
    assert ei.value.code == 400 and "invalid JSON" in ei.value.read().decode()
    assert http(port, "/nope")[0] == 404
    st, out = http(port, "/v1/chat/completions", {"messages": [{"role": "user", "content": "x"}], "echo": True})
    assert st == 400 and "echo" in out


def test_text_prompts_chat_and_defaults(srv):
    port, tok, log = srv
    text = "The GPU cluster's nodes"
    ids = tok.encode(text).ids
    st, body = http(port, "/v1/completions", {"prompt": text, "max_tokens": 4, "temperature": 0})
    c = json.loads(body)
    assert st == 200 and c["usage"]["prompt_tokens"] == len(ids) and c["choices"][0]["token_ids"] == mock_generate(ids, 4)
    st, body = http(port, "/v1/completions", {"prompt": text, "max_tokens": 3, "temperature": 0, "echo": True})
    assert json.loads(body)["choices"][0]["text"] == text + tok.decode(mock_generate(ids, 3), skip_special_tokens=True)
    msgs = [{"role": "system", "content": "Be brief."}, {"role": "user", "content": [{"type": "text", "text": "Hi "}, {"type": "text", "text": "there"}]}]
    rendered = ("<|begin_of_text|><|start_header_id|>system<|end_header_id|>\n\nBe brief.<|eot_id|><|start_header_id|>user<|end_header_id|>\n\n"
                "Hi there<|eot_id|><|start_header_id|>assistant<|end_header_id|>\n\n")
    pids = tok.encode(rendered, add_special_tokens=False).ids
    st, body = http(port, "/v1/chat/completions", {"messages": msgs, "max_completion_tokens": 5, "temperature": 0})
    r = json.loads(body)
    assert st == 200 and r["object"] == "chat.completion" and r["id"].startswith("chatcmpl-") and r["usage"]["prompt_tokens"] == len(pids)
    ch = r["choices"][0]
    assert ch["message"]["role"] == "assistant" and ch["token_ids"] == mock_generate(pids, 5)
    assert ch["message"]["content"] == tok.decode(ch["token_ids"], skip_special_tokens=True)
    # a chat request without a token limit may use the rest of the context (vLLM: max_model_len - prompt length)
    st, body = http(port, "/v1/chat/completions", {"messages": msgs, "temperature": 0})
    assert last_logged(log)["max_tokens"] == 256 - len(pids)
    assert json.loads(body)["choices"][0]["finish_reason"] == "length"


def test_streaming_sse(srv):
    port, tok, log = srv
    base = {"prompt": [5, 6, 7], "max_tokens": 6, "temperature": 0}
    ev = sse(port, "/v1/completions", dict(base, stream=True, stream_options={"include_usage": True}))
    chunks, usage = ev[:-1], ev[-1]
    want = mock_generate([5, 6, 7], 6)
    assert [c["choices"][0]["token_ids"][0] for c in chunks] == want and all(c["object"] == "text_completion" for c in chunks)
    assert "".join(c["choices"][0]["text"] for c in chunks) == tok.decode(want, skip_special_tokens=True)
    assert [c["choices"][0]["finish_reason"] for c in chunks] == [None] * 5 + ["length"]
    assert usage["choices"] == [] and usage["usage"] == {"prompt_tokens": 3, "total_tokens": 9, "completion_tokens": 6}
    ev = sse(port, "/v1/chat/completions", {"messages": [{"role": "user", "content": "Hi"}], "max_tokens": 3, "temperature": 0, "stream": True})
    assert len(ev) == 3 and all(c["object"] == "chat.completion.chunk" and c["choices"][0]["delta"]["role"] == "assistant" for c in ev)
    ev = sse(port, "/v1/completions", dict(base, stream=True, n=2))
    assert sorted(c["choices"][0]["index"] for c in ev) == [0] * 6 + [1] * 6


def test_stop_strings_stop_ids_and_logprobs(srv):
    port, tok, log = srv
    base = {"prompt": "The GPU cluster's nodes", "max_tokens": 24, "temperature": 0, "ignore_eos": True}
    full = json.loads(http(port, "/v1/completions", base)[1])["choices"][0]
    text, ids = full["text"], full["token_ids"]
    assert len(ids) == 24 and len(text) >= 8
    mid = len(text) // 2
    stop_s = text[mid:mid + 3]
    cut = text.index(stop_s)
    c0 = json.loads(http(port, "/v1/completions", dict(base, stop=[stop_s]))[1])["choices"][0]
    assert c0["text"] == text[:cut] and c0["finish_reason"] == "stop" and c0["stop_reason"] == stop_s and c0["token_ids"] == ids[:len(c0["token_ids"])]
    c0 = json.loads(http(port, "/v1/completions", dict(base, stop=stop_s, include_stop_str_in_output=True))[1])["choices"][0]
    assert c0["text"] == text[:cut + len(stop_s)] and c0["stop_reason"] == stop_s
    ev = sse(port, "/v1/completions", dict(base, stop=[stop_s], stream=True))
    assert "".join(c["choices"][0]["text"] for c in ev) == text[:cut]
    assert ev[-1]["choices"][0]["finish_reason"] == "stop" and ev[-1]["choices"][0]["stop_reason"] == stop_s
    sid = ids[5]
    k = ids.index(sid)
    c0 = json.loads(http(port, "/v1/completions", dict(base, stop_token_ids=[sid]))[1])["choices"][0]
    assert c0["token_ids"] == ids[:k + 1] and c0["finish_reason"] == "stop" and c0["stop_reason"] == sid
    assert c0["text"] == tok.decode(ids[:k], skip_special_tokens=True)           # the stop token's text is left out
    # min_tokens keeps both kinds of stop from firing early
    c0 = json.loads(http(port, "/v1/completions", dict(base, stop_token_ids=[sid], min_tokens=k + 2, max_tokens=k + 4))[1])["choices"][0]
    assert len(c0["token_ids"]) > k + 1
    # logprobs: the completions and the chat shape, streamed too (values are the mock's: -0.5 for the sampled token)
    lp = json.loads(http(port, "/v1/completions", dict(base, max_tokens=4, logprobs=2))[1])["choices"][0]["logprobs"]
    assert lp["token_logprobs"] == [-0.5] * 4 and len(lp["tokens"]) == len(lp["top_logprobs"]) == 4
    offs, acc = [], 0
    for t in lp["tokens"]:
        offs.append(acc)
        acc += len(t)
    assert lp["text_offset"] == offs and all(1 <= len(t) <= 3 for t in lp["top_logprobs"])
    assert all(len(t) == 1 for t in json.loads(http(port, "/v1/completions", dict(base, max_tokens=3, logprobs=0))[1])["choices"][0]["logprobs"]["top_logprobs"])
    content = json.loads(http(port, "/v1/chat/completions", {"messages": [{"role": "user", "content": "Hi"}], "max_tokens": 3, "temperature": 0,
                                                             "logprobs": True, "top_logprobs": 2})[1])["choices"][0]["logprobs"]["content"]
    assert len(content) == 3
    for e in content:
        assert e["logprob"] == -0.5 and bytes(e["bytes"]).decode("utf-8", "replace") == e["token"] and len(e["top_logprobs"]) == 2
        assert e["top_logprobs"][0]["token"] == e["token"]
    ev = sse(port, "/v1/completions", dict(base, max_tokens=4, logprobs=1, stream=True))
    assert [c["choices"][0]["logprobs"]["text_offset"][0] for c in ev] == offs
    assert json.loads(http(port, "/v1/completions", dict(base, max_tokens=2))[1])["choices"][0]["logprobs"] is None


def test_generation_config_defaults_and_extra_eos(mock_lib, tmp_path):
    """generation_config.json: sampling defaults for silent requests, and extra end-of-sequence ids that end a request like
    a stop token (Llama-3.1-Instruct's <|eot_id|>), unless ignore_eos."""
    prompt = [5, 6, 7]
    want = mock_generate(prompt, 12)
    eos2 = want[4]
    cfg = dict(om.CONFIGS["tiny-llama"].to_hf(), eos_token_id=[want[9], eos2])
    (tmp_path / "config.json").write_text(json.dumps(cfg))
    (tmp_path / "generation_config.json").write_text(json.dumps({"temperature": 0.6, "top_p": 0.9, "top_k": 20, "eos_token_id": [want[9], eos2, want[7]]}))
    log = tmp_path / "log"
    proc, port = start(mock_lib, tmp_path, log)
    try:
        c0 = json.loads(http(port, "/v1/completions", {"prompt": prompt, "max_tokens": 12})[1])["choices"][0]
        lg = last_logged(log)
        assert (round(lg["temperature"], 3), round(lg["top_p"], 3), lg["top_k"]) == (0.6, 0.9, 20)
        assert sorted(lg["stop_token_ids"]) == sorted([eos2, want[7]])
        assert c0["token_ids"] == want[:5] and c0["finish_reason"] == "stop" and c0["stop_reason"] == eos2
        http(port, "/v1/completions", {"prompt": prompt, "max_tokens": 2, "temperature": 0, "top_p": 0.5, "top_k": 3})
        lg = last_logged(log)
        assert (lg["temperature"], lg["top_p"], lg["top_k"]) == (0, 0.5, 3)           # what the request says wins
        c0 = json.loads(http(port, "/v1/completions", {"prompt": prompt, "max_tokens": 12, "ignore_eos": True})[1])["choices"][0]
        assert c0["token_ids"] == want and last_logged(log)["stop_token_ids"] == []
    finally:
        stop(proc)
    proc, port = start(mock_lib, tmp_path, log, "--generation-config", "vllm")
    try:
        http(port, "/v1/completions", {"prompt": prompt, "max_tokens": 2})
        lg = last_logged(log)
        assert (lg["temperature"], lg["top_p"], lg["top_k"]) == (1, 1, 0) and sorted(lg["stop_token_ids"]) == sorted([eos2, want[7]])
    finally:
        stop(proc)


def test_without_a_tokenizer_and_client_disconnect(mock_lib, tmp_path):
    (tmp_path / "config.json").write_text(json.dumps(om.CONFIGS["tiny-llama"].to_hf()))
    log = tmp_path / "log"
    proc, port = start(mock_lib, tmp_path, log, env_extra={"MOCK_KGE_STEP_US": "20000"})
    try:
        st, out = http(port, "/v1/completions", {"prompt": "hello", "max_tokens": 2})
        assert st == 400 and "tokenizer" in out
        st, out = http(port, "/v1/completions", {"prompt": [5], "max_tokens": 2, "stop": ["x"]})
        assert st == 400 and "tokenizer" in out
        st, out = http(port, "/v1/chat/completions", {"messages": [{"role": "user", "content": "hi"}]})
        assert st == 400 and "chat template" in out
        c0 = json.loads(http(port, "/v1/completions", {"prompt": [5, 6], "max_tokens": 3, "temperature": 0, "logprobs": 1})[1])["choices"][0]
        assert c0["text"] == "" and c0["logprobs"]["tokens"] == [f"token_id:{t}" for t in c0["token_ids"]]
        # a streaming client that goes away: the request is aborted and its KV blocks are released
        s = socket.create_connection(("127.0.0.1", port), timeout=5)
        body = json.dumps({"prompt": [5, 6, 7], "max_tokens": 200, "temperature": 0, "ignore_eos": True, "stream": True}).encode()
        s.sendall(b"POST /v1/completions HTTP/1.1\r\nHost: x\r\nContent-Type: application/json\r\nContent-Length: %d\r\n\r\n%s" % (len(body), body))
        assert b"200 OK" in s.recv(4096)
        s.close()
        for _ in range(100):
            m = http(port, "/metrics")[1]
            running = float([ln for ln in m.splitlines() if ln.startswith("vllm:num_requests_running")][0].split()[-1])
            usage = float([ln for ln in m.splitlines() if ln.startswith("vllm:kv_cache_usage_perc")][0].split()[-1])
            if running == 0 and usage == 0:
                break
            time.sleep(0.05)
        assert running == 0 and usage == 0
        assert json.loads(http(port, "/v1/completions", {"prompt": [5, 6], "max_tokens": 2, "temperature": 0})[1])["choices"][0]["token_ids"] == mock_generate([5, 6], 2)
    finally:
        stop(proc)


def test_server_survives_garbage_under_sanitizers(mock_lib, tmp_path):
    """host/server.cc compiled with ASan + UBSan against the mock library: malformed, truncated, oversized and hostile
    requests are answered or dropped, nothing crashes or trips a sanitizer, and well-formed traffic keeps flowing."""
    import random
    exe = str(tmp_path / "kge-serve-asan")
    p = subprocess.run(["g++", "-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-g", "-O1", "-std=c++17",
                        f"-I{os.path.join(ROOT, 'include')}", f"-I{os.path.join(PKG, 'csrc')}", f"-I{os.path.join(PKG, 'host')}",
                        os.path.join(PKG, "host", "server.cc"), "-o", exe, f"-L{mock_lib}", "-lkge", "-lpthread", "-ldl"],
                       capture_output=True, text=True, timeout=900)
    if p.returncode and "cannot find" in p.stderr:
        pytest.skip("sanitizer runtime not installed")
    assert p.returncode == 0, p.stderr[-3000:]
    from test_tokenizer_cpu import build
    d = tmp_path / "model"
    d.mkdir()
    (d / "config.json").write_text(json.dumps(om.CONFIGS["tiny-llama"].to_hf()))
    build("llama3", str(d))
    port = free_port()
    env = dict(os.environ, LD_LIBRARY_PATH=mock_lib, ASAN_OPTIONS="detect_leaks=0")
    proc = subprocess.Popen([exe, "serve", str(d), "--host", "127.0.0.1", "--port", str(port), "--max-model-len", "256", "--served-model-name", "m"],
                            stderr=subprocess.PIPE, env=env)
    try:
        for _ in range(300):
            try:
                if http(port, "/health", timeout=1)[0] == 200:
                    break
            except Exception:
                time.sleep(0.05)
        rng = random.Random(0)
        good = json.dumps({"prompt": "Hello there", "max_tokens": 4, "temperature": 0, "stop": ["zz"], "logprobs": 2}).encode()
        payloads = [b"", b"\r\n\r\n", b"GET\r\n\r\n", b"POST /v1/completions HTTP/1.1\r\nContent-Length: 99999999999\r\n\r\n{}",
                    b"POST /v1/completions HTTP/1.1\r\nContent-Length: 50\r\n\r\n{}", b"GET /" + b"a" * 70000 + b" HTTP/1.1\r\n\r\n",
                    b"X" * (2 << 20)]
        payloads += [bytes(rng.randrange(256) for _ in range(rng.randrange(1, 300))) for _ in range(30)]
        bodies = [b"[" * 5000, b'{"prompt": 1e999}', b'{"prompt":[1e30,-5,2.5],"max_tokens":1e40}', b'{"prompt":"\\ud800 \xff\xfe","max_tokens":2}',
                  b'{"prompt":"x","stop":["",""],"max_tokens":3,"n":-5}', b'{"messages":[1,2,{"role":5,"content":[{"type":"text"}]}]}',
                  b'{"prompt":[[5],[6],"a"]}', b'{"prompt":"x","logprobs":-7,"top_k":-9,"min_tokens":1e9,"seed":-1}']
        for _ in range(60):
            b = bytearray(good)
            for _ in range(rng.randrange(1, 4)):
                b[rng.randrange(len(b))] = rng.randrange(256)
            bodies.append(bytes(b))
        for path in (b"/v1/completions", b"/v1/chat/completions"):
            payloads += [b"POST " + path + b" HTTP/1.1\r\nContent-Length: %d\r\n\r\n%s" % (len(x), x) for x in bodies]
        for pl in payloads:
            s = socket.create_connection(("127.0.0.1", port), timeout=5)
            try:
                s.sendall(pl)
                s.shutdown(socket.SHUT_WR)
                s.settimeout(5)
                while s.recv(65536):
                    pass
            except OSError:
                pass
            finally:
                s.close()
        assert proc.poll() is None, proc.stderr.read().decode()[-3000:]
        st, body = http(port, "/v1/completions", {"prompt": [5, 6], "max_tokens": 3, "temperature": 0})
        assert st == 200 and json.loads(body)["choices"][0]["token_ids"] == mock_generate([5, 6], 3)
    finally:
        stop(proc)
        err = proc.stderr.read().decode()
    assert "AddressSanitizer" not in err and "runtime error" not in err, err[-3000:]


def test_tokenize_detokenize_and_api_key(mock_lib, tmp_path):
    from test_tokenizer_cpu import build
    (tmp_path / "config.json").write_text(json.dumps(om.CONFIGS["tiny-llama"].to_hf()))
    tok = build("llama3", str(tmp_path))
    proc, port = start(mock_lib, tmp_path, tmp_path / "log", "--api-key", "s3cret")
    try:
        text = "The GPU cluster's nodes aren't ready."
        st, out = http(port, "/tokenize", {"prompt": text})                      # outside /v1: no key needed, as in vLLM
        r = json.loads(out)
        assert st == 200 and r["tokens"] == tok.encode(text).ids and r["count"] == len(r["tokens"]) and r["max_model_len"] == 256
        assert json.loads(http(port, "/tokenize", {"prompt": text, "add_special_tokens": False})[1])["tokens"] == tok.encode(text, add_special_tokens=False).ids
        st, out = http(port, "/detokenize", {"tokens": r["tokens"]})
        assert st == 200 and json.loads(out)["prompt"] == text
        msgs = [{"role": "user", "content": "Hi"}]
        rendered = "<|begin_of_text|><|start_header_id|>user<|end_header_id|>\n\nHi<|eot_id|><|start_header_id|>assistant<|end_header_id|>\n\n"
        assert json.loads(http(port, "/tokenize", {"messages": msgs})[1])["tokens"] == tok.encode(rendered, add_special_tokens=False).ids
        assert http(port, "/tokenize", {"prompt": 5})[0] == 400 and http(port, "/detokenize", {"tokens": "x"})[0] == 400
        # /v1 routes need the key
        st, out = http(port, "/v1/completions", {"prompt": [5], "max_tokens": 1})
        assert st == 401 and json.loads(out) == {"error": "Unauthorized"}
        assert http(port, "/v1/models")[0] == 401 and http(port, "/health")[0] == 200 and http(port, "/metrics")[0] == 200
        for key, code in (("s3cret", 200), ("wrong", 401)):
            req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/completions", data=json.dumps({"prompt": [5], "max_tokens": 1}).encode(),
                                         headers={"Content-Type": "application/json", "Authorization": f"Bearer {key}"})
            try:
                with urllib.request.urlopen(req, timeout=10) as resp:
                    got = resp.status
            except urllib.error.HTTPError as e:
                got = e.code
            assert got == code
    finally:
        stop(proc)
    proc, port = start(mock_lib, tmp_path, tmp_path / "log", env_extra={"VLLM_API_KEY": "envkey"})
    try:
        assert http(port, "/v1/models")[0] == 401
    finally:
        stop(proc)


def test_latency_histograms_in_metrics(mock_lib, tmp_path):
    """vllm:time_to_first_token_seconds, inter_token_latency_seconds (and its v0.8 name), e2e_request_latency_seconds:
    Prometheus histograms with vLLM's names; counts follow the traffic."""
    (tmp_path / "config.json").write_text(json.dumps(om.CONFIGS["tiny-llama"].to_hf()))
    proc, port = start(mock_lib, tmp_path, tmp_path / "log", env_extra={"MOCK_KGE_STEP_US": "2000"})
    try:
        for n in (3, 5, 1):
            assert http(port, "/v1/completions", {"prompt": [5, 6, 7], "max_tokens": n, "temperature": 0, "ignore_eos": True})[0] == 200
        m = http(port, "/metrics")[1]

        def series(name):
            rows = [ln for ln in m.splitlines() if ln.startswith(name)]
            buckets = [(ln.split('le="')[1].split('"')[0], int(ln.split()[-1])) for ln in rows if "_bucket" in ln]
            count = int([ln for ln in rows if ln.startswith(name + "_count")][0].split()[-1])
            total = float([ln for ln in rows if ln.startswith(name + "_sum")][0].split()[-1])
            return buckets, count, total
        for name, want in (("vllm:time_to_first_token_seconds", 3), ("vllm:inter_token_latency_seconds", 6),
                           ("vllm:time_per_output_token_seconds", 6), ("vllm:e2e_request_latency_seconds", 3)):
            assert f"# TYPE {name} histogram" in m
            buckets, count, total = series(name)
            assert count == want and buckets[-1] == ("+Inf", want) and total > 0
            assert all(a[1] <= b[1] for a, b in zip(buckets, buckets[1:]))
        b, c, total = series("vllm:inter_token_latency_seconds")
        assert 0.002 * 6 <= total < 1.0                                   # six gaps of one 2 ms mock step each
    finally:
        stop(proc)


def test_concurrent_clients_under_thread_sanitizer(mock_lib, tmp_path):
    """host/server.cc built with ThreadSanitizer: 24 clients at once - blocking, streaming, some disconnecting mid-stream,
    /metrics and /health polled meanwhile.  Every complete answer is right and TSan reports no data race."""
    import threading
    exe = str(tmp_path / "kge-serve-tsan")
    p = subprocess.run(["g++", "-fsanitize=thread", "-g", "-O1", "-std=c++17", f"-I{os.path.join(ROOT, 'include')}", f"-I{os.path.join(PKG, 'csrc')}",
                        f"-I{os.path.join(PKG, 'host')}", os.path.join(PKG, "host", "server.cc"), "-o", exe, f"-L{mock_lib}", "-lkge", "-lpthread", "-ldl"],
                       capture_output=True, text=True, timeout=900)
    if p.returncode and "cannot find" in p.stderr:
        pytest.skip("sanitizer runtime not installed")
    assert p.returncode == 0, p.stderr[-3000:]
    (tmp_path / "config.json").write_text(json.dumps(om.CONFIGS["tiny-llama"].to_hf()))
    port = free_port()
    env = dict(os.environ, LD_LIBRARY_PATH=mock_lib, MOCK_KGE_STEP_US="500", TSAN_OPTIONS="halt_on_error=0")
    proc = subprocess.Popen([exe, "serve", str(tmp_path), "--host", "127.0.0.1", "--port", str(port), "--max-model-len", "256", "--served-model-name", "m",
                             "--max-num-seqs", "8"], stderr=subprocess.PIPE, env=env)
    errors = []
    try:
        for _ in range(400):
            try:
                if http(port, "/health", timeout=1)[0] == 200:
                    break
            except Exception:
                time.sleep(0.05)

        def client(i):
            try:
                prompt = [5 + i, 6, 7 + i % 3]
                n = 5 + i % 7
                if i % 3 == 0:
                    ev = sse(port, "/v1/completions", {"prompt": prompt, "max_tokens": n, "temperature": 0, "ignore_eos": True, "stream": True})
                    got = [c["choices"][0]["token_ids"][0] for c in ev]
                elif i % 3 == 1:
                    got = json.loads(http(port, "/v1/completions", {"prompt": prompt, "max_tokens": n, "temperature": 0, "ignore_eos": True})[1])["choices"][0]["token_ids"]
                else:                                    # goes away mid-stream
                    s = socket.create_connection(("127.0.0.1", port), timeout=5)
                    body = json.dumps({"prompt": prompt, "max_tokens": 150, "temperature": 0, "ignore_eos": True, "stream": True}).encode()
                    s.sendall(b"POST /v1/completions HTTP/1.1\r\nContent-Length: %d\r\n\r\n%s" % (len(body), body))
                    s.recv(512)
                    s.close()
                    return
                if got != mock_generate(prompt, n):
                    errors.append((i, got))
            except Exception as ex:
                errors.append((i, repr(ex)))

        def poller():
            for _ in range(40):
                http(port, "/metrics")
                http(port, "/health")
        threads = [threading.Thread(target=client, args=(i,)) for i in range(24)] + [threading.Thread(target=poller)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=120)
        assert not errors, errors[:3]
        for _ in range(200):                              # the abandoned requests are gone
            m = http(port, "/metrics")[1]
            if float([ln for ln in m.splitlines() if ln.startswith("vllm:num_requests_running")][0].split()[-1]) == 0:
                break
            time.sleep(0.05)
        assert proc.poll() is None
    finally:
        stop(proc)
        err = proc.stderr.read().decode()
    assert "WARNING: ThreadSanitizer" not in err, err[-4000:]


def test_replicas_behind_the_router(mock_lib, tmp_path):
    """BASELINE config 5 in miniature, without GPUs: two server processes (mock engines) behind kge-router - requests
    alternate between the replicas, SSE passes through, /v1/models is merged."""
    router = os.path.join(PKG, "bin", "kge-router")
    if not os.path.exists(router):
        pytest.skip("router binary not built")
    (tmp_path / "config.json").write_text(json.dumps(om.CONFIGS["tiny-llama"].to_hf()))
    servers = [start(mock_lib, tmp_path, tmp_path / f"log{i}") for i in range(2)]
    rport = free_port()
    rproc = subprocess.Popen([router, "--host", "127.0.0.1", "--port", str(rport), "--service-discovery", "static", "--static-backends",
                              ",".join(f"http://127.0.0.1:{p}" for _, p in servers), "--static-models", "m,m"], stderr=subprocess.PIPE)
    try:
        for _ in range(100):
            try:
                if http(rport, "/health", timeout=1)[0] == 200:
                    break
            except Exception:
                time.sleep(0.05)
        for i in range(6):
            st, body = http(rport, "/v1/completions", {"model": "m", "prompt": [5, 6, 7 + i], "max_tokens": 3, "temperature": 0})
            assert st == 200 and json.loads(body)["choices"][0]["token_ids"] == mock_generate([5, 6, 7 + i], 3)
        ev = sse(rport, "/v1/completions", {"model": "m", "prompt": [9, 9], "max_tokens": 4, "temperature": 0, "stream": True})
        assert [c["choices"][0]["token_ids"][0] for c in ev] == mock_generate([9, 9], 4)
        assert [m["id"] for m in json.loads(http(rport, "/v1/models")[1])["data"]] == ["m"]
        served = []
        for _, p in servers:
            m = http(p, "/metrics")[1]
            served.append(sum(float(ln.split()[-1]) for ln in m.splitlines() if ln.startswith("vllm:request_success_total")))
        assert sorted(served) == [3, 4]                       # 7 requests, round-robin over two replicas
    finally:
        stop(rproc)
        for proc, _ in servers:
            stop(proc)


def test_context_length_comes_from_the_engine(mock_lib, tmp_path):
    cfg = dict(om.CONFIGS["tiny-llama"].to_hf(), max_position_embeddings=4096)
    (tmp_path / "config.json").write_text(json.dumps(cfg))
    proc, port = start(mock_lib, tmp_path, tmp_path / "log", max_model_len=None)
    try:
        assert json.loads(http(port, "/v1/models")[1])["data"][0]["max_model_len"] == 4096
        http(port, "/v1/chat/completions", {"prompt_token_ids": [5, 6, 7], "temperature": 0, "max_tokens": 2})
        st, out = http(port, "/v1/completions", {"prompt": [5] * 5000, "max_tokens": 1})
        assert st == 400 and "max_model_len" in out
    finally:
        stop(proc)
