"""The drop-in server end to end on a GPU: `vllm serve <dir>` (kge-serve), OpenAI-compatible HTTP,
token-id prompts (no tokenizer offline), greedy ids equal to the in-process engine's."""
import json
import math
import os
import socket
import subprocess
import time
import urllib.error
import urllib.request

import pytest

from oracle import model as om

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "kubernetes-gpu-cluster_b200", "bin", "vllm")


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def http(port, path, body=None, timeout=60):
    req = urllib.request.Request(f"http://127.0.0.1:{port}{path}", data=None if body is None else json.dumps(body).encode(),
                                 headers={"Content-Type": "application/json"})
    try:
        with urllib.request.urlopen(req, timeout=timeout) as r:
            return r.status, r.read().decode()
    except urllib.error.HTTPError as e:
        return e.code, e.read().decode()


def start_server(tmp_path_factory, with_tokenizer):
    assert os.path.exists(BIN), "build the server: make -C kubernetes-gpu-cluster_b200"
    cfg = om.CONFIGS["tiny-llama"]
    w = om.make_weights(cfg, seed=0)
    d = str(tmp_path_factory.mktemp("srv-model"))
    om.save_model_dir(d, cfg, w)
    if with_tokenizer:
        from test_tokenizer_cpu import build
        build("llama3", d)            # a 700-entry Llama-3-style tokenizer.json (vocab fits the 1024-row embedding)
    port = free_port()
    proc = subprocess.Popen([BIN, "serve", d, "--host", "127.0.0.1", "--port", str(port), "--max-model-len", "256",
                             "--max-num-seqs", "8", "--max-num-batched-tokens", "128", "--num-gpu-blocks-override", "128",
                             "--served-model-name", "tiny", "--enforce-eager", "--trust-remote-code"],
                            stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    t0 = time.time()
    while True:
        try:
            if http(port, "/health", timeout=2)[0] == 200:
                break
        except Exception:
            pass
        assert proc.poll() is None, proc.stderr.read()[-2000:]
        assert time.time() - t0 < 120, "server did not come up"
        time.sleep(0.2)
    return proc, (port, cfg, w, d)


def stop_server(proc):
    proc.terminate()
    try:
        proc.wait(timeout=20)
    except subprocess.TimeoutExpired:
        proc.kill()


@pytest.fixture(scope="module")
def server(tmp_path_factory):
    proc, info = start_server(tmp_path_factory, with_tokenizer=False)
    yield info
    stop_server(proc)


@pytest.fixture(scope="module")
def server_tok(tmp_path_factory):
    proc, info = start_server(tmp_path_factory, with_tokenizer=True)
    yield info
    stop_server(proc)


def test_models_health_metrics(server):
    port = server[0]
    st, body = http(port, "/v1/models")
    m = json.loads(body)
    assert st == 200 and m["object"] == "list" and m["data"][0]["id"] == "tiny" and m["data"][0]["owned_by"] == "vllm"
    assert m["data"][0]["max_model_len"] == 256
    assert http(port, "/health") == (200, "")
    st, txt = http(port, "/metrics")
    assert st == 200 and "vllm:num_requests_running" in txt and "vllm:kv_cache_usage_perc" in txt
    assert http(port, "/nope")[0] == 404


def test_completions_match_in_process_engine(server, kge):
    port, cfg, w, d = server
    prompts = om.make_prompts(cfg, 3, 20, seed=11)
    llm = kge.LLM(d, max_model_len=256, max_num_seqs=8, max_num_batched_tokens=128, num_gpu_blocks_override=128)
    ref = [o.outputs[0].token_ids for o in llm.generate(prompts, kge.SamplingParams(max_tokens=10, temperature=0.0, ignore_eos=True))]
    llm.close()
    st, body = http(port, "/v1/completions", {"model": "tiny", "prompt": prompts, "max_tokens": 10, "temperature": 0,
                                               "ignore_eos": True})
    assert st == 200, body
    r = json.loads(body)
    assert r["object"] == "text_completion" and r["id"].startswith("cmpl-") and r["model"] == "tiny"
    assert [c["index"] for c in r["choices"]] == [0, 1, 2]
    assert [c["token_ids"] for c in r["choices"]] == ref
    assert all(c["finish_reason"] == "length" for c in r["choices"])
    assert r["usage"] == {"prompt_tokens": 60, "total_tokens": 90, "completion_tokens": 30}
    # single [int] prompt + stop_token_ids
    st, body = http(port, "/v1/completions", {"prompt": prompts[0], "max_tokens": 10, "temperature": 0, "ignore_eos": True,
                                               "stop_token_ids": [ref[0][3]]})
    c = json.loads(body)["choices"][0]
    first = ref[0].index(ref[0][3])
    assert c["finish_reason"] == "stop" and c["token_ids"] == ref[0][:first + 1]


def test_streaming_sse(server, kge):
    port, cfg, w, d = server
    prompt = om.make_prompts(cfg, 1, 12, seed=5)[0]
    req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/completions",
                                 data=json.dumps({"prompt": prompt, "max_tokens": 6, "temperature": 0, "ignore_eos": True,
                                                  "stream": True, "stream_options": {"include_usage": True}}).encode(),
                                 headers={"Content-Type": "application/json"})
    with urllib.request.urlopen(req, timeout=60) as r:
        assert r.headers["Content-Type"].startswith("text/event-stream")
        raw = r.read().decode()
    events = [e for e in raw.split("\n\n") if e.strip()]
    assert all(e.startswith("data: ") for e in events) and events[-1] == "data: [DONE]"
    chunks = [json.loads(e[6:]) for e in events[:-1]]
    toks = [t for c in chunks if c["choices"] for t in c["choices"][0]["token_ids"]]
    assert len(toks) == 6
    assert chunks[-2]["choices"][0]["finish_reason"] == "length"
    assert chunks[-1]["choices"] == [] and chunks[-1]["usage"]["completion_tokens"] == 6
    st, body = http(port, "/v1/completions", {"prompt": prompt, "max_tokens": 6, "temperature": 0, "ignore_eos": True})
    assert json.loads(body)["choices"][0]["token_ids"] == toks


def test_errors_use_the_openai_error_shape(server):
    port = server[0]
    st, body = http(port, "/v1/completions", {"prompt": "hello", "max_tokens": 4})
    e = json.loads(body)["error"]
    assert st == 400 and e["code"] == 400 and "tokenizer" in e["message"] and e["type"] == "BadRequestError"
    st, body = http(port, "/v1/completions", {"model": "other", "prompt": [5, 6], "max_tokens": 4})
    assert st == 404 and "does not exist" in json.loads(body)["error"]["message"]
    st, body = http(port, "/v1/completions", {"prompt": [5, 10 ** 7]})
    assert st == 400 and "out of vocabulary" in body
    st, body = http(port, "/v1/completions", {"prompt": list(range(3, 300)), "max_tokens": 4})
    assert st == 400 and "max_model_len" in body
    st, body = http(port, "/v1/chat/completions", {"messages": [{"role": "user", "content": "hi"}]})
    assert st == 400 and "chat template" in body
    st, body = http(port, "/v1/chat/completions", {"prompt_token_ids": [5, 6, 7], "max_tokens": 3, "temperature": 0})
    r = json.loads(body)
    assert st == 200 and r["object"] == "chat.completion" and r["choices"][0]["message"]["role"] == "assistant"
    assert len(r["choices"][0]["token_ids"]) >= 1


def test_string_prompts_and_chat_with_a_tokenizer(server_tok):
    """With a tokenizer.json in the model directory the server takes text like the engine it replaces: the ids
    it generates for a string prompt equal those for the same prompt sent as token ids, `text` is their decoding,
    and chat messages are rendered with the Llama-3 header format."""
    from tokenizers import Tokenizer
    port, cfg, w, d = server_tok
    tok = Tokenizer.from_file(os.path.join(d, "tokenizer.json"))
    text = "The GPU cluster's nodes aren't ready yet."
    ids = tok.encode(text).ids
    st, a = http(port, "/v1/completions", {"prompt": text, "max_tokens": 8, "temperature": 0, "ignore_eos": True})
    st2, b = http(port, "/v1/completions", {"prompt": ids, "max_tokens": 8, "temperature": 0, "ignore_eos": True})
    assert st == 200 and st2 == 200, (a, b)
    a, b = json.loads(a), json.loads(b)
    assert a["usage"]["prompt_tokens"] == len(ids)
    assert a["choices"][0]["token_ids"] == b["choices"][0]["token_ids"]
    assert a["choices"][0]["text"] == tok.decode(a["choices"][0]["token_ids"], skip_special_tokens=True)
    st, c = http(port, "/v1/completions", {"prompt": [text, "Hello"], "max_tokens": 4, "temperature": 0})
    assert st == 200 and len(json.loads(c)["choices"]) == 2
    msgs = [{"role": "user", "content": "Hi there"}]
    st, c = http(port, "/v1/chat/completions", {"messages": msgs, "max_tokens": 6, "temperature": 0, "ignore_eos": True})
    c = json.loads(c)
    rendered = "<|begin_of_text|><|start_header_id|>user<|end_header_id|>\n\nHi there<|eot_id|><|start_header_id|>assistant<|end_header_id|>\n\n"
    assert st == 200 and c["object"] == "chat.completion"
    assert c["usage"]["prompt_tokens"] == len(tok.encode(rendered, add_special_tokens=False).ids)
    assert c["choices"][0]["message"]["content"] == tok.decode(c["choices"][0]["token_ids"], skip_special_tokens=True)
    # streamed text pieces concatenate to the non-streamed text
    req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/completions",
                                 data=json.dumps({"prompt": text, "max_tokens": 8, "temperature": 0, "ignore_eos": True,
                                                  "stream": True}).encode(), headers={"Content-Type": "application/json"})
    with urllib.request.urlopen(req, timeout=60) as r:
        raw = r.read().decode()
    pieces = [json.loads(e[6:])["choices"][0]["text"] for e in raw.split("\n\n") if e.startswith("data: {")]
    assert "".join(pieces) == a["choices"][0]["text"]


def test_stop_strings_and_stop_reason(server_tok):
    """`stop` strings end a request the way vLLM's detokenizer does (text cut before the stop string, or after it with
    include_stop_str_in_output; finish_reason "stop", stop_reason = the string; the request is aborted in the engine),
    in both response modes; a stop TOKEN id is reported as an integer stop_reason and its text is left out."""
    port, cfg, w, d = server_tok
    base = {"prompt": "The GPU cluster's nodes", "max_tokens": 24, "temperature": 0, "ignore_eos": True}
    st, a = http(port, "/v1/completions", base)
    assert st == 200, a
    full = json.loads(a)["choices"][0]
    text, ids = full["text"], full["token_ids"]
    assert len(ids) == 24 and full["finish_reason"] == "length" and full["stop_reason"] is None and len(text) >= 8
    mid = len(text) // 2
    stop = text[mid:mid + 3]
    cut = text.index(stop)
    st, b = http(port, "/v1/completions", dict(base, stop=[stop]))
    b = json.loads(b)
    c0 = b["choices"][0]
    assert st == 200 and c0["text"] == text[:cut] and c0["finish_reason"] == "stop" and c0["stop_reason"] == stop
    assert c0["token_ids"] == ids[:len(c0["token_ids"])] and b["usage"]["completion_tokens"] == len(c0["token_ids"]) <= 24
    st, b = http(port, "/v1/completions", dict(base, stop=stop, include_stop_str_in_output=True))       # a bare string is accepted
    c0 = json.loads(b)["choices"][0]
    assert c0["text"] == text[:cut + len(stop)] and c0["finish_reason"] == "stop" and c0["stop_reason"] == stop
    st, b = http(port, "/v1/completions", dict(base, stop=["\x00 never there"]))
    assert json.loads(b)["choices"][0]["text"] == text
    # streaming: nothing beyond the cut is ever sent, the last chunk carries the reason
    req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/completions", data=json.dumps(dict(base, stop=[stop], stream=True)).encode(),
                                 headers={"Content-Type": "application/json"})
    with urllib.request.urlopen(req, timeout=60) as r:
        raw = r.read().decode()
    chunks = [json.loads(e[6:])["choices"][0] for e in raw.split("\n\n") if e.startswith("data: {")]
    assert "".join(c["text"] for c in chunks) == text[:cut] and raw.rstrip().endswith("data: [DONE]")
    assert chunks[-1]["finish_reason"] == "stop" and chunks[-1]["stop_reason"] == stop
    assert all(c["finish_reason"] is None for c in chunks[:-1])
    # stop token id: first occurrence of the 6th generated token ends the request; its text is not part of the output
    sid = ids[5]
    k = ids.index(sid)
    st, b = http(port, "/v1/completions", dict(base, stop_token_ids=[sid]))
    c0 = json.loads(b)["choices"][0]
    assert c0["token_ids"] == ids[:k + 1] and c0["finish_reason"] == "stop" and c0["stop_reason"] == sid
    from tokenizers import Tokenizer
    tok = Tokenizer.from_file(os.path.join(d, "tokenizer.json"))
    assert c0["text"] == tok.decode(ids[:k], skip_special_tokens=True)
    # echo: the prompt's text in front of the completion
    st, b = http(port, "/v1/completions", dict(base, echo=True, max_tokens=4))
    c0 = json.loads(b)["choices"][0]
    assert st == 200 and c0["text"] == base["prompt"] + tok.decode(ids[:4], skip_special_tokens=True)
    # the server keeps serving and has released the aborted requests' KV blocks
    st, m = http(port, "/metrics")
    assert st == 200 and "vllm:num_requests_running" in m
    st, a2 = http(port, "/v1/completions", base)
    assert json.loads(a2)["choices"][0]["token_ids"] == ids


def test_logprobs_in_both_api_shapes(server_tok, server):
    """`logprobs` (completions: N; chat: logprobs=true + top_logprobs=N) in the OpenAI shapes vLLM emits: the sampled
    token's log-probability, the top-N alternatives, text offsets; greedy sampling makes the sampled token the most
    likely one.  Without a tokenizer, tokens are spelled token_id:N."""
    port, cfg, w, d = server_tok
    base = {"prompt": "The GPU cluster's nodes", "max_tokens": 6, "temperature": 0, "ignore_eos": True}
    st, a = http(port, "/v1/completions", dict(base, logprobs=3))
    assert st == 200, a
    c0 = json.loads(a)["choices"][0]
    lp = c0["logprobs"]
    assert len(lp["tokens"]) == len(lp["token_logprobs"]) == len(lp["top_logprobs"]) == len(lp["text_offset"]) == 6
    off = 0
    for t, v, top, o in zip(lp["tokens"], lp["token_logprobs"], lp["top_logprobs"], lp["text_offset"]):
        assert o == off and v <= 0 and 3 <= len(top) <= 4
        assert abs(max(top.values()) - v) < 1e-6            # greedy: nothing is more likely than the sampled token
        assert sum(math.exp(x) for x in top.values()) <= 1.0 + 1e-3
        off += len(t)
    st, b = http(port, "/v1/completions", base)
    assert json.loads(b)["choices"][0]["logprobs"] is None
    st, b = http(port, "/v1/completions", dict(base, logprobs=0))
    assert all(len(t) == 1 for t in json.loads(b)["choices"][0]["logprobs"]["top_logprobs"])
    st, c = http(port, "/v1/chat/completions", {"messages": [{"role": "user", "content": "Hi"}], "max_tokens": 4, "temperature": 0,
                                                "ignore_eos": True, "logprobs": True, "top_logprobs": 2})
    assert st == 200, c
    content = json.loads(c)["choices"][0]["logprobs"]["content"]
    assert len(content) == 4
    for e in content:
        assert e["logprob"] <= 0 and bytes(e["bytes"]).decode("utf-8", "replace") == e["token"] and len(e["top_logprobs"]) == 2
        assert e["top_logprobs"][0]["token"] == e["token"] and abs(e["top_logprobs"][0]["logprob"] - e["logprob"]) < 1e-6
    # streaming carries one entry per chunk
    req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/completions", data=json.dumps(dict(base, logprobs=2, stream=True)).encode(),
                                 headers={"Content-Type": "application/json"})
    with urllib.request.urlopen(req, timeout=60) as r:
        raw = r.read().decode()
    chunks = [json.loads(e[6:])["choices"][0] for e in raw.split("\n\n") if e.startswith("data: {")]
    assert len(chunks) == 6 and all(len(ch["logprobs"]["tokens"]) == 1 for ch in chunks)
    assert [ch["logprobs"]["token_logprobs"][0] for ch in chunks] == pytest.approx(lp["token_logprobs"], abs=1e-5)
    assert [ch["logprobs"]["text_offset"][0] for ch in chunks] == lp["text_offset"]
    st, e = http(port, "/v1/completions", dict(base, logprobs=50))
    assert st == 400 and "logprobs" in e
    # no tokenizer: ids in, token_id:N out
    port2 = server[0]
    st, f = http(port2, "/v1/completions", {"prompt": [5, 6, 7], "max_tokens": 3, "temperature": 0, "ignore_eos": True, "logprobs": 1})
    c1 = json.loads(f)["choices"][0]
    assert st == 200 and c1["logprobs"]["tokens"] == [f"token_id:{t}" for t in c1["token_ids"]]
