"""kge-router (the request router in front of engine replicas, BASELINE config 5) against fake engine backends:
round-robin, session stickiness, model filtering, failover when a backend dies, verbatim relay of SSE streams,
/v1/models aggregation, /health and /metrics.  CPU only."""
import http.server
import json
import os
import socket
import subprocess
import threading
import time
import urllib.error
import urllib.request

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "kubernetes-gpu-cluster_b200", "bin", "kge-router")
pytestmark = pytest.mark.skipif(not os.path.exists(BIN), reason="router binary not built")


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class FakeEngine(http.server.BaseHTTPRequestHandler):
    protocol_version = "HTTP/1.1"

    def log_message(self, *a):
        pass

    def _send(self, code, body, ctype="application/json"):
        b = body.encode()
        self.send_response(code)
        self.send_header("Content-Type", ctype)
        self.send_header("Content-Length", str(len(b)))
        self.send_header("Connection", "close")
        self.end_headers()
        self.wfile.write(b)

    def do_GET(self):
        name = self.server.model_name
        if self.path == "/health":
            self.send_response(200)
            self.send_header("Content-Length", "0")
            self.send_header("Connection", "close")
            self.end_headers()
        elif self.path == "/v1/models":
            self._send(200, json.dumps({"object": "list", "data": [{"id": name, "object": "model", "owned_by": "vllm", "root": name,
                                                                     "max_model_len": 4096, "created": 1}]}))
        else:
            self._send(404, "{}")

    def do_POST(self):
        n = int(self.headers.get("Content-Length", "0"))
        try:
            req = json.loads(self.rfile.read(n))
            assert isinstance(req, dict)
        except Exception:
            return self._send(400, '{"error":{"message":"bad json","type":"BadRequestError","param":null,"code":400}}')
        port = self.server.server_address[1]
        self.server.seen.append(req)
        if req.get("stream"):
            self.send_response(200)
            self.send_header("Content-Type", "text/event-stream")
            self.send_header("Transfer-Encoding", "chunked")
            self.send_header("Connection", "close")
            self.end_headers()
            for i in range(3):
                data = f"data: {json.dumps({'port': port, 'i': i})}\n\n".encode()
                self.wfile.write(b"%x\r\n%s\r\n" % (len(data), data))
                self.wfile.flush()
                time.sleep(0.05)
            data = b"data: [DONE]\n\n"
            self.wfile.write(b"%x\r\n%s\r\n0\r\n\r\n" % (len(data), data))
        else:
            self._send(200, json.dumps({"port": port, "path": self.path, "model": req.get("model"), "auth": self.headers.get("Authorization")}))


def start_backend(model):
    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), FakeEngine)
    srv.model_name = model
    srv.seen = []
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    return srv


def start_router(backends, extra=()):
    port = free_port()
    urls = ",".join(f"http://127.0.0.1:{b.server_address[1]}" for b in backends)
    proc = subprocess.Popen([BIN, "--host", "127.0.0.1", "--port", str(port), "--service-discovery", "static", "--static-backends", urls,
                             *extra], stderr=subprocess.PIPE)
    for _ in range(100):
        try:
            with urllib.request.urlopen(f"http://127.0.0.1:{port}/health", timeout=1) as r:
                if r.status == 200:
                    return proc, port
        except Exception:
            time.sleep(0.05)
    proc.kill()
    raise RuntimeError("router did not start: " + proc.stderr.read().decode())


def post(port, path, body, headers=None):
    req = urllib.request.Request(f"http://127.0.0.1:{port}{path}", data=json.dumps(body).encode(),
                                 headers={"Content-Type": "application/json", **(headers or {})})
    try:
        with urllib.request.urlopen(req, timeout=10) as r:
            return r.status, r.read().decode()
    except urllib.error.HTTPError as e:
        return e.code, e.read().decode()


def test_round_robin_failover_models_metrics():
    b1, b2 = start_backend("m"), start_backend("m")
    proc, port = start_router([b1, b2])
    try:
        ports = [json.loads(post(port, "/v1/completions", {"model": "m", "prompt": [1]}, {"Authorization": "Bearer k"})[1])["port"] for _ in range(6)]
        p1, p2 = b1.server_address[1], b2.server_address[1]
        assert sorted(set(ports)) == sorted([p1, p2]) and all(a != b for a, b in zip(ports, ports[1:]))       # strict alternation
        st, body = post(port, "/v1/chat/completions", {"model": "m", "messages": []}, {"Authorization": "Bearer k"})
        assert st == 200 and json.loads(body)["path"] == "/v1/chat/completions" and json.loads(body)["auth"] == "Bearer k"
        with urllib.request.urlopen(f"http://127.0.0.1:{port}/v1/models", timeout=5) as r:
            assert [m["id"] for m in json.loads(r.read())["data"]] == ["m"]                                     # merged, no duplicates
        # one backend dies: requests keep succeeding on the other, immediately (failover) and after the health check
        b2.shutdown()
        b2.server_close()
        got = [json.loads(post(port, "/v1/completions", {"model": "m", "prompt": [1]})[1])["port"] for _ in range(4)]
        assert got == [p1] * 4
        with urllib.request.urlopen(f"http://127.0.0.1:{port}/metrics", timeout=5) as r:
            m = r.read().decode()
        assert "vllm_router:healthy_pods_total 1" in m and f'vllm_router:requests_total{{server="http://127.0.0.1:{p1}"}}' in m
        st, body = post(port, "/nope", {})
        assert st == 404 and json.loads(body)["error"]["code"] == 404
    finally:
        proc.terminate()
        proc.wait(timeout=5)
        b1.shutdown()


def test_session_routing_and_model_filter_and_streaming():
    a, b, c = start_backend("x"), start_backend("x"), start_backend("y")
    pa, pb, pc = (s.server_address[1] for s in (a, b, c))
    proc, port = start_router([a, b, c], ["--static-models", "x,x,y", "--routing-logic", "session", "--session-key", "x-user-id"])
    try:
        for user in ("alice", "bob", "carol", "dave"):
            ports = {json.loads(post(port, "/v1/completions", {"model": "x", "prompt": [1]}, {"x-user-id": user})[1])["port"] for _ in range(4)}
            assert len(ports) == 1 and ports <= {pa, pb}                     # sticky, and only backends that serve the model
        assert json.loads(post(port, "/v1/completions", {"model": "y", "prompt": [1]}, {"x-user-id": "alice"})[1])["port"] == pc
        no_session = {json.loads(post(port, "/v1/completions", {"model": "x", "prompt": [1]})[1])["port"] for _ in range(4)}
        assert no_session == {pa, pb}                                         # without the header: round-robin
        st, body = post(port, "/v1/completions", {"model": "zzz", "prompt": [1]})
        assert st == 404 and "does not exist" in json.loads(body)["error"]["message"]
        with urllib.request.urlopen(f"http://127.0.0.1:{port}/v1/models", timeout=5) as r:
            assert sorted(m["id"] for m in json.loads(r.read())["data"]) == ["x", "y"]
        # SSE: chunks are relayed as they come and arrive complete
        req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/completions", data=json.dumps({"model": "y", "prompt": [1], "stream": True}).encode(),
                                     headers={"Content-Type": "application/json"})
        t0 = time.time()
        first = None
        events = []
        with urllib.request.urlopen(req, timeout=10) as r:
            assert r.headers["Content-Type"] == "text/event-stream"
            for line in r:
                if line.startswith(b"data: "):
                    if first is None:
                        first = time.time() - t0
                    events.append(line[6:].strip().decode())
        assert events[-1] == "[DONE]" and [json.loads(e)["i"] for e in events[:-1]] == [0, 1, 2]
        for s in (a, b, c):
            s.shutdown()
            s.server_close()
        time.sleep(0.1)
        st, body = post(port, "/v1/completions", {"model": "x", "prompt": [1]})
        assert st == 503 and json.loads(body)["error"]["code"] == 503
    finally:
        proc.terminate()
        proc.wait(timeout=5)


def test_bad_arguments():
    assert subprocess.run([BIN, "--service-discovery", "k8s", "--static-backends", "http://127.0.0.1:1"], capture_output=True).returncode == 2
    assert subprocess.run([BIN, "--port", "1"], capture_output=True).returncode == 2
    assert subprocess.run([BIN, "--static-backends", "http://a:1,http://b:2", "--static-models", "m"], capture_output=True).returncode == 2


def test_router_survives_garbage_under_sanitizers(tmp_path):
    """Malformed, truncated and oversized requests against a router built with ASan + UBSan: every connection is answered
    or dropped, nothing crashes, and well-formed traffic keeps flowing."""
    import random
    import shutil
    if shutil.which("g++") is None:
        pytest.skip("needs g++")
    exe = str(tmp_path / "kge-router-asan")
    src = os.path.join(ROOT, "kubernetes-gpu-cluster_b200", "router", "router.cc")
    p = subprocess.run(["g++", "-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-g", "-O1", "-std=c++17", src, "-o", exe, "-lpthread"],
                       capture_output=True, text=True, timeout=600)
    if p.returncode and "cannot find" in p.stderr:
        pytest.skip("sanitizer runtime not installed")
    assert p.returncode == 0, p.stderr[-2000:]
    global BIN
    b = start_backend("m")
    saved, BIN = BIN, exe
    try:
        proc, port = start_router([b])
    finally:
        BIN = saved
    rng = random.Random(0)
    try:
        payloads = [b"", b"\r\n\r\n", b"GET\r\n\r\n", b"POST /v1/completions HTTP/1.1\r\nContent-Length: 999999999999\r\n\r\n{}",
                    b"POST /v1/completions HTTP/1.1\r\nContent-Length: 5\r\n\r\n{}", b"GET /" + b"a" * 70000 + b" HTTP/1.1\r\n\r\n",
                    b"POST /v1/completions HTTP/1.1\r\nContent-Length: 3\r\n\r\n\xff\xfe\xfd", b"X" * (2 << 20)]
        payloads += [bytes(rng.randrange(256) for _ in range(rng.randrange(1, 400))) for _ in range(40)]
        payloads += [b"POST /v1/chat/completions HTTP/1.1\r\nContent-Length: %d\r\n\r\n%s" % (len(x), x)
                     for x in (b"[" * 5000, b'{"model":' + b'"' + b"\\" * 999 + b'"}', b'{"model": 1e999}')]
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
        st, body = post(port, "/v1/completions", {"model": "m", "prompt": [1]})
        assert st == 200 and json.loads(body)["port"] == b.server_address[1]
    finally:
        proc.terminate()
        try:
            proc.wait(timeout=10)
        except subprocess.TimeoutExpired:
            proc.kill()
        err = proc.stderr.read().decode()
        b.shutdown()
    assert "ERROR: AddressSanitizer" not in err and "runtime error" not in err, err[-3000:]
