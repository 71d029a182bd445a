"""The native front end of `dllama-api` (csrc/host/server.cpp) on the CPU engine: same contract as the Python
front end (tests/test_serving.py), which also serves as the oracle for the generated text."""
import json
import socket
import threading
import time
import urllib.error
import urllib.request

import pytest

from dllama_b200.formats import tfile
from dllama_b200.models.cpu_engine import CpuEngine
from dllama_b200.serving import api_types as A
from dllama_b200.serving.api import ApiServer, ApiState
from dllama_b200.serving.batch_engine import BatchEngine
from dllama_b200.serving.native_server import NativeApiServer
from dllama_b200.text import native as T


@pytest.fixture(scope="module")
def tok(tmp_path_factory):
    p = str(tmp_path_factory.mktemp("tok") / "tiny.t")
    tfile.write_synthetic_tokenizer(p, 1024)
    return T.Tokenizer(p)


def _native(tiny_model_path, tok, engine=None, **kw):
    eng = engine or CpuEngine(tiny_model_path, max_rows=8, n_slots=4, max_seq_len=256)
    defaults = A.InferenceParams(messages=[], temperature=0.0, top_p=0.9, seed=1)
    srv = NativeApiServer(eng, tok, "tiny.m", defaults, host="127.0.0.1", port=0, verbose=False, **kw)
    srv.start()
    return srv


def _post(port, body, timeout=120):
    req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/chat/completions", data=json.dumps(body).encode(),
                                 headers={"Content-Type": "application/json"}, method="POST")
    return urllib.request.urlopen(req, timeout=timeout)


def _sse_events(raw: str):
    return [e[len("data: "):] for e in raw.split("\r\n\r\n") if e.startswith("data: ")]


def test_native_server_openai_shape_sse_and_errors(tiny_model_path, tok):
    """Routes, CORS, concurrent users in one batch, chat.completion / SSE chunk shapes, usage, 400s, metrics --
    and text identical to the Python front end's."""
    srv = _native(tiny_model_path, tok)
    port = srv.port
    eng2 = CpuEngine(tiny_model_path, max_rows=8, n_slots=4, max_seq_len=256)
    be = BatchEngine(eng2, tok)
    be.start()
    pstate = ApiState(be, tok, "tiny.m", A.InferenceParams(messages=[], temperature=0.0, top_p=0.9, seed=1))
    psrv = ApiServer(pstate, "127.0.0.1", 0)
    psrv.start()
    try:
        with urllib.request.urlopen(f"http://127.0.0.1:{port}/v1/models") as r:
            models = json.load(r)
        assert models["object"] == "list" and models["data"][0]["id"] == "tiny.m"
        opt = urllib.request.Request(f"http://127.0.0.1:{port}/v1/chat/completions", method="OPTIONS")
        with urllib.request.urlopen(opt) as r:
            assert r.status == 204 and r.headers["Access-Control-Allow-Origin"] == "*"
        with pytest.raises(urllib.error.HTTPError) as ei:
            urllib.request.urlopen(f"http://127.0.0.1:{port}/nope")
        assert ei.value.code == 404
        with urllib.request.urlopen(f"http://127.0.0.1:{port}/health") as r:
            assert json.load(r)["status"] == "ok"
        results = {}

        def user(i, text):
            with _post(port, {"messages": [{"role": "user", "content": text}], "max_tokens": 6}) as r:
                results[i] = json.load(r)

        texts = ["hello world", "the of and", "hello world", "a \"quoted\"\n\\ line é世 \U0001F600"]
        threads = [threading.Thread(target=user, args=(i, t)) for i, t in enumerate(texts)]
        [t.start() for t in threads]
        [t.join(timeout=180) for t in threads]
        assert len(results) == 4
        for i, d in results.items():
            assert d["object"] == "chat.completion" and d["choices"][0]["message"]["role"] == "assistant"
            assert d["id"].startswith("cmpl-") and d["model"] == "tiny.m"
            assert 1 <= d["usage"]["completion_tokens"] <= 6
            assert d["usage"]["total_tokens"] == d["usage"]["prompt_tokens"] + d["usage"]["completion_tokens"]
            assert d["generated_text"] == d["choices"][0]["message"]["content"]
            assert d["choices"][0]["finish_reason"] in ("stop", "length")
            with _post(psrv.port, {"messages": [{"role": "user", "content": texts[i]}], "max_tokens": 6}) as r:
                want = json.load(r)
            assert d["choices"][0]["message"]["content"] == want["choices"][0]["message"]["content"]
            assert d["usage"] == want["usage"] and d["choices"][0]["finish_reason"] == want["choices"][0]["finish_reason"]
        # streaming: SSE chunks concatenate to the non-streamed answer
        with _post(port, {"messages": [{"role": "user", "content": "hello world"}], "max_tokens": 6,
                          "stream": True}) as r:
            assert r.headers["Content-Type"].startswith("text/event-stream")
            raw = r.read().decode("utf-8")
        events = _sse_events(raw)
        assert events[-1] == "[DONE]"
        chunks = [json.loads(e) for e in events[:-1]]
        assert all(c["object"] == "chat.completion.chunk" for c in chunks)
        assert chunks[0]["choices"][0]["delta"] == {"role": "assistant", "content": ""}
        text = "".join(c["choices"][0]["delta"].get("content", "") for c in chunks)
        assert text == results[0]["choices"][0]["message"]["content"]
        assert "usage" not in chunks[-1] and chunks[-1]["choices"][0]["finish_reason"] in ("stop", "length")
        with _post(port, {"messages": [{"role": "user", "content": "hello world"}], "max_tokens": 6, "stream": True,
                          "stream_options": {"include_usage": True}}) as r:
            last = json.loads(_sse_events(r.read().decode("utf-8"))[-2])
        assert last["usage"] == results[0]["usage"]
        # OpenAI content parts + system message + max_completion_tokens
        with _post(port, {"messages": [{"role": "system", "content": "be brief"},
                                       {"role": "user", "content": [{"type": "text", "text": "hello "},
                                                                    {"type": "text", "text": "world"}]}],
                          "max_completion_tokens": 3}) as r:
            assert json.load(r)["usage"]["completion_tokens"] <= 3
        # bad requests: wrong type, empty messages, missing content, prompt longer than the context window
        for body in ({"messages": "nope"}, {"messages": []}, {"messages": [{"role": "user"}]},
                     {"messages": [{"role": "user", "content": "hello " * 400}]}):
            with pytest.raises(urllib.error.HTTPError) as ei:
                _post(port, body)
            assert ei.value.code == 400 and "error" in json.load(ei.value)
        req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/chat/completions", data=b"{\"messages\": [",
                                     method="POST")
        with pytest.raises(urllib.error.HTTPError) as ei:
            urllib.request.urlopen(req)
        assert ei.value.code == 400
        with urllib.request.urlopen(f"http://127.0.0.1:{port}/metrics") as r:
            m = r.read()
        assert b"dllama_requests_total 7" in m and b"dllama_errors_total 5" in m
        st = srv.stats()
        assert st["waiting"] == 0 and st["running"] == 0 and st["free_slots"] == 4
    finally:
        srv.close()
        psrv.shutdown()
        be.stop()


def test_native_server_sampling_stop_strings_disconnect_and_engine_failure(tiny_model_path, tok):
    eng = CpuEngine(tiny_model_path, max_rows=8, n_slots=3, max_seq_len=256)
    real_forward = eng.forward
    fail = {"on": False}

    def flaky(*a, **kw):
        if fail["on"]:
            fail["on"] = False
            raise RuntimeError("KV cache exhausted")
        return real_forward(*a, **kw)

    eng.forward = flaky
    srv = _native(tiny_model_path, tok, engine=eng)
    port = srv.port
    msg = [{"role": "user", "content": "hello world"}]
    try:
        def ask(**kw):
            with _post(port, {"messages": msg, **kw}) as r:
                return json.load(r)
        # per-request seeded sampling (host sampler on the CPU engine): same seed -> same text
        a = ask(max_tokens=8, temperature=1.0, top_p=0.95, seed=11)
        b = ask(max_tokens=8, temperature=1.0, top_p=0.95, seed=11)
        c = ask(max_tokens=8, temperature=1.0, top_p=0.95, seed=12)
        assert a["choices"][0]["message"]["content"] == b["choices"][0]["message"]["content"]
        assert a["choices"][0]["message"]["content"] != c["choices"][0]["message"]["content"]
        # stop string truncates the answer
        full = ask(max_tokens=12, temperature=0.0)["choices"][0]["message"]["content"]
        assert len(full) >= 4
        stop = full[2:4]
        cut = ask(max_tokens=12, temperature=0.0, stop=[stop])
        assert cut["choices"][0]["message"]["content"] == full[:full.index(stop)]
        assert cut["choices"][0]["finish_reason"] == "stop"
        # a client that hangs up mid-stream frees its slot
        generated_before = srv.stats()["generated"]
        body = json.dumps({"messages": msg, "max_tokens": 200, "stream": True}).encode()
        s = socket.create_connection(("127.0.0.1", port))
        s.sendall(b"POST /v1/chat/completions HTTP/1.1\r\nHost: x\r\nContent-Type: application/json\r\n"
                  b"Content-Length: " + str(len(body)).encode() + b"\r\n\r\n" + body)
        assert s.recv(64).startswith(b"HTTP/1.1 200")
        s.close()
        t0 = time.time()
        while time.time() - t0 < 60:
            st = srv.stats()
            if st["running"] == 0 and st["waiting"] == 0 and st["free_slots"] == 3:
                break
            time.sleep(0.05)
        assert st["free_slots"] == 3 and st["generated"] - generated_before < 200
        # a failing forward answers 500 for that batch only; the next request is served
        fail["on"] = True
        with pytest.raises(urllib.error.HTTPError) as ei:
            ask(max_tokens=4)
        assert ei.value.code == 500 and "exhausted" in json.load(ei.value)["error"]["message"]
        ok = ask(max_tokens=4, temperature=0.0)
        assert ok["usage"]["completion_tokens"] >= 1
        st = srv.stats()
        assert st["engine_errors"] == 1 and st["free_slots"] == 3 and srv.n_errors == 1
        # the same failure inside an SSE stream: an `error` object in the last chunk
        fail["on"] = True
        with _post(port, {"messages": msg, "max_tokens": 4, "stream": True}) as r:
            ev = _sse_events(r.read().decode("utf-8"))
        last = json.loads(ev[-2])
        assert ev[-1] == "[DONE]" and last["choices"][0]["finish_reason"] == "error"
        assert "exhausted" in last["error"]["message"]
    finally:
        srv.close()


def test_native_server_many_concurrent_streams(tiny_model_path, tok):
    """More users than batch rows and slots: everybody is served, nothing leaks."""
    eng = CpuEngine(tiny_model_path, max_rows=4, n_slots=3, max_seq_len=256)
    srv = _native(tiny_model_path, tok, engine=eng)
    port = srv.port
    out = {}

    def user(i):
        with _post(port, {"messages": [{"role": "user", "content": "hello world " + "the " * (i % 5)}],
                          "max_tokens": 5, "stream": i % 2 == 0}) as r:
            raw = r.read().decode("utf-8")
        if i % 2 == 0:
            ev = _sse_events(raw)
            out[i] = "".join(json.loads(e)["choices"][0]["delta"].get("content", "") for e in ev[:-1])
        else:
            out[i] = json.loads(raw)["choices"][0]["message"]["content"]

    try:
        threads = [threading.Thread(target=user, args=(i,)) for i in range(12)]
        [t.start() for t in threads]
        [t.join(timeout=240) for t in threads]
        assert len(out) == 12
        assert out[0] == out[5] == out[10] and out[1] == out[6] == out[11]          # same prompt -> same text
        st = srv.stats()
        assert st["requests"] == 12 and st["running"] == 0 and st["waiting"] == 0 and st["free_slots"] == 3
    finally:
        srv.close()


class _FakeEngine:
    """Engine surface without a model (fixed tokens): lets the front end be hammered quickly."""

    class _Pages:
        def __init__(self):
            self.released = []

        def release_pages(self, slot):
            self.released.append(slot)

    class _H:
        vocab_size = 1024

    def __init__(self, max_rows, n_slots, seq_len):
        import torch
        self.h, self.pages = self._H(), self._Pages()
        self.max_rows, self.n_slots, self.seq_len = max_rows, n_slots, seq_len
        self.device = torch.device("cpu")

    def forward(self, tokens, positions, slots, logits_from=0, greedy=False):
        import torch
        n = len(tokens) - logits_from
        lg = torch.zeros(n, 1024)
        for i in range(n):
            lg[i, 5 + (positions[logits_from + i] * 7 + slots[logits_from + i]) % 900] = 10.0
        return lg

    def decode_step_greedy(self, tokens, positions, slots, logits_from=0):
        return [5 + (positions[i] * 7 + slots[i]) % 900 for i in range(logits_from, len(tokens))]


def test_native_server_survives_a_storm_of_partial_requests_and_disconnects(tok):
    """Clients that connect and leave at every stage (no bytes, half a header, half a body, mid-stream), pipelined
    garbage, oversized declarations -- while well-behaved users are being served.  Nothing may leak or wedge."""
    eng = _FakeEngine(max_rows=8, n_slots=4, seq_len=256)
    defaults = A.InferenceParams(messages=[], temperature=0.0, top_p=0.9, seed=1)
    srv = NativeApiServer(eng, tok, "fake.m", defaults, host="127.0.0.1", port=0, verbose=False)
    srv.start()
    port = srv.port
    body = json.dumps({"messages": [{"role": "user", "content": "hello world the of and"}], "max_tokens": 40,
                       "stream": True}).encode()
    head = (b"POST /v1/chat/completions HTTP/1.1\r\nHost: x\r\nContent-Type: application/json\r\nContent-Length: "
            + str(len(body)).encode() + b"\r\n\r\n")
    good = {}

    def good_user(i):
        with _post(port, {"messages": [{"role": "user", "content": "hello world"}], "max_tokens": 16}) as r:
            good[i] = json.load(r)

    def bad_client(kind):
        try:
            s = socket.create_connection(("127.0.0.1", port), timeout=10)
            if kind == 0:
                pass                                            # connect and leave
            elif kind == 1:
                s.sendall(head[:25])                            # half a request line / header
            elif kind == 2:
                s.sendall(head + body[:10])                     # half a body
            elif kind == 3:
                s.sendall(head + body)                          # full request, leave while it streams
                s.recv(200)
            elif kind == 4:
                s.sendall(b"\x00\xff garbage\r\n\r\n")          # not HTTP
                s.recv(200)
            elif kind == 5:
                s.sendall(b"POST /v1/chat/completions HTTP/1.1\r\nContent-Length: 99999999999\r\n\r\n")
                assert b"413" in s.recv(200)
            elif kind == 6:
                s.sendall(head + body + head + body)            # pipelined second request is ignored
                s.recv(200)
            s.close()
        except OSError:
            pass

    try:
        threads = [threading.Thread(target=bad_client, args=(k % 7,)) for k in range(70)]
        threads += [threading.Thread(target=good_user, args=(i,)) for i in range(10)]
        import random
        random.Random(3).shuffle(threads)
        [t.start() for t in threads]
        [t.join(timeout=120) for t in threads]
        assert len(good) == 10 and all(g["usage"]["completion_tokens"] >= 1 for g in good.values())
        t0 = time.time()
        while time.time() - t0 < 60:
            st = srv.stats()
            if st["running"] == 0 and st["waiting"] == 0 and st["free_slots"] == 4 and st["connections"] <= 20:
                break
            time.sleep(0.05)
        assert st["running"] == 0 and st["waiting"] == 0 and st["free_slots"] == 4, st
        # half-open clients (kinds 1, 2) are still counted until their deadline; everything else is gone
        assert st["connections"] <= 20, st
        with _post(port, {"messages": [{"role": "user", "content": "hello world"}], "max_tokens": 4}) as r:
            assert json.load(r)["usage"]["completion_tokens"] >= 1
    finally:
        srv.close()


def test_full_native_stack_http_server_on_the_tcp_tensor_parallel_cpu_backend(tiny_model_path, tok):
    """HTTP -> C++ server -> C++ host backend on two nodes over loopback TCP: the whole no-GPU product with no
    Python on any per-token path except the engine thread's three calls.  Text must equal the single-node server's."""
    from dllama_b200.models import native_cpu_engine as NC
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        wport = s.getsockname()[1]
    w = threading.Thread(target=NC.serve_worker, args=(wport, 1), daemon=True)
    w.start()
    time.sleep(0.3)
    eng2 = NC.NativeCpuEngine(tiny_model_path, workers=[("127.0.0.1", wport)], n_threads=1, max_rows=8, n_slots=4, max_seq_len=256)
    eng1 = NC.NativeCpuEngine(tiny_model_path, n_threads=2, max_rows=8, n_slots=4, max_seq_len=256)
    srv2, srv1 = _native(tiny_model_path, tok, engine=eng2), _native(tiny_model_path, tok, engine=eng1)
    try:
        out = {}

        def user(key, port, text):
            with _post(port, {"messages": [{"role": "user", "content": text}], "max_tokens": 8}) as r:
                out[key] = json.load(r)

        texts = ["hello world", "the of and", "token model fast"]
        threads = [threading.Thread(target=user, args=((p, i), p, t)) for p in (srv1.port, srv2.port) for i, t in enumerate(texts)]
        [t.start() for t in threads]
        [t.join(timeout=180) for t in threads]
        assert len(out) == 6
        for i in range(3):
            a, b = out[(srv1.port, i)], out[(srv2.port, i)]
            assert a["choices"][0]["message"]["content"] == b["choices"][0]["message"]["content"] and a["usage"] == b["usage"]
        assert eng2.network_stats()["sent_bytes"] > 0
    finally:
        srv2.close()
        srv1.close()
        eng2.close()
        eng1.close()
        w.join(timeout=30)
    assert not w.is_alive()


def test_native_server_keep_alive_connections(tok):
    """HTTP/1.1 clients that keep their connection (httpx, requests, browsers) reuse it for several requests;
    streams, errors and `Connection: close` requests end it."""
    import http.client
    eng = _FakeEngine(max_rows=8, n_slots=4, seq_len=256)
    defaults = A.InferenceParams(messages=[], temperature=0.0, top_p=0.9, seed=1)
    srv = NativeApiServer(eng, tok, "fake.m", defaults, host="127.0.0.1", port=0, verbose=False)
    srv.start()
    try:
        c = http.client.HTTPConnection("127.0.0.1", srv.port, timeout=60)
        body = json.dumps({"messages": [{"role": "user", "content": "hello world"}], "max_tokens": 5})
        c.request("GET", "/v1/models")
        r = c.getresponse()
        assert r.status == 200 and r.getheader("Connection") == "keep-alive" and json.load(r)["object"] == "list"
        sock = c.sock
        texts = []
        for _ in range(3):
            c.request("POST", "/v1/chat/completions", body=body, headers={"Content-Type": "application/json"})
            r = c.getresponse()
            assert r.status == 200 and r.getheader("Connection") == "keep-alive"
            texts.append(json.load(r)["choices"][0]["message"]["content"])
            assert c.sock is sock                                   # same TCP connection, not a silent reconnect
        assert srv.stats()["connections"] == 1 and texts[0] == texts[1] == texts[2]
        c.request("OPTIONS", "/v1/chat/completions")
        r = c.getresponse()
        assert r.status == 204 and r.read() == b"" and c.sock is sock
        # a stream has no length: it ends with the connection
        c.request("POST", "/v1/chat/completions", body=json.dumps({"messages": [{"role": "user", "content": "hello"}],
                                                                   "max_tokens": 3, "stream": True}))
        r = c.getresponse()
        assert r.getheader("Connection") == "close" and r.read().endswith(b"data: [DONE]\r\n\r\n")
        c.close()
        # errors close too; so does an explicit Connection: close
        c = http.client.HTTPConnection("127.0.0.1", srv.port, timeout=60)
        c.request("POST", "/v1/chat/completions", body="{\"messages\": 1}")
        r = c.getresponse()
        assert r.status == 400 and r.getheader("Connection") == "close"
        r.read()
        c.close()
        c = http.client.HTTPConnection("127.0.0.1", srv.port, timeout=60)
        c.request("GET", "/health", headers={"Connection": "close"})
        r = c.getresponse()
        assert r.status == 200 and r.getheader("Connection") == "close"
        r.read()
        c.close()
        t0 = time.time()
        while srv.stats()["connections"] and time.time() - t0 < 10:
            time.sleep(0.05)
        assert srv.stats()["connections"] == 0
    finally:
        srv.close()


def test_native_server_answers_every_mutated_request_body(tok):
    """Mutation fuzzing of the JSON body (truncations, byte flips, deep nesting, huge numbers, odd escapes, wrong
    types): every request gets a well-formed 200 or 400 and the server keeps its slots.  (scripts/host_sanitizers.sh
    runs this file under ASan / TSan.)"""
    import random
    eng = _FakeEngine(max_rows=8, n_slots=4, seq_len=256)
    defaults = A.InferenceParams(messages=[], temperature=0.0, top_p=0.9, seed=1)
    srv = NativeApiServer(eng, tok, "fake.m", defaults, host="127.0.0.1", port=0, verbose=False)
    srv.start()
    rnd = random.Random(7)
    base = json.dumps({"messages": [{"role": "system", "content": "be \"brief\" \\ \u00e9\u4e16\U0001F600"},
                                    {"role": "user", "content": [{"type": "text", "text": "hello world"}]}],
                       "max_tokens": 3, "temperature": 0.5, "top_p": 0.9, "seed": 42, "stop": ["zz", "q"],
                       "stream": False, "stream_options": {"include_usage": True}}).encode()
    bodies = [b"", b"{}", b"[]", b"null", b"[" * 200 + b"]" * 200, b"{\"messages\": " + b"[" * 100 + b"]" * 100 + b"}",
              b"{\"messages\": [{\"role\": \"user\", \"content\": \"\\ud800\"}], \"max_tokens\": 2}",
              b"{\"messages\": [{\"role\": \"user\", \"content\": \"a\\u0000b\"}], \"max_tokens\": 2}",
              b"{\"messages\": [{\"role\": \"user\", \"content\": \"hi\"}], \"max_tokens\": 1e400}",
              b"{\"messages\": [{\"role\": \"user\", \"content\": \"hi\"}], \"seed\": 99999999999999999999999, \"max_tokens\": 2}",
              b"{\"messages\": [{\"role\": \"user\", \"content\": \"hi\"}], \"temperature\": \"hot\"}",
              b"{\"messages\": [{\"role\": 5, \"content\": \"hi\"}]}", b"\xff\xfe\x00\x01"]
    for _ in range(150):
        b = bytearray(base)
        kind = rnd.randrange(4)
        if kind == 0:
            b = b[:rnd.randrange(len(b))]
        elif kind == 1:
            for _ in range(rnd.randrange(1, 4)):
                b[rnd.randrange(len(b))] = rnd.randrange(256)
        elif kind == 2:
            i = rnd.randrange(len(b))
            b[i:i] = bytes(rnd.randrange(32, 127) for _ in range(rnd.randrange(1, 6)))
        else:
            i, j = sorted((rnd.randrange(len(b)), rnd.randrange(len(b))))
            del b[i:j]
        bodies.append(bytes(b))
    codes = {}
    try:
        for body in bodies:
            req = urllib.request.Request(f"http://127.0.0.1:{srv.port}/v1/chat/completions", data=body, method="POST")
            try:
                with urllib.request.urlopen(req, timeout=60) as r:
                    d = json.load(r)
                    assert d["object"] == "chat.completion"
                    codes[200] = codes.get(200, 0) + 1
            except urllib.error.HTTPError as e:
                assert e.code == 400 and "error" in json.load(e), (e.code, body)
                codes[400] = codes.get(400, 0) + 1
        assert codes.get(400, 0) > 50 and codes.get(200, 0) >= 3, codes
        st = srv.stats()
        assert st["running"] == 0 and st["waiting"] == 0 and st["free_slots"] == 4 and st["engine_errors"] == 0
    finally:
        srv.close()


def test_native_server_sheds_load_when_the_admission_queue_is_full(tok, monkeypatch):
    """DLLAMA_MAX_WAITING bounds the admission queue: requests beyond it are answered 503 at once instead of
    queueing without limit; the admitted ones finish normally."""
    monkeypatch.setenv("DLLAMA_MAX_WAITING", "2")
    eng = _FakeEngine(max_rows=4, n_slots=1, seq_len=256)
    real = eng.decode_step_greedy

    def slow(*a, **kw):
        time.sleep(0.02)
        return real(*a, **kw)

    eng.decode_step_greedy = slow
    defaults = A.InferenceParams(messages=[], temperature=0.0, top_p=0.9, seed=1)
    srv = NativeApiServer(eng, tok, "fake.m", defaults, host="127.0.0.1", port=0, verbose=False)
    srv.start()
    codes = []

    def user():
        try:
            with _post(srv.port, {"messages": [{"role": "user", "content": "hello"}], "max_tokens": 20}) as r:
                json.load(r)
                codes.append(200)
        except urllib.error.HTTPError as e:
            codes.append(e.code)
            assert "overloaded" in json.load(e)["error"]["message"]

    try:
        threads = [threading.Thread(target=user) for _ in range(12)]
        [t.start() for t in threads]
        [t.join(timeout=120) for t in threads]
        assert len(codes) == 12 and set(codes) == {200, 503} and 2 <= codes.count(200) <= 6, codes
        st = srv.stats()
        assert st["free_slots"] == 1 and st["waiting"] == 0 and st["running"] == 0
    finally:
        srv.close()
