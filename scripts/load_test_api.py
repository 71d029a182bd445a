#!/usr/bin/env python3
"""HTTP-level load test of `dllama-api` (BASELINE.json config 3): U concurrent users POST streaming
chat completions at a server running in this process (engine thread + threaded HTTP front end);
reports aggregate generated tok/s over the wall-clock window and the p50 / p95 time to the first
streamed token as each client sees it.

    python scripts/load_test_api.py --users 32 --max-tokens 64 [--model llama3.1-8b]
    torchrun --nproc-per-node N ... scripts/load_test_api.py ...      (workers replay rank 0's batches)

Random-init weights and a synthetic vocabulary of the model's size (no network): the text is noise,
the work is real.  temperature 0 -> device argmax (no logits leave the GPU); --temperature 0.8 ->
host sampling of every row (D2H of the logits)."""
import argparse
import json
import os
import sys
import threading
import time
import urllib.error
import urllib.request

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


class FakeEngine:
    """Engine surface with a fixed step time and no model: measures the front end alone (no GPU needed).
    `time.sleep` stands in for the GPU step (it releases the interpreter lock like a stream synchronize)."""

    class _Pages:
        def release_pages(self, slot):
            pass

    class _H:
        vocab_size = 0

    def __init__(self, vocab, max_rows, n_slots, seq_len, step_ms):
        self.h = self._H()
        self.h.vocab_size = vocab
        self.max_rows, self.n_slots, self.seq_len = max_rows, n_slots, seq_len
        self.device = torch.device("cpu")
        self.pages = self._Pages()
        self.step_s = step_ms * 1e-3
        self._k = 0

    def forward(self, tokens, positions, slots, logits_from=0, greedy=False):
        time.sleep(self.step_s)
        n = len(tokens) - logits_from
        lg = torch.zeros(n, self.h.vocab_size)
        for i in range(n):
            lg[i, 5 + (positions[logits_from + i] * 7 + slots[logits_from + i]) % 900] = 10.0
        return lg

    def decode_step_greedy(self, tokens, positions, slots, logits_from=0):
        time.sleep(self.step_s)
        return [5 + (positions[i] * 7 + slots[i]) % 900 for i in range(logits_from, len(tokens))]

    def close(self):
        pass


def _client_proc(port, users, first_user, max_tokens, prompt_words, temperature, q, go):
    """U streaming users of one client process (own interpreter: the server's engine thread is not slowed by them)."""
    words = ["the", "of", "and", "hello", "world", "llama", "token", "model", "fast", "gpu", "blackwell", "tensor"]
    ttft = [None] * users
    ntok = [0] * users

    def user(k):
        i = first_user + k
        text = " ".join(words[(i * 7 + j) % len(words)] for j in range(prompt_words))
        body = {"messages": [{"role": "user", "content": text}], "max_tokens": max_tokens, "stream": True,
                "temperature": temperature, "seed": 100 + i, "stream_options": {"include_usage": True}}
        req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/chat/completions", data=json.dumps(body).encode(),
                                     headers={"Content-Type": "application/json"}, method="POST")
        t0 = time.perf_counter()
        try:
            r = urllib.request.urlopen(req, timeout=600)
        except urllib.error.HTTPError as e:            # e.g. 400: the prompt does not fit --max-seq-len
            sys.stderr.write(f"[client {i}] HTTP {e.code}: {e.read()[:200]!r}\n")
            return
        with r:
            buf = b""
            while True:
                chunk = r.read1(65536)
                if not chunk:
                    break
                buf += chunk
                while b"\r\n\r\n" in buf:
                    ev, buf = buf.split(b"\r\n\r\n", 1)
                    if not ev.startswith(b"data: ") or ev == b"data: [DONE]":
                        continue
                    if ttft[k] is None and b'"content": ""' not in ev and b'"content"' in ev:
                        ttft[k] = time.perf_counter() - t0         # first non-empty text delta
                    if b'"usage"' in ev:
                        ntok[k] = json.loads(ev[6:])["usage"]["completion_tokens"]

    threads = [threading.Thread(target=user, args=(k,)) for k in range(users)]
    # first use of urllib in a fresh interpreter imports http.client, email, socket ... (60+ ms): keep that out of
    # the users' time to first token
    urllib.request.urlopen(f"http://127.0.0.1:{port}/health", timeout=30).read()
    q.put("ready")
    go.wait()                      # interpreter start-up (imports) stays outside the timed window
    [t.start() for t in threads]
    [t.join() for t in threads]
    q.put((ttft, sum(ntok)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.1-8b")
    ap.add_argument("--model-dir", default="/dev/shm")
    ap.add_argument("--users", type=int, default=32)
    ap.add_argument("--max-tokens", type=int, default=64)
    ap.add_argument("--prompt-words", type=int, default=24)
    ap.add_argument("--temperature", type=float, default=0.0)
    ap.add_argument("--max-batch", type=int, default=64)
    ap.add_argument("--max-seq-len", type=int, default=1024)
    ap.add_argument("--rounds", type=int, default=2, help="round 0 warms up (graph capture), later rounds are timed")
    ap.add_argument("--backend", default="native", choices=["native", "python"],
                    help="front end: csrc/host/server.cpp (default) or serving/api.py")
    ap.add_argument("--engine", default="cuda", choices=["cuda", "fake"],
                    help="fake = no model, fixed step time: measures the front end alone on any machine")
    ap.add_argument("--fake-step-ms", type=float, default=3.2)
    ap.add_argument("--client-procs", type=int, default=4, help="client processes (0 = client threads in this process)")
    a = ap.parse_args()
    import multiprocessing as mp
    from dllama_b200.formats import mfile, tfile
    from dllama_b200.serving import api_types as A
    from dllama_b200.serving.api import ApiServer, ApiState
    from dllama_b200.serving.batch_engine import BatchEngine, worker_loop
    from dllama_b200.serving.native_server import NativeApiServer
    from dllama_b200.text import native as T
    ctl = None
    if a.engine == "fake":
        rank, world = 0, 1
        eng = FakeEngine(32000, a.max_batch, a.users + 2, a.max_seq_len, a.fake_step_ms)
        path = "fake.m"
    else:
        from dllama_b200.models.llama import LlamaEngine
        from dllama_b200.parallel.symm import init_distributed
        rank, world, _ = init_distributed()
        path = os.path.join(a.model_dir, f"dllama_b200_{a.model}_q40.m")
        if rank == 0 and not os.path.exists(path):
            mfile.write_random_model(path, a.model, seed=1)
        if world > 1:
            torch.distributed.barrier()
        eng = LlamaEngine(path, rank=rank, world=world, max_rows=a.max_batch, n_slots=a.users + 2, max_seq_len=a.max_seq_len,
                          bucket_rows=True)
        if world > 1:
            from dllama_b200.parallel.control import ControlChannel
            ctl = ControlChannel(os.environ.get("MASTER_PORT", "0"), rank, world, a.max_batch)
        if rank != 0:
            worker_loop(eng, ctl)
            return
    tpath = os.path.join(a.model_dir, f"dllama_b200_synth_{eng.h.vocab_size}.t")
    if not os.path.exists(tpath):
        tfile.write_synthetic_tokenizer(tpath, eng.h.vocab_size)
    tok = T.Tokenizer(tpath)
    defaults = A.InferenceParams(messages=[], temperature=a.temperature, top_p=0.9, seed=1)
    t0 = time.perf_counter()
    if a.backend == "native":
        srv = NativeApiServer(eng, tok, os.path.basename(path), defaults, control=ctl, host="127.0.0.1", port=0,
                              max_default_tokens=a.max_tokens, verbose=False)
        srv.warmup()
        srv.start()
        be = None
        counters = lambda: (lambda st: (st["generated"], st["steps"], st["rows"]))(srv.stats())
    else:
        be = BatchEngine(eng, tok, control=ctl)
        be.warmup()
        be.start()
        state = ApiState(be, tok, os.path.basename(path), defaults, max_default_tokens=a.max_tokens)
        srv = ApiServer(state, "127.0.0.1", 0)
        srv.start()
        counters = lambda: (be.n_generated, be.n_steps, be.n_rows)
    print(f"graph warm-up: {time.perf_counter() - t0:.2f} s ({len(getattr(eng, '_graphs', {}))} graphs)", flush=True)
    out = {}
    ctx = mp.get_context("spawn")
    for rnd in range(a.rounds):
        c0 = counters()
        q = ctx.Queue() if a.client_procs > 0 else __import__("queue").Queue()
        nproc = max(1, min(a.client_procs, a.users))
        shares = [a.users // nproc + (1 if k < a.users % nproc else 0) for k in range(nproc)]
        firsts = [sum(shares[:k]) for k in range(nproc)]
        mk = ctx.Process if a.client_procs > 0 else threading.Thread
        go = ctx.Event() if a.client_procs > 0 else threading.Event()
        procs = [mk(target=_client_proc, args=(srv.port, shares[k], firsts[k], a.max_tokens, a.prompt_words,
                                               a.temperature, q, go)) for k in range(nproc)]
        [p.start() for p in procs]
        for _ in procs:
            assert q.get(timeout=300) == "ready"
        t0 = time.perf_counter()
        go.set()
        got = [q.get(timeout=900) for _ in procs]
        wall = time.perf_counter() - t0
        [p.join() for p in procs]
        log = getattr(srv, "step_log", None)
        if log and rnd == a.rounds - 1:
            mine = [e for e in log if e[0] >= t0]
            print("first steps of the last round (ms after go, rows, first logits row, sampled rows, step ms):")
            for e in mine[:24]:
                print(f"   {1e3 * (e[0] - t0):8.2f}  rows {e[1]:5d}  logits_from {e[2]:5d}  sampled {e[3]:3d}  {1e3 * e[4]:7.2f} ms")
        tt = sorted(x for g in got for x in g[0] if x is not None)
        c1 = counters()
        total = c1[0] - c0[0]              # server-side counter (also exported at /metrics)
        assert total == sum(g[1] for g in got), (total, [g[1] for g in got])
        out = {"metric": "dllama-api aggregate generated tok/s (HTTP, streaming)", "backend": a.backend, "engine": a.engine,
               "users": a.users, "world": world, "client_procs": a.client_procs,
               "max_tokens": a.max_tokens, "temperature": a.temperature, "round": rnd, "wall_s": round(wall, 3),
               "completion_tokens": total, "aggregate_tokens_per_s": round(total / wall, 1),
               "ttft_p50_ms": round(tt[len(tt) // 2] * 1e3, 2) if tt else None,
               "ttft_p95_ms": round(tt[min(len(tt) - 1, int(len(tt) * 0.95))] * 1e3, 2) if tt else None,
               "steps": c1[1] - c0[1], "rows": c1[2] - c0[2],
               "ms_per_step": round(wall * 1e3 / max(1, c1[1] - c0[1]), 3)}
        if a.engine == "fake":
            out["fake_step_ms"] = a.fake_step_ms
        if be is not None:
            out["engine_thread"] = {"sched_s": round(be.t_sched, 3), "engine_s": round(be.t_engine, 3), "emit_s": round(be.t_emit, 3)}
        print(json.dumps(out), flush=True)
    if a.backend == "native":
        srv.close()
    else:
        srv.shutdown()
        be.stop()
    eng.close()


if __name__ == "__main__":
    main()
