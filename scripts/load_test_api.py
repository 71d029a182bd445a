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
import urllib.request

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.1-8b")
    ap.add_argument("--model-dir", default="/dev/shm")
    ap.add_argument("--users", type=int, default=32)
    ap.add_argument("--max-tokens", type=int, default=64)
    ap.add_argument("--prompt-words", type=int, default=24)
    ap.add_argument("--temperature", type=float, default=0.0)
    ap.add_argument("--max-batch", type=int, default=64)
    ap.add_argument("--rounds", type=int, default=2, help="round 0 warms up (graph capture), later rounds are timed")
    a = ap.parse_args()
    from dllama_b200.formats import mfile, tfile
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.parallel.symm import init_distributed
    from dllama_b200.serving import api_types as A
    from dllama_b200.serving.api import ApiServer, ApiState
    from dllama_b200.serving.batch_engine import BatchEngine, worker_loop
    from dllama_b200.text import native as T
    rank, world, _ = init_distributed()
    path = os.path.join(a.model_dir, f"dllama_b200_{a.model}_q40.m")
    if rank == 0 and not os.path.exists(path):
        mfile.write_random_model(path, a.model, seed=1)
    if world > 1:
        torch.distributed.barrier()
    eng = LlamaEngine(path, rank=rank, world=world, max_rows=a.max_batch, n_slots=a.users + 2, max_seq_len=1024,
                      bucket_rows=True)
    ctl = None
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
    be = BatchEngine(eng, tok, control=ctl)
    t0 = time.perf_counter()
    be.warmup()
    print(f"graph warm-up: {time.perf_counter() - t0:.2f} s ({len(getattr(eng, '_graphs', {}))} graphs)", flush=True)
    be.start()
    state = ApiState(be, tok, os.path.basename(path), A.InferenceParams(messages=[], temperature=a.temperature, top_p=0.9, seed=1),
                     max_default_tokens=a.max_tokens)
    srv = ApiServer(state, "127.0.0.1", 0)
    srv.start()
    words = ["the", "of", "and", "hello", "world", "llama", "token", "model", "fast", "gpu", "blackwell", "tensor"]
    out = {}
    for rnd in range(a.rounds):
        ttft = [None] * a.users
        gen0 = be.n_generated
        st0 = (be.n_steps, be.n_rows, be.t_sched, be.t_engine, be.t_emit)

        def user(i):
            text = " ".join(words[(i * 7 + k) % len(words)] for k in range(a.prompt_words))
            body = {"messages": [{"role": "user", "content": text}], "max_tokens": a.max_tokens, "stream": True,
                    "temperature": a.temperature, "seed": 100 + i}
            req = urllib.request.Request(f"http://127.0.0.1:{srv.port}/v1/chat/completions", data=json.dumps(body).encode(),
                                         headers={"Content-Type": "application/json"}, method="POST")
            t0 = time.perf_counter()
            with urllib.request.urlopen(req, timeout=600) as r:
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
                        d = json.loads(ev[6:])
                        if d["choices"][0]["delta"].get("content"):      # first non-empty text delta (the role chunk is empty)
                            if ttft[i] is None:
                                ttft[i] = time.perf_counter() - t0

        t0 = time.perf_counter()
        threads = [threading.Thread(target=user, args=(i,)) for i in range(a.users)]
        [t.start() for t in threads]
        [t.join() for t in threads]
        wall = time.perf_counter() - t0
        tt = sorted(x for x in ttft if x is not None)
        total = be.n_generated - gen0      # server-side counter (also exported at /metrics)
        out = {"metric": "dllama-api aggregate generated tok/s (HTTP, streaming)", "users": a.users, "world": world,
               "max_tokens": a.max_tokens, "temperature": a.temperature, "round": rnd, "wall_s": round(wall, 3),
               "completion_tokens": total, "aggregate_tokens_per_s": round(total / wall, 1),
               "ttft_p50_ms": round(tt[len(tt) // 2] * 1e3, 2) if tt else None,
               "ttft_p95_ms": round(tt[min(len(tt) - 1, int(len(tt) * 0.95))] * 1e3, 2) if tt else None}
        out["engine_thread"] = {"steps": be.n_steps - st0[0], "rows": be.n_rows - st0[1], "sched_s": round(be.t_sched - st0[2], 3),
                                "engine_s": round(be.t_engine - st0[3], 3), "emit_s": round(be.t_emit - st0[4], 3)}
        print(json.dumps(out), flush=True)
    srv.shutdown()
    be.stop()
    eng.close()


if __name__ == "__main__":
    main()
