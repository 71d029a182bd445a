"""Scheduler, continuous batching and the HTTP API on the CPU engine (no GPU needed)."""
import json
import threading
import urllib.request

import pytest

from dllama_b200.formats import mfile, tfile
from dllama_b200.models.cpu_engine import CpuEngine
from dllama_b200.models.oracle import OracleLlama
from dllama_b200.serving import api_types as A
from dllama_b200.serving.api import ApiServer, ApiState
from dllama_b200.serving.batch_engine import BatchEngine, GenerationRequest
from dllama_b200.text import native as T


@pytest.fixture(scope="module")
def tok(tmp_path_factory):
    p = str(tmp_path_factory.mktemp("tok") / "tiny.t")
    tfile.write_synthetic_tokenizer(p, 1024)
    return T.Tokenizer(p)


def test_scheduler_batches_prefill_first_and_decodes_together():
    s = T.Scheduler(max_rows=8, n_slots=2, prefill_chunk=4, max_seq_len=64)
    assert s.submit(1, [10, 11, 12, 13, 14, 15], 3) == 0
    assert s.submit(2, [20, 21], 2) == 0
    assert s.submit(3, [30], 2) == 0                      # waits: only 2 slots
    assert s.submit(4, [], 2) == -3 and s.submit(5, list(range(64)), 2) == -1 and s.submit(1, [1], 1) == -2
    # step 1: 4 of request 1's 6 prompt tokens (prefill_chunk) + all of request 2's prompt; the row that
    # completes prompt 2 is moved to the tail so the rows needing logits are one contiguous range
    ids, toks, pos, slots, lf, samp = s.next_batch()
    assert ids == [1, 1, 1, 1, 2, 2] and toks == [10, 11, 12, 13, 20, 21] and pos == [0, 1, 2, 3, 0, 1]
    assert lf == 5 and samp == [2]
    assert slots[0] == slots[3] != slots[4] == slots[5]
    assert s.commit([200], [False]) == [False]
    # step 2: rest of request 1 (its last row samples) + request 2 decoding
    ids, toks, pos, slots, lf, samp = s.next_batch()
    assert ids == [1, 1, 2] and toks == [14, 15, 200] and pos == [4, 5, 2] and lf == 1 and samp == [1, 2]
    assert s.commit([100, 201], [False, False]) == [False, True]      # request 2's budget (2 tokens) is spent
    assert len(s.pop_released()) == 1
    # step 3: request 3 admitted into the freed slot, prefilled while request 1 decodes
    ids, toks, pos, slots, lf, samp = s.next_batch()
    assert ids == [3, 1] and toks == [30, 100] and pos == [0, 6] and lf == 0 and samp == [3, 1]
    assert s.commit([7, 101], [True, False]) == [True, False]          # stop flag ends request 3 at once
    ids, toks, pos, slots, lf, samp = s.next_batch()
    assert ids == [1] and toks == [101] and pos == [7]
    assert s.commit([102], [False]) == [True]                           # 3 new tokens: done
    assert s.next_batch()[0] == []
    s.pop_released()
    assert s.stats()["free_slots"] == 2


def test_scheduler_admits_by_free_kv_pages():
    """Page-aware admission: a request holds the pages of its whole life (prompt + max_new tokens) while it runs, so
    the engine's page allocator can never run dry in the middle of a step; requests that do not fit wait (FIFO), and
    one that could never fit is refused at submit."""
    s = T.Scheduler(max_rows=64, n_slots=4, prefill_chunk=64, max_seq_len=512)
    s.set_pages(6, 32)                                     # 6 pages of 32 tokens
    assert s.submit(1, list(range(40)), 60) == 0           # 100 tokens -> 4 pages
    assert s.submit(2, list(range(20)), 40) == 0           # 60 tokens  -> 2 pages
    assert s.submit(3, [1, 2, 3], 5) == 0                  # 1 page: free slots exist, but no pages are left
    assert s.submit(4, list(range(100)), 200) == -1        # 10 pages: can never be admitted
    ids, *_ = s.next_batch()
    assert set(ids) == {1, 2} and s.stats()["waiting"] == 1
    # finish request 2: its pages come back and request 3 gets in
    while True:
        ids, toks, pos, slots, lf, samp = s.next_batch()
        if not samp:
            s.commit([], [])
            continue
        done = s.commit([7] * len(samp), [r == 2 for r in samp])
        if any(d and r == 2 for d, r in zip(done, samp)):
            break
    s.pop_released()
    ids, *_ = s.next_batch()
    assert 3 in ids and s.stats()["waiting"] == 0


def test_scheduler_cancel_and_round_robin():
    s = T.Scheduler(max_rows=2, n_slots=3, prefill_chunk=2, max_seq_len=64)
    for i in (1, 2, 3):
        s.submit(i, [i], 10)
    while True:                                           # prefill (2 rows per step) until all three decode
        ids, toks, pos, slots, lf, samp = s.next_batch()
        s.commit([50 + x for x in samp], [False] * len(samp))
        if s.stats()["running"] == 3 and lf == 0 and len(samp) == len(ids) and len(ids) == 2:
            break
    seen = []
    for _ in range(3):                                    # 3 decoders, 2 rows: round-robin must serve all
        ids, *_rest, samp = s.next_batch()
        seen += ids
        s.commit([60] * len(samp), [False] * len(samp))
    assert set(seen) == {1, 2, 3}
    s.cancel(2)
    ids = s.next_batch()[0]
    assert 2 not in ids
    assert s.pop_released() != []


def _greedy_reference(path, prompt, n_new):
    o = OracleLlama(mfile.ModelFile(path), max_seq_len=256)
    return o.generate_greedy(prompt, n_new)


def test_continuous_batching_matches_single_user_runs(tiny_model_path, tok):
    eng = CpuEngine(tiny_model_path, max_rows=8, n_slots=3, max_seq_len=256)
    be = BatchEngine(eng, tok, prefill_chunk=4)
    prompts = [[5, 6, 7, 8, 9, 10], [100, 200, 300], [9, 8, 7, 6, 5, 4, 3, 2, 1], [42]]
    reqs = [be.submit(GenerationRequest(prompt_tokens=p, max_tokens=5, temperature=0.0)) for p in prompts]
    be.start()
    for r in reqs:
        r.result(timeout=120)
    be.stop()
    for r, p in zip(reqs, prompts):
        want = _greedy_reference(tiny_model_path, p, 5)
        eos = [i for i, t in enumerate(want) if tok.is_eos(t)]
        if eos:
            want = want[:eos[0] + 1]
        assert r.tokens == want, (p, r.tokens, want)
        assert r.finish_reason in ("length", "stop") and r.t_first_token is not None
    assert be.sched.stats() == {"waiting": 0, "running": 0, "free_slots": 3}


@pytest.mark.gpu
def test_gpu_engine_behind_the_batch_engine(tiny_model_path, tok):
    """The serving path on the CUDA engine: power-of-two batch buckets + pre-captured graphs, several prompts
    packed per step (one long enough for the prefill attention), device greedy pick and device sampler."""
    from dllama_b200.models.llama import LlamaEngine
    eng = LlamaEngine(tiny_model_path, max_rows=32, n_slots=6, max_seq_len=256, bucket_rows=True)
    be = BatchEngine(eng, tok)
    be.warmup()
    n_graphs = len(eng._graphs)
    assert n_graphs >= 12
    rng = __import__("numpy").random.default_rng(4)
    prompts = [[5, 6, 7, 8, 9, 10], [100, 200, 300], rng.integers(1, 1000, 40).tolist(), [42]]
    reqs = [be.submit(GenerationRequest(prompt_tokens=p, max_tokens=5, temperature=0.0)) for p in prompts]
    mk = lambda seed: GenerationRequest(prompt_tokens=[3, 4, 5], max_tokens=6, temperature=1.0, top_p=0.95, seed=seed)
    a, b, c = be.submit(mk(11)), be.submit(mk(11)), be.submit(mk(12))
    be.start()
    for r in reqs + [a]:
        r.result(timeout=120)
    for r in (b, c):
        r.result(timeout=120)
    be.stop()
    o = OracleLlama(mfile.ModelFile(tiny_model_path), max_seq_len=256)
    for r, p in zip(reqs, prompts):
        # greedy tokens: every generated token must be (near-)argmax of the oracle's logits given the same prefix
        assert 1 <= len(r.tokens) <= 5 and r.finish_reason in ("length", "stop")
        o.reset()
        lg = o.forward(p, 0).numpy()[-1]
        for i, t in enumerate(r.tokens):
            assert lg[t] >= lg.max() - 2e-2 * max(1.0, float(abs(lg).max())), (p, i, t)
            lg = o.forward([t], len(p) + i).numpy()[0]
    assert a.tokens == b.tokens and len(a.tokens) >= 1          # same seed -> same sample path on the device sampler
    assert len(eng._graphs) == n_graphs                         # no shape outside the warmed-up buckets
    assert be.sched.stats() == {"waiting": 0, "running": 0, "free_slots": 6}
    eng.close()


def test_sampling_is_per_request_and_seeded(tiny_model_path, tok):
    eng = CpuEngine(tiny_model_path, max_rows=8, n_slots=4, max_seq_len=256)
    be = BatchEngine(eng, tok)
    mk = lambda seed: GenerationRequest(prompt_tokens=[3, 4, 5], max_tokens=6, temperature=1.0, top_p=0.95, seed=seed)
    a, b, c = be.submit(mk(11)), be.submit(mk(11)), be.submit(mk(12))
    be.start()
    for r in (a, b, c):
        r.result(timeout=120)
    be.stop()
    assert a.tokens == b.tokens                 # same seed, same prompt -> same sample path, even when batched
    assert a.tokens != c.tokens


def test_stop_string_truncates_output(tiny_model_path, tok):
    eng = CpuEngine(tiny_model_path, max_rows=4, n_slots=2, max_seq_len=256)
    be = BatchEngine(eng, tok)
    free = be.submit(GenerationRequest(prompt_tokens=[5, 6, 7], max_tokens=12, temperature=0.0))
    be.start()
    full = free.result(timeout=120)
    assert len(full) >= 4
    stop = full[2:4]
    r = be.submit(GenerationRequest(prompt_tokens=[5, 6, 7], max_tokens=12, temperature=0.0, stop=[stop]))
    got = r.result(timeout=120)
    be.stop()
    assert got == full[:full.index(stop)] and r.finish_reason == "stop"


def test_engine_failure_ends_the_batch_but_not_the_server(tiny_model_path, tok):
    """A forward that raises (KV pool exhausted, CUDA error...) ends the requests of that batch with
    finish_reason "error"; the engine thread survives and serves the next request."""
    eng = CpuEngine(tiny_model_path, max_rows=8, n_slots=4, max_seq_len=256)
    real_forward = eng.forward
    calls = {"n": 0}

    def flaky(*a, **kw):
        calls["n"] += 1
        if calls["n"] == 2:
            raise RuntimeError("KV cache exhausted")
        return real_forward(*a, **kw)

    eng.forward = flaky
    be = BatchEngine(eng, tok)
    be.start()
    try:
        r1 = be.submit(GenerationRequest(prompt_tokens=[5, 6, 7], max_tokens=6, temperature=0.0))
        while r1.out.get(timeout=60) is not None:
            pass
        assert r1.finish_reason == "error" and "exhausted" in r1.error
        r2 = be.submit(GenerationRequest(prompt_tokens=[5, 6, 7], max_tokens=4, temperature=0.0))
        while r2.out.get(timeout=60) is not None:
            pass
        assert r2.finish_reason in ("length", "stop") and len(r2.tokens) >= 1
        assert be.n_errors == 1 and be.sched.stats()["free_slots"] == 4
    finally:
        be.stop()


def _post(port, body, timeout=120):
    req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/chat/completions", data=json.dumps(body).encode(),
                                 headers={"Content-Type": "application/json"}, method="POST")
    return urllib.request.urlopen(req, timeout=timeout)


def test_http_api_concurrent_users_openai_shape_and_sse(tiny_model_path, tok):
    eng = CpuEngine(tiny_model_path, max_rows=8, n_slots=4, max_seq_len=256)
    be = BatchEngine(eng, tok)
    be.start()
    state = ApiState(be, tok, "tiny.m", A.InferenceParams(messages=[], temperature=0.0, top_p=0.9, seed=1))
    srv = ApiServer(state, "127.0.0.1", 0)
    srv.start()
    port = srv.port
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
        # three users at once; different prompts must give different, self-consistent answers
        results = {}

        def user(i, text):
            with _post(port, {"messages": [{"role": "user", "content": text}], "max_tokens": 6}) as r:
                results[i] = json.load(r)

        threads = [threading.Thread(target=user, args=(i, t)) for i, t in
                   enumerate(["hello world", "the of and", "hello world"])]
        [t.start() for t in threads]
        [t.join(timeout=180) for t in threads]
        assert len(results) == 3
        for d in results.values():
            assert d["object"] == "chat.completion" and d["choices"][0]["message"]["role"] == "assistant"
            assert d["usage"]["completion_tokens"] <= 6
            assert d["usage"]["total_tokens"] == d["usage"]["prompt_tokens"] + d["usage"]["completion_tokens"]
            assert d["generated_text"] == d["choices"][0]["message"]["content"]
        assert results[0]["choices"][0]["message"]["content"] == results[2]["choices"][0]["message"]["content"]
        assert results[0]["usage"]["prompt_tokens"] != results[1]["usage"]["prompt_tokens"] or \
            results[0]["choices"][0]["message"]["content"] != results[1]["choices"][0]["message"]["content"]
        # streaming: SSE chunks concatenate to the non-streamed answer
        with _post(port, {"messages": [{"role": "user", "content": "hello world"}], "max_tokens": 6,
                          "stream": True}) as r:
            assert r.headers["Content-Type"].startswith("text/event-stream")
            raw = r.read().decode("utf-8")
        events = [e[len("data: "):] for e in raw.split("\r\n\r\n") if e.startswith("data: ")]
        assert events[-1] == "[DONE]"
        chunks = [json.loads(e) for e in events[:-1]]
        assert all(c["object"] == "chat.completion.chunk" for c in chunks)
        text = "".join(c["choices"][0]["delta"].get("content", "") for c in chunks)
        assert text == results[0]["choices"][0]["message"]["content"]
        assert "usage" not in chunks[-1]
        # stream_options.include_usage: the last chunk carries the token counts
        with _post(port, {"messages": [{"role": "user", "content": "hello world"}], "max_tokens": 6, "stream": True,
                          "stream_options": {"include_usage": True}}) as r:
            raw = r.read().decode("utf-8")
        last = json.loads([e[len("data: "):] for e in raw.split("\r\n\r\n") if e.startswith("data: ")][-2])
        assert last["usage"] == results[0]["usage"] and last["choices"][0]["finish_reason"] in ("stop", "length")
        assert chunks[-1]["choices"][0]["finish_reason"] in ("stop", "length")
        # bad request
        with pytest.raises(urllib.error.HTTPError) as ei:
            _post(port, {"messages": "nope"})
        assert ei.value.code == 400
        with urllib.request.urlopen(f"http://127.0.0.1:{port}/metrics") as r:
            assert b"dllama_requests_total 5" in r.read()
    finally:
        srv.shutdown()
        be.stop()


def test_scheduler_randomised_simulation_keeps_its_invariants():
    """A few hundred requests with random prompt lengths, budgets, stop flags and cancellations driven through the C++
    scheduler: every step respects the row budget, gives every sequence strictly consecutive positions starting at 0,
    never shares a slot between live requests, keeps the sampling rows a contiguous tail, and in the end every request
    is finished and every slot is back."""
    import random
    rnd = random.Random(11)
    R, S, CH, L = 16, 5, 6, 96
    s = T.Scheduler(max_rows=R, n_slots=S, prefill_chunk=CH, max_seq_len=L)
    prompts, budget, next_pos, fed, produced, done, cancelled = {}, {}, {}, {}, {}, set(), set()
    next_id, steps = 1, 0
    slot_owner = {}
    while steps < 4000:
        steps += 1
        if next_id <= 300 and rnd.random() < 0.5:
            n = rnd.choice([1, 2, 3, 7, 20, 50, 94])
            prompts[next_id] = [rnd.randrange(1, 1000) for _ in range(n)]
            budget[next_id] = rnd.choice([1, 2, 5, 30])
            rc = s.submit(next_id, prompts[next_id], budget[next_id])
            assert rc == 0
            next_pos[next_id], fed[next_id], produced[next_id] = 0, [], 0
            next_id += 1
        live = [i for i in prompts if i not in done and i not in cancelled]
        if live and rnd.random() < 0.05:
            victim = rnd.choice(live)
            s.cancel(victim)
            cancelled.add(victim)
        ids, toks, pos, slots, lf, samp = s.next_batch()
        for sl in s.pop_released():
            slot_owner.pop(sl, None)
        if not ids:
            if next_id > 300 and not [i for i in prompts if i not in done and i not in cancelled]:
                break
            continue
        assert len(ids) <= R and 0 <= lf <= len(ids)
        assert ids[lf:] == samp and len(set(samp)) == len(samp)               # sampling rows = contiguous tail, one per request
        seen_rows = {}
        for i, t, p, sl in zip(ids, toks, pos, slots):
            assert i not in cancelled or True
            assert p == next_pos[i], (i, p, next_pos[i])                      # consecutive positions per sequence
            next_pos[i] += 1
            assert p < L
            assert slot_owner.setdefault(sl, i) == i, "slot shared by two live requests"
            fed[i].append(t)
            seen_rows[i] = seen_rows.get(i, 0) + 1
        for i, c in seen_rows.items():
            assert c <= CH or len(fed[i]) > len(prompts[i])                    # at most prefill_chunk prompt rows per step
            k = min(len(fed[i]), len(prompts[i]))
            assert fed[i][:k] == prompts[i][:k]                                # the prompt is fed in order
        sampled = [rnd.randrange(1, 1000) for _ in samp]
        stops = [rnd.random() < 0.1 for _ in samp]
        fin = s.commit(sampled, stops)
        for i, tok_id, st, f in zip(samp, sampled, stops, fin):
            produced[i] += 1
            assert len(fed[i]) >= len(prompts[i])                              # a request samples only after its whole prompt
            if st:
                assert f
            if f:
                done.add(i)
                s.forget(i)
            else:
                assert produced[i] < budget[i]
                prompts[i] = prompts[i] + [tok_id]                             # the sampled token is fed next
        for i in list(samp):
            if i in done:
                assert produced[i] <= max(budget[i], 1)
    assert next_id > 300 and steps < 4000
    s.next_batch()
    s.pop_released()
    assert s.stats() == {"waiting": 0, "running": 0, "free_slots": S}
    assert len(done) + len(cancelled) >= 300 and len(done) > 150


def test_engine_row_buckets_and_logits_tail():
    """Serving batch shapes (LlamaEngine._bucket / _buckets / _tail_rows, pure host logic): powers of two up to 256 rows,
    multiples of 256 above, never beyond max_rows; every bucket a batch can ask for is in the warm-up list; the logits
    tail covers one row per sequence slot and is only used when it saves at least half of the rows."""
    from types import SimpleNamespace
    from dllama_b200.models.llama import LlamaEngine
    for max_rows, n_slots in ((32, 8), (64, 34), (256, 34), (1000, 34), (2048, 34), (2048, 300)):
        e = SimpleNamespace(max_rows=max_rows, n_slots=n_slots)
        buckets = LlamaEngine._buckets(e)
        assert buckets == sorted(set(buckets)) and buckets[-1] == max_rows
        for M in range(1, max_rows + 1):
            Mp = LlamaEngine._bucket(e, M)
            assert M <= Mp <= max_rows and Mp in buckets, (max_rows, M, Mp)
            assert Mp < 2 * M or Mp <= 1 or Mp == max_rows or M <= 256          # padding never doubles a large batch
            if M > 256:
                assert Mp - M < 256
        for Mp in buckets:
            tail = LlamaEngine._tail_rows(e, Mp)
            assert tail == 0 or (tail >= min(n_slots, 256) and 2 * tail <= Mp and tail & (tail - 1) == 0)
