"""The native `dllama-api` front end (csrc/host/server.cpp) driving the CUDA engine: bucketed batch shapes with
pre-captured graphs, device greedy pick, device sampler fed by the server's per-request coins.  Collected late
on purpose (it exercises the whole serving stack on top of everything the earlier files test)."""
import json
import threading
import urllib.request

import pytest

from dllama_b200.formats import tfile
from dllama_b200.serving import api_types as A
from dllama_b200.text import native as T


def _post(port, body, timeout=180):
    req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/chat/completions", data=json.dumps(body).encode(),
                                 headers={"Content-Type": "application/json"}, method="POST")
    with urllib.request.urlopen(req, timeout=timeout) as r:
        return json.load(r)


@pytest.mark.gpu
def test_native_server_on_the_cuda_engine(tiny_model_path, tmp_path):
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.serving.native_server import NativeApiServer
    tp = str(tmp_path / "tiny.t")
    tfile.write_synthetic_tokenizer(tp, 1024)
    tok = T.Tokenizer(tp)
    eng = LlamaEngine(tiny_model_path, max_rows=32, n_slots=8, max_seq_len=256, bucket_rows=True)
    defaults = A.InferenceParams(messages=[], temperature=0.0, top_p=0.9, seed=1)
    srv = NativeApiServer(eng, tok, "tiny.m", defaults, host="127.0.0.1", port=0, verbose=False)
    srv.warmup()
    n_graphs = len(eng._graphs)
    srv.start()
    port = srv.port
    try:
        texts = ["hello world", "the of and hello world llama token model fast gpu " * 3, "hello world", "token"]
        out = {}

        def user(i):
            out[i] = _post(port, {"messages": [{"role": "user", "content": texts[i]}], "max_tokens": 5})

        def sampled(key, seed):
            out[key] = _post(port, {"messages": [{"role": "user", "content": "hello world"}], "max_tokens": 6,
                                    "temperature": 1.0, "top_p": 0.95, "seed": seed})

        threads = [threading.Thread(target=user, args=(i,)) for i in range(4)]
        threads += [threading.Thread(target=sampled, args=("c", 12))]
        [t.start() for t in threads]
        [t.join(timeout=300) for t in threads]
        # alone, one after the other: identical batch shapes, so identical kernels and bit-identical logits
        sampled("a", 11)
        sampled("b", 11)
        user(0)
        assert len(out) == 7
        for d in out.values():
            assert d["object"] == "chat.completion" and d["usage"]["completion_tokens"] >= 1
            assert d["choices"][0]["finish_reason"] in ("stop", "length")
        # same seed -> same sample path on the device sampler
        assert out["a"]["choices"][0]["message"]["content"] == out["b"]["choices"][0]["message"]["content"]
        # the server tokenises the templated prompt exactly like the Python text runtime
        tmpl = T.ChatTemplateGenerator(T.CHAT_UNKNOWN, tok.chat_template, tok.piece(tok.eos_token_ids[-1]).decode())
        prompt = tok.encode(tmpl.generate([("user", texts[0])], True)[0], add_bos=True, add_special_tokens=True)
        assert out[0]["usage"]["prompt_tokens"] == len(prompt)
        with urllib.request.urlopen(urllib.request.Request(
                f"http://127.0.0.1:{port}/v1/chat/completions", method="POST",
                data=json.dumps({"messages": [{"role": "user", "content": texts[0]}], "max_tokens": 5,
                                 "stream": True}).encode()), timeout=180) as r:
            raw = r.read().decode("utf-8")
        streamed = "".join(json.loads(e[6:])["choices"][0]["delta"].get("content", "")
                           for e in raw.split("\r\n\r\n") if e.startswith("data: {"))
        assert streamed == out[0]["choices"][0]["message"]["content"]
        assert len(eng._graphs) == n_graphs                         # no shape outside the warmed-up buckets
        st = srv.stats()
        assert st["running"] == 0 and st["waiting"] == 0 and st["free_slots"] == 8 and st["engine_errors"] == 0
    finally:
        srv.close()
        eng.close()


@pytest.mark.gpu
def test_back_to_back_prompt_chunks_behind_a_busy_stream(tiny_model_path):
    """Prompt chunks are launched without a host synchronisation between them, so the async copy of a chunk's batch
    descriptor can execute long after it was enqueued.  With the stream kept busy, a descriptor staging buffer that
    is rewritten too early makes every chunk run with the last chunk's rows; the engine has to hold the buffer until
    the copy has been consumed (LlamaEngine.set_batch).  Checked against the CPU oracle."""
    import numpy as np
    import torch
    from dllama_b200.formats.mfile import ModelFile
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.models.oracle import OracleLlama
    o = OracleLlama(ModelFile(tiny_model_path), max_seq_len=512)
    e = LlamaEngine(tiny_model_path, max_rows=32, n_slots=3, max_seq_len=512)
    rng = np.random.default_rng(11)
    warm = rng.integers(1, 1024, 96).tolist()
    e.prefill(e.alloc_slot(), warm, 0, chunk=32)                    # captures the graphs of this batch shape
    torch.cuda.synchronize()
    prompt = rng.integers(1, 1024, 96).tolist()
    slot = e.alloc_slot()
    torch.cuda._sleep(int(3e8))                                     # ~0.15 s of work queued in front of the chunks
    got = e.prefill(slot, prompt, 0, chunk=32).float().cpu().numpy()
    ref = o.forward(prompt, 0).numpy()
    err = float(np.abs(got[0] - ref[-1]).max() / (np.abs(ref[-1]).max() + 1e-9))
    assert err < 3e-2, err
    # and the sequence's KV state is intact: one more token still agrees with the oracle
    tok = int(np.argmax(ref[-1]))
    got1 = e.forward([tok], [96], [slot]).float().cpu().numpy()
    ref1 = o.forward([tok], 96).numpy()
    assert float(np.abs(got1[0] - ref1[0]).max() / (np.abs(ref1[0]).max() + 1e-9)) < 3e-2
    e.close()
