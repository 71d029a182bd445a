"""`dllama inference / chat` and the argument parser (reference: src/app.cpp:33-136, src/dllama.cpp): run
in-process on the CPU engine (and on a GPU with the CUDA engine) with a tiny random model and a
synthetic tokenizer; the output block must keep the reference's benchmark format."""
import io
import os
import re
import sys
from contextlib import redirect_stdout

import pytest

from dllama_b200.cli import dllama as cli
from dllama_b200.cli.app import AppArgs
from dllama_b200.formats import tfile


@pytest.fixture(scope="module")
def tok_path(tmp_path_factory):
    p = str(tmp_path_factory.mktemp("clitok") / "tiny.t")
    tfile.write_synthetic_tokenizer(p, 1024)
    return p


def _run(argv):
    buf = io.StringIO()
    with redirect_stdout(buf):
        rc = cli.main(argv)
    return rc, buf.getvalue()


def test_arg_parser_accepts_the_reference_flags():
    a = AppArgs.parse(["dllama", "inference", "--model", "m.m", "--tokenizer", "t.t", "--prompt", "hi", "--steps", "7",
                       "--buffer-float-type", "q80", "--nthreads", "4", "--max-seq-len", "128", "--temperature", "0.5",
                       "--topp", "0.8", "--seed", "42", "--workers", "127.0.0.1:9999", "10.0.0.2:9998", "--port", "9990",
                       "--net-turbo", "0", "--gpu-index", "1", "--gpu-segments", "0:3", "--chat-template", "llama3"], True)
    assert a.mode == "inference" and a.model_path == "m.m" and a.tokenizer_path == "t.t" and a.steps == 7
    assert a.max_seq_len == 128 and abs(a.temperature - 0.5) < 1e-6 and abs(a.topp - 0.8) < 1e-6 and a.seed == 42
    assert len(a.workers) == 2 and a.workers[0] == ("127.0.0.1", 9999) and a.workers[1] == ("10.0.0.2", 9998)
    assert a.port == 9990 and a.gpu_index == 1
    with pytest.raises(ValueError):
        AppArgs.parse(["dllama", "inference", "--no-such-flag", "1"], True)


def _check_inference_output(out: str, steps: int):
    assert re.search(r"Evaluation\n\s+nBatches: \d+\n\s+nTokens: \d+\n\s+tokens/s: [\d.]+", out), out
    assert re.search(r"Prediction\n\s+nTokens: \d+\n\s+tokens/s: [\d.]+", out), out
    assert len(re.findall(r"Pred\s+\d+ ms", out)) >= 1
    assert "Eval" in out


def test_inference_on_cpu_engine(tiny_model_path, tok_path):
    rc, out = _run(["dllama", "inference", "--model", tiny_model_path, "--tokenizer", tok_path, "--buffer-float-type", "q80",
                    "--prompt", "hello world the of and", "--steps", "12", "--max-seq-len", "128", "--device", "cpu", "--seed", "7"])
    assert rc == 0, out
    _check_inference_output(out, 12)
    rc2, out2 = _run(["dllama", "inference", "--model", tiny_model_path, "--tokenizer", tok_path, "--buffer-float-type", "q80",
                      "--prompt", "hello world the of and", "--steps", "12", "--max-seq-len", "128", "--device", "cpu", "--seed", "7"])
    pieces = lambda s: [ln.rsplit("|", 1)[1] for ln in s.splitlines() if "Pred" in ln and "|" in ln]
    assert pieces(out) and pieces(out) == pieces(out2)    # same seed: same text twice


def test_missing_model_is_a_clean_error(tok_path):
    rc, out = _run(["dllama", "inference", "--model", "/nonexistent.m", "--tokenizer", tok_path, "--prompt", "x", "--steps", "2"])
    assert rc != 0 and "error" in out.lower()


def test_chat_repl_on_cpu_engine(tiny_model_path, tok_path, monkeypatch):
    monkeypatch.setattr(sys, "stdin", io.StringIO("\nhello world\n"))      # empty system prompt, one user turn, then EOF
    rc, out = _run(["dllama", "chat", "--model", tiny_model_path, "--tokenizer", tok_path, "--buffer-float-type", "q80",
                    "--max-seq-len", "96", "--device", "cpu", "--temperature", "0"])
    assert rc == 0, out
    assert "Assistant" in out or "🤖" in out


@pytest.mark.gpu
def test_inference_on_gpu_engine_matches_cpu_text(tiny_model_path, tok_path):
    argv = ["dllama", "inference", "--model", tiny_model_path, "--tokenizer", tok_path, "--buffer-float-type", "q80",
            "--prompt", "hello world the of and", "--steps", "10", "--max-seq-len", "128", "--temperature", "0"]
    rc_g, out_g = _run(argv)
    assert rc_g == 0, out_g
    _check_inference_output(out_g, 10)


@pytest.mark.parametrize("backend", ["native", "python"])
def test_dllama_api_executable_serves_http(tiny_model_path, tok_path, backend):
    """`./dllama-api --model .. --tokenizer .. --port P` as a user starts it (reference: src/dllama-api.cpp:367-410),
    on the CPU engine, with either front end; SIGTERM ends it cleanly."""
    import json
    import signal
    import socket
    import subprocess
    import time
    import urllib.request
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    p = subprocess.Popen([sys.executable, os.path.join(root, "dllama-api"), "--model", tiny_model_path, "--tokenizer", tok_path,
                          "--buffer-float-type", "q80", "--max-seq-len", "128", "--device", "cpu", "--port", str(port),
                          "--temperature", "0", "--api-backend", backend, "--max-batch", "8"],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        t0 = time.time()
        up = False
        while time.time() - t0 < 120 and p.poll() is None:
            try:
                with urllib.request.urlopen(f"http://127.0.0.1:{port}/v1/models", timeout=2) as r:
                    up = json.load(r)["data"][0]["id"] == os.path.basename(tiny_model_path)
                break
            except OSError:
                time.sleep(0.2)
        assert up, p.stdout.read() if p.poll() is not None else "server did not come up"
        body = json.dumps({"messages": [{"role": "user", "content": "hello world"}], "max_tokens": 5}).encode()
        req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/chat/completions", data=body, method="POST",
                                     headers={"Content-Type": "application/json"})
        with urllib.request.urlopen(req, timeout=120) as r:
            d = json.load(r)
        assert d["object"] == "chat.completion" and 1 <= d["usage"]["completion_tokens"] <= 5
    finally:
        p.send_signal(signal.SIGTERM)
        try:
            p.wait(timeout=30)
        except subprocess.TimeoutExpired:
            p.kill()
            raise


def test_root_plus_worker_over_localhost_tcp_on_cpu(tiny_model_path, tok_path):
    """BASELINE.json configuration 1 (plumbing, no GPU): `dllama worker --port P` + `dllama inference --workers
    127.0.0.1:P` as separate processes, q80 buffer; the worker has no model file; the text equals the single-node run."""
    import socket
    import subprocess
    import time
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    exe = os.path.join(root, "dllama")
    worker = subprocess.Popen([sys.executable, exe, "worker", "--port", str(port), "--nthreads", "2", "--device", "cpu"],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd="/")
    try:
        time.sleep(1.0)
        common = ["inference", "--model", tiny_model_path, "--tokenizer", tok_path, "--buffer-float-type", "q80",
                  "--prompt", "hello world the of and", "--steps", "12", "--max-seq-len", "128", "--device", "cpu",
                  "--temperature", "0", "--nthreads", "2"]
        r2 = subprocess.run([sys.executable, exe] + common + ["--workers", f"127.0.0.1:{port}"], capture_output=True,
                            text=True, timeout=180)
        assert r2.returncode == 0, r2.stdout + r2.stderr
        r1 = subprocess.run([sys.executable, exe] + common, capture_output=True, text=True, timeout=180)
        assert r1.returncode == 0, r1.stdout + r1.stderr
        _check_inference_output(r2.stdout, 12)
        pieces = lambda s: [ln.rsplit("|", 1)[1] for ln in s.splitlines() if "Pred" in ln and "|" in ln]
        assert pieces(r2.stdout) and pieces(r2.stdout) == pieces(r1.stdout)        # 2 nodes == 1 node, token for token
        sent = [int(m) for m in re.findall(r"Pred.*Sent\s*(\d+) kB", r2.stdout)]
        assert sent and sent[-1] > 0 and "Sent     0 kB" in r1.stdout                 # real socket counters
    finally:
        worker.terminate()
        out = worker.communicate(timeout=30)[0]
    assert "Worker is ready" in out and "Weights loaded" in out, out


def test_cpu_worker_serves_a_second_root_after_the_first_left(tiny_model_path, tok_path):
    """`dllama worker` goes back to listening when its root stops or dies (reference: runWorkerApp, src/app.cpp:405-464):
    one worker process serves a root that finishes, a root that is killed mid-run, and a third one."""
    import socket
    import subprocess
    import time
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    exe = os.path.join(root, "dllama")
    worker = subprocess.Popen([sys.executable, exe, "worker", "--port", str(port), "--nthreads", "1", "--device", "cpu"],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd="/")
    argv = [sys.executable, exe, "inference", "--model", tiny_model_path, "--tokenizer", tok_path, "--buffer-float-type", "q80",
            "--prompt", "hello world", "--max-seq-len", "128", "--device", "cpu", "--temperature", "0", "--nthreads", "1",
            "--workers", f"127.0.0.1:{port}"]
    try:
        time.sleep(1.0)
        r1 = subprocess.run(argv + ["--steps", "8"], capture_output=True, text=True, timeout=180)
        assert r1.returncode == 0, r1.stdout + r1.stderr
        time.sleep(0.5)
        victim = subprocess.Popen(argv + ["--steps", "120"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        t0 = time.time()
        while time.time() - t0 < 60:                       # wait until it is decoding, then kill it without warning
            line = victim.stdout.readline()
            if "Pred" in line or not line:
                break
        victim.kill()
        victim.wait(timeout=30)
        time.sleep(1.0)
        r3 = subprocess.run(argv + ["--steps", "8"], capture_output=True, text=True, timeout=180)
        assert r3.returncode == 0, r3.stdout + r3.stderr
        pieces = lambda s: [ln.rsplit("|", 1)[1] for ln in s.splitlines() if "Pred" in ln and "|" in ln]
        assert pieces(r1.stdout) and pieces(r1.stdout) == pieces(r3.stdout)
    finally:
        worker.terminate()
        out = worker.communicate(timeout=30)[0]
    assert out.count("Worker is ready") == 3 and out.count("waiting for a new one") >= 2, out


def test_every_environment_variable_is_documented():
    """utils/envvars.py lists every DLLAMA_* variable the sources read (Python, C++, scripts)."""
    import os
    import re
    from dllama_b200.utils import ENV_VARS
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    seen = set()
    for base in ("distributed-llama-multiusers_b200", "scripts", "baseline"):
        for dp, _, files in os.walk(os.path.join(root, base)):
            if "_ref" in dp:
                continue
            for f in files:
                if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h", ".hpp")):
                    seen |= set(re.findall(r"DLLAMA_[A-Z0-9_]+", open(os.path.join(dp, f), errors="ignore").read()))
    seen |= set(re.findall(r"DLLAMA_[A-Z0-9_]+", open(os.path.join(root, "bench.py")).read()))
    missing = sorted(v for v in seen if v not in ENV_VARS)
    assert not missing, f"undocumented environment variables: {missing}"
