#!/usr/bin/env python3
"""BASELINE.json configuration 1 on a machine without a GPU: Llama-3.2-1B-shaped Q40 random-init `.m`, root +
(N-1) workers over localhost TCP, q80 buffer, CPU.  Runs this repo's native host backend (`dllama inference` +
`dllama worker`, separate processes) and the unmodified reference (its own build, same files, same prompt, same
thread count per node) and prints one JSON line per arm: single-stream decode tok/s from each program's own
per-token timers.

    python scripts/cpu_tcp_bench.py [--model llama3.2-1b] [--nodes 2] [--nthreads 2] [--steps 24]
"""
import argparse
import json
import os
import re
import socket
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run_ours(mpath, tpath, prompt, steps, nodes, nthreads):
    exe = os.path.join(ROOT, "dllama")
    ports, workers = [], []
    for _ in range(nodes - 1):
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            ports.append(s.getsockname()[1])
    try:
        for p in ports:
            workers.append(subprocess.Popen([sys.executable, exe, "worker", "--port", str(p), "--nthreads", str(nthreads),
                                             "--device", "cpu"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL))
        time.sleep(1.5)
        cmd = [sys.executable, exe, "inference", "--model", mpath, "--tokenizer", tpath, "--buffer-float-type", "q80",
               "--prompt", prompt, "--steps", str(steps), "--nthreads", str(nthreads), "--max-seq-len", "2048",
               "--temperature", "0", "--seed", "1", "--device", "cpu"]
        if ports:
            cmd += ["--workers"] + [f"127.0.0.1:{p}" for p in ports]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=1800)
        if r.returncode != 0:
            raise RuntimeError(r.stdout[-400:] + r.stderr[-400:])
    finally:
        for w in workers:
            w.terminate()
    m = re.search(r"Prediction\s+nTokens:\s*(\d+)\s+tokens/s:\s*([\d.]+)\s*\(([\d.]+) ms/tok\)", r.stdout)
    e = re.search(r"Evaluation\s+nBatches:\s*\d+\s+nTokens:\s*(\d+)\s+tokens/s:\s*([\d.]+)", r.stdout)
    sync = [int(x) for x in re.findall(r"Pred\s*\d+ ms Sync\s*(\d+) ms", r.stdout)]
    sent = [int(x) for x in re.findall(r"Pred.*Sent\s*(\d+) kB", r.stdout)]
    return {"pred_tokens_per_s": float(m.group(2)), "pred_ms_per_token": float(m.group(3)), "pred_tokens": int(m.group(1)),
            "eval_tokens_per_s": float(e.group(2)) if e else None,
            "sync_ms_per_token": round(sum(sync) / max(1, len(sync)), 2),
            "sent_kb_per_token": round((sent[-1] - sent[0]) / max(1, len(sent) - 1), 1) if len(sent) > 1 else 0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.2-1b")
    ap.add_argument("--model-dir", default="/dev/shm")
    ap.add_argument("--nodes", type=int, default=2)
    ap.add_argument("--nthreads", type=int, default=2)
    ap.add_argument("--steps", type=int, default=24)
    ap.add_argument("--prompt-tokens", type=int, default=16)
    ap.add_argument("--skip-reference", action="store_true")
    ap.add_argument("--repeats", type=int, default=3, help="runs per arm; the fastest is reported (shared, noisy hosts)")
    a = ap.parse_args()
    from baseline import reference_arm as RA
    from dllama_b200.formats import mfile, tfile
    from dllama_b200.ops import quant
    mpath = os.path.join(a.model_dir, f"dllama_b200_{a.model}_q40.m")
    tpath = os.path.join(a.model_dir, f"dllama_b200_{a.model}.t")
    if not os.path.exists(mpath):
        mfile.write_random_model(mpath, a.model, quant.Q40, seed=1)
    if not os.path.exists(tpath):
        tfile.write_synthetic_tokenizer(tpath, mfile.MODEL_SHAPES[a.model]["vocab_size"])
    prompt, n_prompt = RA._prompt_and_count(tpath, a.prompt_tokens)
    steps = n_prompt - 1 + a.steps
    base = {"config": f"{a.model} Q40, root + {a.nodes - 1} worker(s) over localhost TCP, q80 buffer, CPU",
            "nodes": a.nodes, "nthreads_per_node": a.nthreads, "host_cpus": os.cpu_count(), "prompt_tokens": n_prompt,
            "model_bytes": os.path.getsize(mpath)}
    runs = [run_ours(mpath, tpath, prompt, steps, a.nodes, a.nthreads) for _ in range(a.repeats)]
    ours = min(runs, key=lambda r: r["pred_ms_per_token"])
    print(json.dumps({"impl": "b200-dllama native CPU backend", **base, **ours,
                      "all_runs_ms_per_token": [r["pred_ms_per_token"] for r in runs]}), flush=True)
    if not a.skip_reference:
        binary = RA.build_reference()
        runs = []
        for _ in range(a.repeats):
            r = RA._run_once(binary, mpath, tpath, prompt, n_prompt, a.nodes, a.nthreads, 0, a.steps)
            runs.append(r + (RA.LAST_RUN.get("eval_tokens_per_s"),))
        ms, tps, _, ev = min(runs, key=lambda r: r[0])
        print(json.dumps({"impl": "reference (unmodified, own build)", **base, "pred_tokens_per_s": tps,
                          "pred_ms_per_token": round(ms, 2), "eval_tokens_per_s": ev,
                          "all_runs_ms_per_token": [round(r[0], 2) for r in runs]}), flush=True)


if __name__ == "__main__":
    main()
