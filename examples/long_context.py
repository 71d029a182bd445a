#!/usr/bin/env python3
"""Long-context regression (the role of the reference's examples/macbeth.sh: fill the KV cache with a
long greedy continuation and compare exactly).  Runs `--steps` greedy tokens from a fixed prompt and
prints a digest of the generated ids; the digest must be identical on every rank of a TP run and
stable from run to run (the kernels are deterministic: fixed reduction orders, no float atomics).

    python examples/long_context.py --model m.m --steps 2048
    torchrun --nproc-per-node 4 examples/long_context.py --model m.m --steps 2048
    python examples/long_context.py --model m.m --steps 2048 --device cpu --nodes 4 --nthreads 2
        (no GPU: the host backend, root + 3 workers over loopback TCP started by this script; the digest does not
         depend on --nthreads, and a second run must print the same digest)
"""
import argparse
import hashlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", required=True)
    ap.add_argument("--steps", type=int, default=2048)
    ap.add_argument("--expect", default="")
    ap.add_argument("--device", default="cuda", choices=["cuda", "cpu"])
    ap.add_argument("--nodes", type=int, default=1, help="--device cpu: root + (nodes-1) workers over loopback TCP")
    ap.add_argument("--nthreads", type=int, default=0)
    a = ap.parse_args()
    if a.device == "cpu":
        return main_cpu(a)
    import torch
    import torch.distributed as dist
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.parallel.symm import init_distributed
    rank, world, _ = init_distributed()
    eng = LlamaEngine(a.model, rank=rank, world=world, max_rows=8, n_slots=1, max_seq_len=a.steps + 64)
    slot = eng.alloc_slot()
    prompt = [1 + (i * 7919) % (eng.h.vocab_size - 1) for i in range(16)]
    lg = eng.prefill(slot, prompt, 0)
    tok = int(torch.argmax(lg[0]))
    eng.ensure_capacity(slot, len(prompt) + a.steps + 1)
    eng.set_batch([tok], [len(prompt)], [slot])
    out = []
    done = 0
    while done < a.steps:                       # the history ring holds 4096 steps
        n = min(1024, a.steps - done)
        eng.decode_greedy_device(n, 1)
        torch.cuda.synchronize()
        out += eng.read_history(n, 1, eng.current_step() - n)[:, 0].tolist()
        done += n
    digest = hashlib.sha256(",".join(map(str, out)).encode()).hexdigest()[:16]
    all_d = [None] * world
    if world > 1:
        dist.all_gather_object(all_d, digest)
    else:
        all_d = [digest]
    if rank == 0:
        ok = all(d == digest for d in all_d) and (not a.expect or a.expect == digest)
        print(f"tokens={len(out)} world={world} digest={digest} {'✅' if ok else '❌ ' + str(all_d)}")
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def generate_cpu(model, steps, nodes, nthreads):
    """Greedy continuation on the host backend; returns the generated token ids."""
    import socket
    import threading
    import time
    import torch
    from dllama_b200.models import native_cpu_engine as N
    ports, threads = [], []
    for _ in range(nodes - 1):
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            ports.append(s.getsockname()[1])
    for p in ports:
        t = threading.Thread(target=N.serve_worker, args=(p, nthreads), daemon=True)
        t.start()
        threads.append(t)
    if ports:
        time.sleep(0.3)
    eng = N.NativeCpuEngine(model, workers=[("127.0.0.1", p) for p in ports], n_threads=nthreads, max_rows=16, n_slots=1,
                            max_seq_len=steps + 64)
    prompt = [1 + (i * 7919) % (eng.h.vocab_size - 1) for i in range(16)]
    lg = eng.prefill(0, prompt, 0)
    tok = int(torch.argmax(lg[0]))
    out, pos = [], len(prompt)
    for _ in range(steps):
        out.append(tok)
        tok = eng.decode_step_greedy([tok], [pos], [0])[0]
        pos += 1
    eng.close()
    for t in threads:
        t.join(timeout=30)
    return out


def main_cpu(a):
    out = generate_cpu(a.model, a.steps, a.nodes, a.nthreads)
    digest = hashlib.sha256(",".join(map(str, out)).encode()).hexdigest()[:16]
    ok = not a.expect or a.expect == digest
    print(f"tokens={len(out)} nodes={a.nodes} digest={digest} {'✅' if ok else '❌'}")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main() or 0)
