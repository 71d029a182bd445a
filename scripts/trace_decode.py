#!/usr/bin/env python3
"""Timeline of one decode step (every kernel's first-CTA start and last-CTA end, globaltimer ns),
taken WITHOUT a profiler so PDL overlap and peer waits are the real ones.  Works under torchrun
(rank 0 prints).  Usage: python scripts/trace_decode.py [--layers 4] [--pdl 1] [--users 1]"""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

KINDS = ["embed", "rnq", "gemv", "att", "argmax", "tc", "other"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.1-8b")
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--users", type=int, default=1)
    ap.add_argument("--ctx", type=int, default=128)
    ap.add_argument("--pdl", type=int, default=1)
    ap.add_argument("--model-dir", default="/dev/shm")
    ap.add_argument("--max-print", type=int, default=60)
    a = ap.parse_args()
    from dllama_b200.formats import mfile
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.ops import kernels as K, quant
    from dllama_b200.parallel.symm import init_distributed
    import torch.distributed as dist
    rank, world, _ = init_distributed()
    path = os.path.join(a.model_dir, f"dllama_b200_{a.model}-l{a.layers}_q40.m")
    if rank == 0 and not os.path.exists(path):
        mfile.write_random_model(path, a.model, quant.Q40, seed=1, n_layers=a.layers)
    if world > 1:
        dist.barrier()
    U = a.users
    eng = LlamaEngine(path, rank=rank, world=world, max_rows=max(8, U), n_slots=U + 1, max_seq_len=a.ctx + 64,
                      use_graphs=False, use_pdl=bool(a.pdl))
    slots = [eng.alloc_slot() for _ in range(U)]
    for s in slots:
        eng.ensure_capacity(s, a.ctx + 16)
    eng.set_batch([5 + u for u in range(U)], [a.ctx] * U, slots)
    eng.decode_greedy_device(3, U)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    L = K.cuda_lib()
    L.dl_trace_set.argtypes = [C.c_void_p, C.c_int]
    cap = 4096
    buf = torch.zeros(cap, 2, dtype=torch.int64, device="cuda")
    buf[:, 0] = torch.iinfo(torch.int64).max
    torch.cuda.synchronize()
    L.dl_trace_set(buf.data_ptr(), cap)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    eng.decode_greedy_device(2, U)
    ev1.record()
    torch.cuda.synchronize()
    n = L.dl_trace_count()
    kinds = [L.dl_trace_kind(i) for i in range(n)]
    L.dl_trace_set(None, 0)
    t = buf.cpu().numpy()[:n].astype(np.float64)
    if rank == 0:
        half = n // 2
        t0 = t[half, 0]
        print(f"world={world} layers={a.layers} users={U} pdl={a.pdl}: 2 steps in {ev0.elapsed_time(ev1) * 1e3:.1f} us; "
              f"second step = {(t[n - 1, 1] - t[half, 0]) / 1e3:.1f} us over {n - half} kernels")
        prev_end = None
        agg = {}
        for i in range(half, n):
            s, e = t[i, 0] - t0, t[i, 1] - t0
            gap = (s - prev_end) if prev_end is not None else 0.0
            k = KINDS[kinds[i]]
            agg.setdefault(k, [0, 0.0, 0.0])
            agg[k][0] += 1
            agg[k][1] += (e - s)
            agg[k][2] += max(gap, 0.0)
            if i - half < a.max_print:
                print(f"  {i - half:3d} {k:6s} start {s / 1e3:8.2f}  dur {(e - s) / 1e3:6.2f}  gap-from-prev-end {gap / 1e3:6.2f} us")
            prev_end = e
        for k, (c, d, g) in agg.items():
            print(f"  sum {k:6s} n={c:3d} busy {d / 1e3:8.1f} us   positive gaps before {g / 1e3:7.1f} us")
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
