#!/usr/bin/env python3
"""Decode-step profiling target for ncu (1 GPU): loads a model, warms up, then runs a few eager
decode steps between cudaProfilerStart/Stop.

  python scripts/profile_decode.py --layers 4 --steps 3                      (plain run, must exit 0)
  ncu --profile-from-start off --metrics gpu__time_duration.sum --csv ...    (launch list)
  ncu --profile-from-start off --set full -k regex:gemv -c 3 ...            (one kernel in depth)
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.1-8b")
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--users", type=int, default=1)
    ap.add_argument("--ctx", type=int, default=128)
    ap.add_argument("--pdl", type=int, default=0)
    ap.add_argument("--prefill", type=int, default=0, help="profile one prefill forward of this many rows instead of decode steps")
    ap.add_argument("--model-dir", default="/dev/shm")
    a = ap.parse_args()
    from dllama_b200.formats import mfile
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.ops import quant
    path = os.path.join(a.model_dir, f"dllama_b200_{a.model}-l{a.layers}_q40.m")
    if not os.path.exists(path):
        mfile.write_random_model(path, a.model, quant.Q40, seed=1, n_layers=a.layers)
    U = a.users
    eng = LlamaEngine(path, max_rows=max(8, U, a.prefill), n_slots=U + 2, max_seq_len=a.ctx + 64, use_graphs=False,
                      use_pdl=bool(a.pdl))
    slots = [eng.alloc_slot() for _ in range(U)]
    toks = [5 + u for u in range(U)]
    # fill the KV cache up to ctx by running real forwards at increasing positions (cheap: U rows)
    for s in slots:
        eng.ensure_capacity(s, a.ctx + a.steps + 8)
    pos = a.ctx
    eng.set_batch(toks, [pos] * U, slots)
    for _ in range(2):
        eng.decode_greedy_device(1, U)
    torch.cuda.synchronize()
    if a.prefill:
        ps = eng.alloc_slot()
        prompt = [7 + (i % 1000) for i in range(a.prefill)]
        eng.prefill(ps, prompt, 0)                   # warm-up (same shapes)
        eng.free_slot(ps)
        ps = eng.alloc_slot()
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        eng.prefill(ps, prompt, 0)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        print("profile target done")
        eng.close()
        return
    torch.cuda.profiler.start()
    eng.decode_greedy_device(a.steps, U)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("profile target done")
    eng.close()


if __name__ == "__main__":
    main()
