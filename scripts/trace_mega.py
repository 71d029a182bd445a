#!/usr/bin/env python3
"""Per-phase timeline of the persistent decode kernel (CTA 0, globaltimer): python scripts/trace_mega.py [--model llama3.1-8b]
Under torchrun it prints rank 0's timeline of a tensor-parallel step."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.1-8b")
    ap.add_argument("--model-dir", default=os.environ.get("DLLAMA_MODEL_DIR", "/dev/shm"))
    ap.add_argument("--ctx", type=int, default=128)
    ap.add_argument("--rows", type=int, default=1)
    a = ap.parse_args()
    import torch.distributed as dist
    from dllama_b200.formats import mfile
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.ops import quant
    from dllama_b200.parallel.symm import init_distributed
    rank, world, _ = init_distributed()
    path = os.path.join(a.model_dir, f"dllama_b200_{a.model}_q40.m")
    if rank == 0 and not os.path.exists(path):
        mfile.write_random_model(path, a.model, quant.Q40, seed=1)
    if world > 1:
        dist.barrier()
    eng = LlamaEngine(path, rank=rank, world=world, max_rows=256, n_slots=a.rows + 1, max_seq_len=max(512, a.ctx + 64))
    rng = np.random.default_rng(0)
    slots, toks = [], []
    for u in range(a.rows):
        s = eng.alloc_slot()
        lg = eng.prefill(s, rng.integers(1, eng.h.vocab_size, a.ctx).tolist(), 0)
        slots.append(s)
        toks.append(int(torch.argmax(lg[0])))
    eng.enable_mega_trace(True)
    for s in slots:
        eng.ensure_capacity(s, a.ctx + 16)
    eng.set_batch(toks, [a.ctx] * a.rows, slots)
    eng.decode_greedy_device(6, a.rows)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t = eng.mega_trace_ns()
    L = eng.h.n_layers
    if rank == 0:
        names = ["norm+qkv prologue", "qkv gemv", "grid barrier + attention", "wo prologue+gemv(+barrier)", "ffn norm prologue",
                 "w13 gemv + barrier", "w2 prologue+gemv(+barrier)"]
        per = np.zeros((L, 7))
        for l in range(L):
            b = t[l * 8: l * 8 + 8]
            nxt = t[(l + 1) * 8] if l + 1 < L else t[L * 8]
            # stamps: 0 start, 1 after prologue, 2 after qkv+barrier, 3 after attention(+merge), 4 after wo, 5 after ffn prologue, 6 after w13+barrier, 7 after w2
            per[l] = [b[1] - b[0], b[2] - b[1], b[3] - b[2], b[4] - b[3], b[5] - b[4], b[6] - b[5], b[7] - b[6]]
        print(f"[trace_mega] world={world} model={a.model} rows={a.rows} ctx={a.ctx}: us per phase (median over layers 1..)")
        for i, n in enumerate(names):
            print(f"  {n:34s} {np.median(per[1:, i]) / 1e3:7.2f}  (min {per[1:, i].min() / 1e3:.2f}, max {per[1:, i].max() / 1e3:.2f})")
        print(f"  layer total (median)               {np.median(per[1:].sum(1)) / 1e3:7.2f}")
        print(f"  logits phase                       {(t[L * 8] - t[(L - 1) * 8 + 7]) / 1e3:7.2f}")
        print(f"  whole kernel                       {(t[L * 8 + 1] - t[0]) / 1e3:7.2f}")
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
