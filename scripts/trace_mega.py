#!/usr/bin/env python3
"""Per-phase timeline of the persistent decode kernel (one CTA, globaltimer):
    python scripts/trace_mega.py [--model llama3.1-8b] [--cta 0,73,146]
Under torchrun it prints rank 0's timeline of a tensor-parallel step."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

NAMES = ["norm prologue", "qkv gemv", "barrier", "attention", "barrier (+merge)", "att-merge prologue", "wo gemv",
         "barrier (1 GPU)", "ffn norm prologue", "w13 gemv", "barrier", "hidden prologue", "w2 gemv", "barrier (1 GPU)"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.1-8b")
    ap.add_argument("--model-dir", default=os.environ.get("DLLAMA_MODEL_DIR", "/dev/shm"))
    ap.add_argument("--ctx", type=int, default=128)
    ap.add_argument("--rows", type=int, default=1)
    ap.add_argument("--cta", default="0")
    ap.add_argument("--intra", action="store_true")
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
    eng = LlamaEngine(path, rank=rank, world=world, max_rows=256, n_slots=a.rows + 1, max_seq_len=max(512, a.ctx + 128))
    rng = np.random.default_rng(0)
    slots, toks = [], []
    for u in range(a.rows):
        s = eng.alloc_slot()
        lg = eng.prefill(s, rng.integers(1, eng.h.vocab_size, a.ctx).tolist(), 0)
        slots.append(s)
        toks.append(int(torch.argmax(lg[0])))
    for s in slots:
        eng.ensure_capacity(s, a.ctx + 100)
    eng.set_batch(toks, [a.ctx] * a.rows, slots)
    L = eng.h.n_layers
    for cta in [int(x) for x in a.cta.split(",")]:
        eng.enable_mega_trace(True, cta)
        eng.decode_greedy_device(6, a.rows)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t = eng.mega_trace_ns()
        if world > 1:          # one line per rank: who waits in the exchange, who is waited for
            Lr = eng.h.n_layers
            ia = np.array([[t[l * 32 + 16 + k] for k in range(4)] for l in range(1, Lr)], dtype=np.float64)
            mine = (rank, float(np.median(ia[:, 1] - ia[:, 0])) / 1.9e3, float(np.median([t[l * 32 + 1] - t[l * 32] for l in range(1, Lr)])) / 1e3,
                    float(np.median([t[l * 32 + 9] - t[l * 32 + 8] for l in range(1, Lr)])) / 1e3, float(t[Lr * 32 + 2] - t[0]) / 1e3)
            allr = [None] * world
            dist.all_gather_object(allr, mine)
            if rank == 0:
                for r_, mp, n1, n2, tot in allr:
                    print(f"  rank {r_}: ffn-norm merge pass {mp:6.2f} us | att norm prologue {n1:6.2f} | ffn norm prologue {n2:6.2f} | kernel {tot:8.1f} us")
        if rank != 0:
            continue
        per = np.zeros((L, 14))
        intra = np.zeros((L, 16))
        for l in range(L):
            b = t[l * 32: l * 32 + 15]
            per[l] = np.diff(b)
            intra[l] = t[l * 32 + 16: l * 32 + 32]
        print(f"[trace_mega] world={world} model={a.model} rows={a.rows} ctx={a.ctx} cta={cta}: us per phase (median / min / max over layers 1..)")
        for i, n in enumerate(NAMES):
            print(f"  {n:22s} {np.median(per[1:, i]) / 1e3:7.2f} {per[1:, i].min() / 1e3:7.2f} {per[1:, i].max() / 1e3:7.2f}")
        print(f"  layer total (median)   {np.median(per[1:].sum(1)) / 1e3:7.2f}")
        print(f"  logits phase           {(t[L * 32 + 1] - t[(L - 1) * 32 + 14]) / 1e3:7.2f}   final barrier {(t[L * 32 + 2] - t[L * 32 + 1]) / 1e3:.2f}")
        print(f"  whole kernel           {(t[L * 32 + 2] - t[0]) / 1e3:7.2f}")
        if a.intra:
            ghz = 1.9          # intra-phase stamps are SM clocks (sub-us resolution; the global timer ticks every 256 ns)

            def d(i, j):
                return float(np.median(intra[1:, j] - intra[1:, i])) / (ghz * 1e3)
            print("  -- inside ffn prologue_norm: merge pass (x + exchange words) %.2f | sum of squares + bar %.2f | normalise + quantise %.2f" %
                  (d(0, 1), d(1, 2), d(2, 3)))
            print("  -- inside qkv gemv: stage loop %.2f | reduce %.2f | epilogue %.2f | bar %.2f" % (d(4, 5), d(5, 6), d(6, 7), d(7, 8)))
            print("  -- inside w2 gemv:  stage loop %.2f | reduce %.2f | epilogue %.2f | bar %.2f" % (d(9, 10), d(10, 11), d(11, 12), d(12, 13)))
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
