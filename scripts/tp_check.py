#!/usr/bin/env python3
"""Tensor-parallel correctness check, run under torchrun on N GPUs of one node:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29533 scripts/tp_check.py [--model tiny-tp8]

Every rank runs the engine on its weight slice; partial sums travel through peer memory (epilogue
stores + release flags).  Rank 0 compares logits with the CPU oracle run with n_nodes=N (which
Q80-casts each node's partial sum like the reference) and all ranks must agree on greedy tokens.
Exit code 0 = pass.
"""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="tiny-tp8")
    ap.add_argument("--no-pdl", action="store_true")
    a = ap.parse_args()
    from dllama_b200.formats import mfile
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.models.oracle import OracleLlama
    from dllama_b200.parallel.symm import init_distributed
    rank, world, _ = init_distributed()
    path = os.path.join(tempfile.gettempdir(), f"dllama_tpcheck_{a.model}.m")
    if rank == 0 and not os.path.exists(path):
        mfile.write_random_model(path, a.model, exact=True, seed=5)
    if world > 1:
        dist.barrier()
    ok = True

    def check(name, cond, detail=""):
        nonlocal ok
        ok = ok and bool(cond)
        if rank == 0 or not cond:
            print(f"[rank {rank}] {'PASS' if cond else 'FAIL'} {name} {detail}", flush=True)

    for graphs in (False, True):
        eng = LlamaEngine(path, rank=rank, world=world, max_rows=64, n_slots=4, max_seq_len=256,
                          use_graphs=graphs, use_pdl=not a.no_pdl)
        oracle = OracleLlama(mfile.ModelFile(path), n_nodes=world, max_seq_len=256) if rank == 0 else None
        rng = np.random.default_rng(1)
        vocab = eng.h.vocab_size
        # --- GEMV path: prefill 6 rows then teacher-forced decode ---
        slot = eng.alloc_slot()
        prompt = rng.integers(1, vocab, 6).tolist()
        got = eng.prefill(slot, prompt, 0).float().cpu().numpy()[0]
        if rank == 0:
            ref = oracle.forward(prompt, 0).numpy()
            err = float(np.abs(got - ref[-1]).max() / np.abs(ref[-1]).max())
            check(f"prefill logits (graphs={graphs})", err < 3e-2, f"rel err {err:.2e}")
        tok = int(np.argmax(got))
        toks = []
        pos = len(prompt)
        for i in range(6):
            lg = eng.forward([tok], [pos], [slot]).float().cpu().numpy()[0]
            if rank == 0:
                ref = oracle.forward([tok], pos).numpy()[0]
                err = float(np.abs(lg - ref).max() / np.abs(ref).max())
                check(f"decode step {i} logits (graphs={graphs})", err < 3e-2, f"rel err {err:.2e}")
            tok = int(np.argmax(lg))
            toks.append(tok)
            pos += 1
        # --- tensor-core path: 20-row chunk on a second sequence ---
        slot2 = eng.alloc_slot()
        p2 = rng.integers(1, vocab, 20).tolist()
        lg2 = eng.forward(p2, list(range(20)), [slot2] * 20, logits_from=0).float().cpu().numpy()
        if rank == 0:
            oracle.reset()
            ref2 = oracle.forward(p2, 0).numpy()
            err = float(np.abs(lg2 - ref2).max() / np.abs(ref2).max())
            check(f"tcgen05 20-row chunk logits (graphs={graphs})", err < 4e-2, f"rel err {err:.2e}")
        # --- 50-row chunk: more rows than the LL exchange takes, so the partial sums travel as Q80 wire rows under the
        #     flag protocol at every world size; logits of the last 3 rows only (offset into the exchange rows) ---
        slot3 = eng.alloc_slot()
        p3 = rng.integers(1, vocab, 50).tolist()
        lg3 = eng.forward(p3, list(range(50)), [slot3] * 50, logits_from=47).float().cpu().numpy()
        if rank == 0:
            oracle.reset()
            ref3 = oracle.forward(p3, 0).numpy()[47:]
            err = float(np.abs(lg3 - ref3).max() / np.abs(ref3).max())
            check(f"50-row chunk, Q80 wire exchange, last-3 logits (graphs={graphs})", err < 4e-2, f"rel err {err:.2e}")
        eng.free_slot(slot3)
        # --- device-resident greedy loop: all ranks must produce identical tokens ---
        t0 = int(np.argmax(lg2[-1]))
        eng.ensure_capacity(slot2, 20 + 12)
        eng.set_batch([t0], [20], [slot2])
        torch.cuda.synchronize()
        eng.decode_greedy_device(8, 1)
        torch.cuda.synchronize()
        hist = eng.read_history(8, 1, eng.current_step() - 8)[:, 0]
        all_hist = [None] * world
        if world > 1:
            dist.all_gather_object(all_hist, hist.tolist())
        else:
            all_hist = [hist.tolist()]
        check(f"device-loop tokens identical on all ranks (graphs={graphs})",
              all(h == all_hist[0] for h in all_hist), str(all_hist[0]))
        all_t = [None] * world
        if world > 1:
            dist.all_gather_object(all_t, toks)
            check("host-loop tokens identical on all ranks", all(t == all_t[0] for t in all_t))
        eng.close()
        if world > 1:
            dist.barrier()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        dist.destroy_process_group()
    if rank == 0:
        print("TP_CHECK", "PASS" if int(flag) else "FAIL", f"world={world}", flush=True)
    sys.exit(0 if int(flag) else 1)


if __name__ == "__main__":
    main()
