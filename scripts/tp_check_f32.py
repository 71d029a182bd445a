#!/usr/bin/env python3
"""Tensor-parallel check of the native f32 path under torchrun (N GPUs of one node): every rank runs F32Engine on its
weight slice (partial sums travel as f32 peer stores + release flags), rank 0 compares the logits with the CPU oracle
(f32 sync, like the reference's f32 buffer mode) and all ranks must pick the same greedy token.  Exit code 0 = pass."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    from dllama_b200.formats import mfile
    from dllama_b200.models.f32_engine import F32Engine
    from dllama_b200.models.oracle import OracleLlama
    from dllama_b200.ops import quant
    from dllama_b200.parallel.symm import init_distributed
    rank, world, _ = init_distributed()
    path = os.path.join(tempfile.gettempdir(), "dllama_tpcheck_f32_tiny.m")
    if rank == 0 and not os.path.exists(path):
        mfile.write_random_model(path, "tiny", weight_type=quant.F32, seed=3)
    if world > 1:
        dist.barrier()
    e = F32Engine(path, rank=rank, world=world, max_rows=16, n_slots=3, max_seq_len=256)
    assert e.native, "the f32 engine must run its CUDA kernels on a GPU box"
    ok = True
    s = e.alloc_slot()
    prompt = [1, 5, 9, 200, 33, 17, 4]
    got = e.forward(prompt, list(range(len(prompt))), [s] * len(prompt)).float().cpu().numpy().copy()
    toks = []
    tok, pos = int(np.argmax(got[-1])), len(prompt)
    for _ in range(4):
        tok = e.decode_step_greedy([tok], [pos], [s])[0]
        toks.append(tok)
        pos += 1
    if rank == 0:
        o = OracleLlama(mfile.ModelFile(path), sync_q80=False, max_seq_len=256)
        ref = o.forward(prompt, 0).numpy()
        err = float(np.abs(got - ref).max() / np.abs(ref).max())
        print(f"[rank 0] prompt logits rel err {err:.2e}", flush=True)
        ok = ok and err < 1e-4
        t, p, want = int(np.argmax(ref[-1])), len(prompt), []
        for _ in range(4):
            t = int(np.argmax(o.forward([t], p).numpy()[0]))
            want.append(t)
            p += 1
        print(f"[rank 0] greedy tokens {toks} oracle {want}", flush=True)
        ok = ok and toks == want
    allt = [None] * world
    if world > 1:
        dist.all_gather_object(allt, toks)
        ok = ok and all(t == allt[0] for t in allt)
    e.close()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        dist.destroy_process_group()
    if rank == 0:
        print("TP_CHECK_F32", "PASS" if int(flag) else "FAIL", f"world={world}", flush=True)
    sys.exit(0 if int(flag) else 1)


if __name__ == "__main__":
    main()
