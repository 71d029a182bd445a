#!/usr/bin/env python3
"""Where a long prompt's time goes: one forward of --rows prompt tokens as a single batch with the per-launch
timeline on (first-CTA start / last-CTA end of every kernel, globaltimer ns, no profiler), summed by kernel kind.
Usage: python scripts/trace_prefill.py [--rows 4096] [--layers 32] [--wh auto|off]"""
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
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--rows", type=int, default=4096)
    ap.add_argument("--wh", default="auto")
    ap.add_argument("--pt-dbg", type=int, default=0, help="attention_prefill_tc timing experiments (results invalid): 1 no softmax, 2 no KV copies, 4 no MMAs")
    ap.add_argument("--model-dir", default="/dev/shm")
    a = ap.parse_args()
    from dllama_b200.formats import mfile
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.ops import kernels as K, quant
    full = a.layers == mfile.MODEL_SHAPES[a.model]["n_layers"]
    path = os.path.join(a.model_dir, f"dllama_b200_{a.model}_q40.m" if full else f"dllama_b200_{a.model}-l{a.layers}_q40.m")
    if not os.path.exists(path):
        mfile.write_random_model(path, a.model, quant.Q40, seed=1, **({} if full else {"n_layers": a.layers}))
    n = a.rows
    if a.pt_dbg:
        K.cuda_lib().dl_debug_set_pt_mode(a.pt_dbg)
    eng = LlamaEngine(path, max_rows=n, n_slots=2, max_seq_len=n + 64, use_graphs=False, use_pdl=False, prefill_wh=a.wh)
    prompt = np.random.default_rng(1).integers(1, eng.h.vocab_size, n).tolist()
    slot = eng.alloc_slot()
    eng.decode_step_greedy(prompt, list(range(n)), [slot] * n, logits_from=n - 1)       # warm-up
    torch.cuda.synchronize()
    eng.free_slot(slot)
    slot = eng.alloc_slot()
    L = K.cuda_lib()
    L.dl_trace_set.argtypes = [C.c_void_p, C.c_int]
    cap = 4096
    buf = torch.zeros(cap, 2, dtype=torch.int64, device="cuda")
    buf[:, 0] = torch.iinfo(torch.int64).max
    torch.cuda.synchronize()
    L.dl_trace_set(buf.data_ptr(), cap)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    eng.decode_step_greedy(prompt, list(range(n)), [slot] * n, logits_from=n - 1)
    ev1.record()
    torch.cuda.synchronize()
    cnt = L.dl_trace_count()
    kinds = [L.dl_trace_kind(i) for i in range(cnt)]
    L.dl_trace_set(None, 0)
    t = buf.cpu().numpy()[:cnt].astype(np.float64)
    print(f"rows={n} layers={a.layers} weights={eng.wh_mode}: forward {ev0.elapsed_time(ev1):.2f} ms (eager launches, no PDL), {cnt} traced kernels")
    agg = {}
    for i in range(cnt):
        k = KINDS[kinds[i]]
        agg.setdefault(k, [0, 0.0])
        agg[k][0] += 1
        agg[k][1] += t[i, 1] - t[i, 0]
    for k, (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"  {k:6s} n={c:4d}  busy {d / 1e6:8.2f} ms")
    # per-launch detail of the second layer (kernel order inside a layer is fixed)
    per = (cnt - 3) // a.layers if a.layers else 0
    if per > 0:
        print(f"  layer 1, launch by launch ({per} kernels):")
        for i in range(1 + per, 1 + 2 * per):
            print(f"    {KINDS[kinds[i]]:6s} {(t[i, 1] - t[i, 0]) / 1e3:9.1f} us")
    eng.close()


if __name__ == "__main__":
    main()
