#!/usr/bin/env python3
"""Per-KV-step timeline of the tcgen05 prompt attention kernel (CTA (0,0) = the tile with the longest history):
clock64 stamps of the MMA warp, the softmax threads and the KV producers.  Usage: python scripts/trace_attention_tc.py [--rows 4096]"""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from dllama_b200.ops import kernels as K  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=4096)
    ap.add_argument("--heads", type=int, default=32)
    ap.add_argument("--kv-heads", type=int, default=8)
    ap.add_argument("--dbg", type=int, default=0)
    a = ap.parse_args()
    M, H, KV, hs = a.rows, a.heads, a.kv_heads, 128
    max_pages = M // 32 + 2
    q = torch.randn(M, H * hs, device="cuda")
    Kc = torch.randn(max_pages, KV, 32, hs, device="cuda").half()
    Vc = torch.randn(max_pages, KV, 32, hs, device="cuda").half()
    pos = torch.arange(M, dtype=torch.int32, device="cuda")
    slot = torch.zeros(M, dtype=torch.int32, device="cuda")
    bt = torch.arange(max_pages, dtype=torch.int32, device="cuda").reshape(1, max_pages)
    meta = torch.tensor([1, 0, M] + [0] * 16, dtype=torch.int32, device="cuda")
    ncap = (M + 127) // 128 * 128
    xh = K.alloc_xh(H * hs, ncap, "cuda")
    L = K.cuda_lib()
    L.dl_debug_set_pt_trace.argtypes = [C.c_void_p]
    if a.dbg:
        L.dl_debug_set_pt_mode(a.dbg)
    run = lambda tc: K.attention_prefill(q, M, H, KV, hs, Kc, Vc, pos, slot, bt, max_pages, meta, None, xh, ncap, None, tc=tc)
    for tc in (False, True):
        for _ in range(2):
            run(tc)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            run(tc)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 5 * 1e3
        fl = 4.0 * H * hs * M * (M + 1) / 2
        print(f"{'tcgen05' if tc else 'mma.sync'} kernel: {us:8.1f} us  ({fl / us / 1e6:6.1f} TFLOP/s causal)")
    st = torch.zeros(64, 8, dtype=torch.int64, device="cuda")
    L.dl_debug_set_pt_trace(st.data_ptr())
    run(True)
    torch.cuda.synchronize()
    L.dl_debug_set_pt_trace(None)
    s = st.cpu().numpy()
    n = int((s[:, 0] != 0).sum())
    t0 = s[0, 0]
    names = ["kv landed (mma)", "S issued", "P ready (mma)", "S ready (smx)", "max pass done", "P written", "stage free (prod)", "copies issued"]
    print("clk since the first KV tile landed; one row per KV step of CTA (0,0):")
    print("step " + " ".join(f"{x:>17s}" for x in names))
    for j in list(range(min(n, 6))) + list(range(max(6, n - 3), n)):
        print(f"{j:4d} " + " ".join(f"{int(s[j, k] - t0):17d}" for k in range(8)))
    if n > 8:
        d = np.diff(s[4:n - 1, 5])
        print(f"steady state: {d.mean():.0f} clk per KV step (P written -> P written), min {d.min()}, max {d.max()}")


if __name__ == "__main__":
    main()
