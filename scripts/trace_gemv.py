#!/usr/bin/env python3
"""Per-CTA timeline of the GEMV kernel (debug hook dl_debug_set_gemv_trace): where do the
microseconds of a small launch go?  Prints min/median/max of each trace point relative to the
first CTA's start."""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from dllama_b200.ops import kernels as K  # noqa: E402
from dllama_b200.ops.lib import EPI_STORE  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--d", type=int, default=6144)
    ap.add_argument("--n", type=int, default=4096)
    ap.add_argument("--M", type=int, default=1)
    ap.add_argument("--configs", default="auto,1x1,1x2,1x4,1x8,2x1,4x1")
    ap.add_argument("--fused", type=int, default=0)
    a = ap.parse_args()
    L = K.cuda_lib()
    L.dl_debug_set_gemv_trace.argtypes = [C.c_void_p]
    d, n, M = a.d, a.n, a.M
    wbytes = K.q40t_bytes(d, n)
    copies = max(2, (400 << 20) // wbytes)
    ws = [torch.randint(0, 255, (wbytes,), dtype=torch.uint8, device="cuda") for _ in range(copies)]
    for w in ws:
        v = w.view(-1, 9216)[:, 8192:].view(torch.float16)
        v.copy_(torch.full_like(v, 0.01))
    xq = K.alloc_xq(n, M, "cuda")
    K.quant_q80(torch.randn(M, n, device="cuda"), M, n, M, xq, None)
    y = torch.zeros(M, d, device="cuda")
    epi = K.epi_store([y.data_ptr()], d)
    trace = torch.zeros(4096 * 8, dtype=torch.int64, device="cuda")
    nm = None
    if a.fused:
        xin = torch.randn(8, n, device="cuda")
        wn = torch.ones(n, device="cuda")
        nm = K.norm_prolog(xin, None, None, 0, 0, None, None, 0, 0, wn, 1e-5)
    xq_arg = None if a.fused else xq
    for cfg in a.configs.split(","):
        if cfg == "auto":
            R, S, T = K.choose_splits(d, n, EPI_STORE, L.dl_device_sm_count(), M)
        else:
            R, S = [int(x) for x in cfg.split("x")][:2]
            T = int(cfg.split("x")[2]) if cfg.count("x") == 2 else 1
        nct = ((((d + 127) // 128) * R + T - 1) // T) * S
        for i in range(3):
            K.gemv_q40(ws[i % copies], d, n, xq_arg, M, epi, R, S, norm=nm, tiles_per_cta=T)
        torch.cuda.synchronize()
        trace.zero_()
        L.dl_debug_set_gemv_trace(trace.data_ptr())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        K.gemv_q40(ws[3 % copies], d, n, xq_arg, M, epi, R, S, norm=nm, tiles_per_cta=T)
        e1.record()
        torch.cuda.synchronize()
        L.dl_debug_set_gemv_trace(None)
        t = trace.cpu().numpy().reshape(-1, 8)[:nct].astype(np.float64)
        t0 = t[:, 0].min()
        print(f"--- d={d} n={n} M={M} R={R} S={S} T={T} ctas={nct}  event time {e0.elapsed_time(e1) * 1e3:.1f} us "
              f"({wbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9:.0f} GB/s)")
        names = ["start", "act ready", "loop done", "local reduce done", "ssq exchanged", "epilogue done", "prologue done"]
        for k in (0, 4, 6, 1, 2, 3, 5):
            col = t[:, k]
            col = col[col > 0] - t0
            if col.size:
                print(f"   {names[k]:18s} min {col.min() / 1e3:7.2f}  med {np.median(col) / 1e3:7.2f}  max {col.max() / 1e3:7.2f} us  (n={col.size})")


if __name__ == "__main__":
    main()
