#!/usr/bin/env python3
"""Per-tile timeline of CTA 0 of the tcgen05 GEMM (debug hook dl_debug_set_tc_trace): which hand-off
of the producer -> dequant -> MMA pipeline sets the per-k-tile period?  Times are SM clocks (clock64)
relative to the first stamp; columns are the stamp slots of gemm_q40_tc.cu (TC_DBG)."""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from dllama_b200.ops import kernels as K  # noqa: E402

NAMES = ["A issue", "fullA seen", "emptyH seen", "math done", "arrived", "fullH seen", "fullB seen", "committed"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--d", type=int, default=28672)
    ap.add_argument("--n", type=int, default=4096)
    ap.add_argument("--M", type=int, default=32)
    ap.add_argument("--S", type=int, default=1)
    ap.add_argument("--tiles", type=int, default=32)
    ap.add_argument("--mode", type=int, default=0, help="debug timing experiment (kernel dbg_mode)")
    ap.add_argument("--brief", type=int, default=0)
    a = ap.parse_args()
    L = K.cuda_lib()
    L.dl_debug_set_tc_trace.argtypes = [C.c_void_p]
    L.dl_debug_set_tc_mode(a.mode)
    d, n, M = a.d, a.n, a.M
    wbytes = K.q40t_bytes(d, n)
    ws = [torch.randint(0, 255, (wbytes,), dtype=torch.uint8, device="cuda") for _ in range(4)]
    for w in ws:
        v = w.view(-1, 9216)[:, 8192:].view(torch.float16)
        v.copy_(torch.full_like(v, 0.01))
    ncap = 16 if M <= 16 else 32 if M <= 32 else 64 if M <= 64 else (M + 127) // 128 * 128
    xh = K.alloc_xh(n, ncap, "cuda")
    K.quant_q80(torch.randn(M, n, device="cuda"), M, n, ncap, None, xh)
    y = torch.zeros(max(M, 8), d, device="cuda")
    epi = K.epi_store([y.data_ptr()], d)
    for i in range(3):
        K.gemm_q40_tc(ws[i], d, n, xh, ncap, M, epi, a.S)
    torch.cuda.synchronize()
    trace = torch.zeros(64 * 8, dtype=torch.int64, device="cuda")
    L.dl_debug_set_tc_trace(trace.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    K.gemm_q40_tc(ws[3], d, n, xh, ncap, M, epi, a.S)
    e1.record()
    torch.cuda.synchronize()
    L.dl_debug_set_tc_trace(None)
    t = trace.cpu().numpy().reshape(64, 8)
    nk = min(a.tiles, n // 128 // a.S)
    t0 = t[:nk][t[:nk] > 0].min()
    print(f"mode={a.mode} d={d} n={n} M={M} S={a.S}: event time {e0.elapsed_time(e1) * 1e3:.1f} us, {wbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9:.0f} GB/s")
    print("tile " + " ".join(f"{x:>12s}" for x in NAMES))
    for i in range(nk if not a.brief else 0):
        print(f"{i:4d} " + " ".join(f"{(int(v) - int(t0)) if v else -1:12d}" for v in t[i]))
    k = t[63]
    print(f"CTA 0 clocks from kernel entry: setup done {k[1]-k[0]}, first tile stamp {t0-k[0]}, accumulators complete {k[2]-k[0]}, exit {k[3]-k[0]}; epilogue chunk 0: tmem->smem done {k[4]-k[0]}, barrier {k[5]-k[0]}, rows done {k[6]-k[0]}, barrier {k[7]-k[0]}")
    c = t[:nk, 7].astype(np.float64)
    per = np.diff(c[c > 0])
    if per.size:
        print(f"commit-to-commit period: median {np.median(per):.0f} clk, mean {per.mean():.0f} clk")


if __name__ == "__main__":
    main()
