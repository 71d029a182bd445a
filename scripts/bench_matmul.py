#!/usr/bin/env python3
"""Micro-benchmark of the two Q40 matmul kernels on the Llama-3.1-8B layer shapes (1 GPU).

Each shape gets enough distinct weight copies (>= 512 MB total) that consecutive launches never hit
L2; times are CUDA-event averages after warm-up.  Reports achieved HBM bandwidth against
MEASURED_PEAKS.json and, for the tensor-core kernel, achieved TFLOP/s.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from dllama_b200.ops import kernels as K  # noqa: E402
from dllama_b200.ops.lib import EPI_STORE  # noqa: E402


def timed(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3      # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--iters", type=int, default=24)
    ap.add_argument("--gemv-m", default="1,2,4,8")
    ap.add_argument("--tc-m", default="8,16,32,64,128,256")
    ap.add_argument("--wh-m", default="", help="rows for the two-phase prompt GEMM (gemm_wh_tc), e.g. 256,512,1024,2048,4096")
    ap.add_argument("--wh-cg", type=int, default=0, help="1 = force single CTAs (no cta_group::2 pairs)")
    ap.add_argument("--wh-dbg", type=int, default=0, help="timing experiments of gemm_wh_tc (results invalid): 1 no epilogue, 2 no MMAs, 4 no activation copies")
    ap.add_argument("--wh-pf", default="", help="gemm_wh_tc L2 prefetch: distance,mask (k-steps ahead, activation-tile mask)")
    ap.add_argument("--wh-piece", type=int, default=0, help="gemm_wh_tc bytes per bulk copy (4096, 8192, 16384)")
    ap.add_argument("--out", default="")
    ap.add_argument("--shapes", default="qkv,wo,w13,w2,wcls")
    a = ap.parse_args()
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    hbm = peaks["hbm_gbs"]
    dim, hid, N = 4096, 14336, a.tp
    shapes = {"qkv": (6144 // N, dim), "wo": (dim, dim // N), "w13": (2 * hid // N, dim), "w2": (dim, hid // N),
              "wcls": (128256 // N, dim)}
    n_sms = K.cuda_lib().dl_device_sm_count()
    if a.wh_dbg:
        K.cuda_lib().dl_debug_set_wh_mode(a.wh_dbg)
    if a.wh_piece:
        K.cuda_lib().dl_debug_set_wh_piece(a.wh_piece)
    if a.wh_pf:
        K.cuda_lib().dl_debug_set_wh_prefetch(*[int(x) for x in a.wh_pf.split(",")])
    rows = []
    for name, (d, n) in shapes.items():
        if name not in a.shapes.split(","):
            continue
        wbytes = K.q40t_bytes(d, n)
        copies = max(2, min(64, (512 << 20) // wbytes + 1))
        ws = [torch.randint(0, 255, (wbytes,), dtype=torch.uint8, device="cuda") for _ in range(copies)]
        for w in ws:      # sane fp16 scales (avoid inf/nan patterns)
            v = w.view(-1, 9216)[:, 8192:].view(torch.float16)
            v.copy_(torch.full_like(v, 0.01))
        max_m = max([256] + [int(x) for x in (a.tc_m + "," + a.wh_m).split(",") if x])
        y = torch.zeros(max_m, d, device="cuda")
        epi = K.epi_store([y.data_ptr()], d)
        for M in [int(x) for x in a.gemv_m.split(",") if x]:
            xq = K.alloc_xq(n, M, "cuda")
            x = torch.randn(M, n, device="cuda")
            K.quant_q80(x, M, n, M, xq, None)
            R, S, T = K.choose_splits(d, n, EPI_STORE, n_sms, M)
            it = [0]

            def run():
                K.gemv_q40(ws[it[0] % copies], d, n, xq, M, epi, R, S, tiles_per_cta=T)
                it[0] += 1
            us = timed(run, a.iters)
            rows.append(dict(kernel="gemv", shape=name, d=d, n=n, M=M, R=R, S=S, T=T, us=round(us, 2),
                             gbs=round(wbytes / us / 1e3, 1), frac_hbm=round(wbytes / us / 1e3 / hbm, 3)))
            print(rows[-1], flush=True)
        for M in [int(x) for x in a.tc_m.split(",") if x]:
            ncap = 16 if M <= 16 else 32 if M <= 32 else 64 if M <= 64 else (M + 127) // 128 * 128
            xh = K.alloc_xh(n, ncap, "cuda")
            x = torch.randn(M, n, device="cuda")
            K.quant_q80(x, M, n, ncap, None, xh)
            S = K.choose_tc_split(d, n, M, n_sms)
            it = [0]

            def run_tc():
                K.gemm_q40_tc(ws[it[0] % copies], d, n, xh, ncap, M, epi, S)
                it[0] += 1
            us = timed(run_tc, a.iters)
            rows.append(dict(kernel="tc", shape=name, d=d, n=n, M=M, S=S, us=round(us, 2),
                             gbs=round(wbytes / us / 1e3, 1), frac_hbm=round(wbytes / us / 1e3 / hbm, 3),
                             tflops=round(2.0 * d * n * M / us / 1e6, 1)))
            print(rows[-1], flush=True)
        wh_ms = [int(x) for x in a.wh_m.split(",") if x]
        if wh_ms:
            whs = [K.q40t_to_wh(ws[i], d, n) for i in range(2)]
            it = [0]

            def run_deq():
                K.q40t_to_wh(ws[it[0] % copies], d, n, whs[it[0] % 2])
                it[0] += 1
            us = timed(run_deq, a.iters)
            rows.append(dict(kernel="q40t_to_wh", shape=name, d=d, n=n, us=round(us, 2),
                             gbs=round((wbytes + K.wh_bytes(d, n)) / us / 1e3, 1)))
            print(rows[-1], flush=True)
            for M in wh_ms:
                ncap = (M + 255) // 256 * 256
                xh = K.alloc_xh(n, ncap, "cuda")
                x = torch.randn(M, n, device="cuda")
                K.quant_q80(x, M, n, ncap, None, xh)
                it = [0]

                def run_wh():
                    K.gemm_wh_tc(whs[it[0] % 2], d, n, xh, ncap, M, epi, a.wh_cg)
                    it[0] += 1
                us = timed(run_wh, a.iters)
                rows.append(dict(kernel="wh_tc", shape=name, d=d, n=n, M=M, cg=1 if (a.wh_cg == 1 or ((d + 127) // 128) % 2) else 2,
                                 us=round(us, 2), tflops=round(2.0 * d * n * M / us / 1e6, 1)))
                print(rows[-1], flush=True)
            del whs
        del ws
        torch.cuda.empty_cache()
    if a.out:
        json.dump(rows, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
