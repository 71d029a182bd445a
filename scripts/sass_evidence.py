#!/usr/bin/env python3
"""Count the SASS mnemonics that prove which Blackwell features each kernel uses (CPU only: reads the
built libdllama_b200_cuda.so with cuobjdump).  Output -> profiles/sass_evidence.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed-llama-multiusers_b200", "lib", "libdllama_b200_cuda.so")
KEYS = ["UTCHMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UBLKCP", "SYNCS", "ACQBULK", "IDP.4A", "UCGABAR", "MAPA", "ELECT",
        "NANOSLEEP", "ST.E.STRONG.SYS", "LD.E.STRONG.SYS", "MEMBAR", "CCTL", "HMMA", "HFMA2", "HMUL2", "HADD2", "LDS.128", "STS.128",
        "REDUX", "SHFL", "ATOMG", "RED."]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            for k in KEYS:
                if op.startswith(k):
                    kernels[cur][k] += 1
                    break
    lines = ["SASS mnemonic counts per kernel (cuobjdump -sass libdllama_b200_cuda.so), sm_100a; regenerate with scripts/sass_evidence.py",
             "UTCHMMA = tcgen05.mma kind::f16 (operand `tmem[..]` in the A slot = A from tensor memory), STTM = tcgen05.st, LDTM = tcgen05.ld,",
             "UTCATOMSWS = tcgen05.alloc/dealloc, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (TMA 1-D), SYNCS = mbarrier ops,",
             "IDP.4A = dp4a, UCGABAR = cluster barrier, MAPA = DSMEM address map, ELECT = elect.sync, ST/LD.E.STRONG.SYS = system-scope flags", ""]
    for name, c in kernels.items():
        lines.append(name)
        lines.append("    " + ", ".join(f"{k}: {v}" for k, v in c.items() if v))
    # one UTCHMMA line verbatim, to show the TMEM A operand
    m = re.search(r"UTCHMMA[^;]*;", out)
    if m:
        lines += ["", "first UTCHMMA of gemm_q40_tc_kernel, verbatim:", "    " + m.group(0)]
    path = os.path.join(ROOT, "profiles", "sass_evidence.txt")
    open(path, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[-14:]))


if __name__ == "__main__":
    sys.exit(main())
