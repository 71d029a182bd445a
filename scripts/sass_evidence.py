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
KEYS = ["UTCHMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UBLKCP", "UTMALDG", "SYNCS", "ACQBULK", "IDP.4A", "UCGABAR", "MAPA", "ELECT",
        "NANOSLEEP", "ST.E.STRONG.SYS", "LD.E.STRONG.SYS", "MEMBAR", "CCTL", "HMMA", "HFMA2", "HMUL2", "HADD2", "LDS.128", "STS.128",
        "REDUX", "SHFL", "ATOMG", "RED."]


def write_listings(out: str):
    """One SASS listing per kernel under profiles/sass/ (instruction text only: the 128-bit encodings are dropped to keep
    the files reviewable).  Template families keep every instantiation."""
    ddir = os.path.join(ROOT, "profiles", "sass")
    os.makedirs(ddir, exist_ok=True)
    for f in os.listdir(ddir):
        if f.endswith(".sass"):
            os.remove(os.path.join(ddir, f))
    demangle = {}
    try:
        names = sorted(set(re.findall(r"Function : (\S+)", out)))
        dm = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
        demangle = dict(zip(names, dm))
    except Exception:
        pass
    cur, buf, index = None, [], []
    # template families: one or two representative instantiations are enough for review (the others differ in unroll counts)
    keep_variants = {"attention_kernel": ("int)128, (int)1, __half", "int)64, (int)1, float"), "decode_mega_kernel": ("(int)1>",),
                     "gemv_q40_kernel": ("(int)1>", "(int)4>"), "gemm_f32_kernel": ("(int)8>",), "attention_prefill_kernel": ("(int)128>",)}

    def flush():
        if cur is None:
            return
        nice = demangle.get(cur, cur)
        for fam, pats in keep_variants.items():
            if fam + "<" in nice and not any(p in nice for p in pats):
                return
        base = nice.replace("void ", "")
        base = base[:base.rfind("(")] if "(" in base else base                   # drop the parameter list, keep template arguments
        fname = re.sub(r"[^A-Za-z0-9_]+", "_", base).strip("_")[:80] + ".sass"
        with open(os.path.join(ddir, fname), "w") as f:
            f.write(f"// {nice}\n// cuobjdump -sass libdllama_b200_cuda.so (sm_100a), encodings stripped\n" + "\n".join(buf) + "\n")
        index.append((fname, nice, len(buf)))

    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            flush()
            cur, buf = m.group(1), []
            continue
        if cur is None:
            continue
        if re.match(r"\s+/\* 0x[0-9a-f]{16} \*/\s*$", line):          # second half of an encoding
            continue
        line = re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", line)
        if line.strip():
            buf.append(line.rstrip())
    flush()
    with open(os.path.join(ddir, "INDEX.md"), "w") as f:
        f.write("# SASS listings (sm_100a), one file per kernel -- regenerate with `python scripts/sass_evidence.py`\n\n")
        for fname, nice, n in sorted(index):
            f.write(f"* `{fname}` ({n} lines): `{nice}`\n")
    return index


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    idx = write_listings(out)
    print(f"wrote {len(idx)} listings to profiles/sass/")
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
             "UTCATOMSWS = tcgen05.alloc/dealloc, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (TMA 1-D), UTMALDG = cp.async.bulk.tensor (TMA tensor copy), SYNCS = mbarrier ops,",
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
