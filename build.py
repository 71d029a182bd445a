#!/usr/bin/env python3
"""Build the native libraries in-tree (no pip, no JIT cache):

  distributed-llama-multiusers_b200/lib/libdllama_b200_cuda.so   nvcc, sm_100a only
  distributed-llama-multiusers_b200/lib/libdllama_b200_host.so   g++, CPU runtime (tokenizer,
                                                                 sampler, codecs, scheduler ...)

Usage: python build.py [--force] [--verbose]
Objects are rebuilt only when a source or header is newer (so repeated calls are cheap).
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "distributed-llama-multiusers_b200")
CUDA_SRC = os.path.join(PKG, "csrc", "cuda")
HOST_SRC = os.path.join(PKG, "csrc", "host")
LIB = os.path.join(PKG, "lib")
OBJ = os.path.join(ROOT, "build", "obj")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v"]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-Wall", "-pthread"]
# Debug / sanitizer builds (reference: `make DEBUG=1` -> -g -fsanitize=address, Makefile:8-9): extra flags from the
# environment are appended, e.g.  CXXFLAGS="-g -O1 -fsanitize=address" python build.py --force   (run python with
# LD_PRELOAD=$(g++ -print-file-name=libasan.so)), NVCCFLAGS="-G" for device debug info.
CXX_FLAGS += os.environ.get("CXXFLAGS", "").split()
NVCC_FLAGS += os.environ.get("NVCCFLAGS", "").split()


def _find_nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd, verbose, log_path=None):
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log_path:
        with open(log_path, "w") as f:
            f.write(" ".join(cmd) + "\n" + p.stdout)
    if p.returncode != 0:
        sys.stderr.write(p.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd))
    if verbose:
        sys.stdout.write(p.stdout)


def build_cuda(force=False, verbose=False) -> str:
    nvcc = _find_nvcc()
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIB, exist_ok=True)
    headers = [os.path.join(CUDA_SRC, f) for f in os.listdir(CUDA_SRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "dllama_b200.h"))
    srcs = sorted(f for f in os.listdir(CUDA_SRC) if f.endswith(".cu"))
    objs, jobs = [], []
    for s in srcs:
        src = os.path.join(CUDA_SRC, s)
        obj = os.path.join(OBJ, s[:-3] + ".o")
        objs.append(obj)
        if force or _newer(obj, [src] + headers):
            jobs.append(([nvcc] + NVCC_FLAGS + ["-c", src, "-o", obj], obj[:-2] + ".ptxas.log"))
    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        list(ex.map(lambda j: _run(j[0], verbose, j[1]), jobs))
    out = os.path.join(LIB, "libdllama_b200_cuda.so")
    if force or jobs or _newer(out, objs):
        _run([nvcc, "-shared", "-o", out] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static"], verbose)
    return out


def build_host(force=False, verbose=False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIB, exist_ok=True)
    if not os.path.isdir(HOST_SRC):
        return ""
    headers = [os.path.join(HOST_SRC, f) for f in os.listdir(HOST_SRC) if f.endswith((".hpp", ".h"))]
    srcs = sorted(f for f in os.listdir(HOST_SRC) if f.endswith(".cpp"))
    if not srcs:
        return ""
    objs, jobs = [], []
    for s in srcs:
        src = os.path.join(HOST_SRC, s)
        obj = os.path.join(OBJ, "host_" + s[:-4] + ".o")
        objs.append(obj)
        if force or _newer(obj, [src] + headers):
            jobs.append((["g++"] + CXX_FLAGS + ["-c", src, "-o", obj], None))
    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        list(ex.map(lambda j: _run(j[0], verbose, j[1]), jobs))
    out = os.path.join(LIB, "libdllama_b200_host.so")
    if force or jobs or _newer(out, objs):
        _run(["g++", "-shared", "-o", out] + objs + ["-pthread"] + [f for f in CXX_FLAGS if f.startswith("-fsanitize")], verbose)
    return out


def build_all(force=False, verbose=False):
    return build_cuda(force, verbose), build_host(force, verbose)


if __name__ == "__main__":
    a, b = build_all("--force" in sys.argv, "--verbose" in sys.argv)
    print("built:", a, b)
