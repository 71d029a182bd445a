#!/usr/bin/env python3
"""Headline benchmark: Llama-3.1-8B Q40 single-stream decode tok/s, TP = --gpus (strong scaling).

    python bench.py --gpus N --steps K --warmup W          (N>1: launched under torchrun)
    python bench.py --impl reference ...                   (the unmodified reference, CPU + TCP)

Metric/config are the ones BASELINE.json names.  Weights: random-init `.m` of the named
architecture (no network); prompt: synthetic token ids.  `value` is timed on the device with CUDA
events (max over ranks) around K decode steps replayed from a CUDA graph in a device-resident
greedy loop; `e2e.value` is the same K steps through the public API with a pinned-host -> device
descriptor copy and a device -> host token read every step.  Every step streams this rank's full
Q40 weight image (4.5 GB at TP=1, 0.53 GB at TP=8) -- far larger than the 126 MB L2 -- so no
explicit L2 flush is needed between iterations.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BASELINE_PUBLISHED = None       # BASELINE.json "published": {} -> vs_baseline is null
METRIC = "Llama-3.1-8B Q40 single-stream decode tok/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "nccl_cublas"])
    ap.add_argument("--model", default="llama3.1-8b")
    ap.add_argument("--layers", type=int, default=0, help="debug only: override layer count (invalid as a bench)")
    ap.add_argument("--prompt-len", type=int, default=128)
    ap.add_argument("--prefill-chunk", type=int, default=256, help="rows per prefill forward (engine max_rows)")
    ap.add_argument("--b2-cache", type=int, default=1, help="nccl_cublas arm: keep bf16 copies of the weights (1) or dequantise per call (0)")
    ap.add_argument("--b2-graph", type=int, default=1, help="nccl_cublas arm: replay the decode step from a CUDA graph with the argmax on the device (1) or run eagerly (0)")
    ap.add_argument("--no-pdl", action="store_true")
    ap.add_argument("--no-graphs", action="store_true")
    ap.add_argument("--users", type=int, default=32, help="also measure an N-user batched decode (key multi_user; 0 = skip)")
    ap.add_argument("--long-prompt", type=int, default=4096, help="also measure the TTFT of one long prompt run as a single batch on fp16-expanded weights (key ttft_long; 0 = skip)")
    ap.add_argument("--model-dir", default=os.environ.get("DLLAMA_MODEL_DIR", "/dev/shm"))
    ap.add_argument("--nthreads", type=int, default=0, help="reference arm: threads per node (0 = auto)")
    return ap.parse_args()


# --------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.samples = []
        self._stop = threading.Event()
        self._t = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.idx)], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 8:
                    self.samples.append(parts)
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        self._t = threading.Thread(target=self._loop, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm = sorted(float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit())
        mx = [float(s[2]) for s in self.samples if s[2].replace(".", "").isdigit()]
        reasons = set()
        for s in self.samples:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.samples)}


def ensure_model(args, rank: int) -> str:
    from dllama_b200.formats import mfile
    from dllama_b200.ops import quant
    name = args.model + (f"-l{args.layers}" if args.layers else "")
    path = os.path.join(args.model_dir, f"dllama_b200_{name}_q40.m")
    if rank == 0 and not os.path.exists(path):
        t0 = time.time()
        mfile.write_random_model(path, args.model, quant.Q40, seed=1, n_layers=args.layers or None)
        sys.stderr.write(f"[bench] wrote {path} ({os.path.getsize(path) / 1e9:.2f} GB) in {time.time() - t0:.1f}s\n")
    return path


def run_b200(args):
    import torch
    import torch.distributed as dist
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.ops import lib
    from dllama_b200.parallel.symm import init_distributed

    rank, world, local_rank = init_distributed()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    path = ensure_model(args, rank)
    if world > 1:
        dist.barrier()
    K_, W = args.steps, args.warmup
    ctx = args.prompt_len + 2 * (K_ + W) + 64
    eng = LlamaEngine(path, rank=rank, world=world, max_rows=max(args.prefill_chunk, args.users), n_slots=max(2, args.users + 2),
                      max_seq_len=max(512, ctx), use_pdl=not args.no_pdl, use_graphs=not args.no_graphs)
    h = eng.h
    rng = __import__("numpy").random.default_rng(0)
    prompt = rng.integers(1, h.vocab_size, args.prompt_len).tolist()
    slot = eng.alloc_slot()
    logits = eng.prefill(slot, prompt, 0)
    tok = int(torch.argmax(logits[0]))
    torch.cuda.synchronize()
    pos = len(prompt)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-timed: K graph replays of the decode step ----------------
    eng.ensure_capacity(slot, pos + W + K_ + 1)
    eng.set_batch([tok], [pos], [slot])
    sampler = ClockSampler(torch.cuda.current_device())
    sampler.start()
    eng.decode_greedy_device(W, 1)
    barrier()
    n0 = lib.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    eng.decode_greedy_device(K_, 1)
    ev1.record()
    barrier()
    launches = lib.launch_count - n0
    dev_ms = ev0.elapsed_time(ev1)
    # (the first use of a batch shape runs one untimed, uncaptured forward before the graph: count back from the end)
    tokens_dev = eng.read_history(W + K_, 1, eng.current_step() - (W + K_))[:, 0].tolist()

    # ---------------- end to end through the public API, on a second sequence with the same prompt ----------------
    # every step: pinned-host descriptor -> device, forward, device argmax, token id -> host.  Starting from the same
    # prompt the host loop must reproduce the tokens of the device-resident loop exactly (same kernels, greedy).
    slot2 = eng.alloc_slot()
    logits2 = eng.prefill(slot2, prompt, 0)
    tok_host = int(torch.argmax(logits2[0]))
    pos2 = len(prompt)
    eng.ensure_capacity(slot2, pos2 + W + K_ + 1)
    tokens_host = []
    for _ in range(W):
        tok_host = eng.decode_greedy_host([tok_host], [pos2], [slot2])[0]
        tokens_host.append(tok_host)
        pos2 += 1
    barrier()
    t0 = time.perf_counter()
    for _ in range(K_):
        tok_host = eng.decode_greedy_host([tok_host], [pos2], [slot2])[0]
        tokens_host.append(tok_host)
        pos2 += 1
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()
    tokens_match = tokens_host == tokens_dev
    if not tokens_match:
        bad = next(i for i, (x, y) in enumerate(zip(tokens_host, tokens_dev)) if x != y)
        raise SystemExit(f"[bench] device-resident loop and host loop disagree at step {bad}: {tokens_dev[bad]} vs {tokens_host[bad]}")
    eng.free_slot(slot2)
    pos += W + K_

    t = torch.tensor([dev_ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = float(t[0]), float(t[1])
    ttft = bench_ttft(eng, args, barrier, world, dist, prompt)
    multi = bench_users(eng, args, barrier, world, dist) if args.users > 1 else None
    clocks = sampler.stop()
    ms_per_step = dev_ms / K_
    value = 1000.0 / ms_per_step
    wbytes = eng.weight_bytes()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0)
    result = {
        "metric": METRIC, "value": round(value, 2), "unit": "tokens/s", "n_gpus": world, "steps": K_, "warmup": W,
        "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": BASELINE_PUBLISHED, "dtype": "q40 weights x q80 activations (int8 dp4a, f32 accumulate)",
        "data": "synthetic (random-init .m of the named architecture, random prompt token ids)",
        "config": {"model": args.model + (f"-{args.layers}layers-DEBUG" if args.layers else ""),
                   "global_batch": 1, "seq_len": args.prompt_len + W + K_, "parallelism": f"tp{world}",
                   "kv_cache": "fp16 paged (32-token pages)", "sync": "q80 partial sums over NVLink peer stores",
                   "l2": "inputs larger than L2: every step streams %.0f MB of weights per GPU" % (wbytes / 1e6),
                   "pdl": not args.no_pdl, "cuda_graph": not args.no_graphs},
        "clocks": clocks,
        "e2e": {"value": round(1000.0 * K_ / e2e_ms, 2), "unit": "tokens/s", "h2d_bytes_per_step": eng._h_desc3.numel() * 4,
                "d2h_bytes_per_step": eng.max_rows * 4, "ms_per_step": round(e2e_ms / K_, 4)},
        "gpu_launches": launches,
        "roofline": {"weight_bytes_per_gpu_per_step": wbytes, "hbm_gbs_measured": hbm,
                     "hbm_floor_ms": round(wbytes / (hbm * 1e9) * 1e3, 4),
                     "achieved_fraction_of_measured_hbm": round(wbytes / (hbm * 1e9) * 1e3 / ms_per_step, 4)},
        "tokens_match_host_loop": tokens_match,
        "persistent_kernel": bool(eng.use_mega),
        "ttft": ttft,
        "impl": "b200",
    }
    if multi is not None:
        result["multi_user"] = multi
    eng.close()
    del eng
    torch.cuda.empty_cache()
    if args.long_prompt > 0:
        result["ttft_long"] = bench_ttft_long(args, path, rank, world, barrier, dist)
    if rank == 0:
        print(json.dumps(result), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def bench_ttft(eng, args, barrier, world, dist, prompt):
    """p50 time-to-first-token of a warm engine: H2D of the prompt descriptor, prefill in chunks of
    --prefill-chunk rows through the tcgen05 GEMM, device argmax, D2H of the token.  Host clock around
    a synchronised region (it is a latency as a client sees it), max over ranks."""
    import torch
    times = []
    for rep in range(6):
        slot = eng.alloc_slot()
        barrier()
        t0 = time.perf_counter()
        out = None
        for i in range(0, len(prompt), args.prefill_chunk):
            part = prompt[i:i + args.prefill_chunk]
            last = i + args.prefill_chunk >= len(prompt)
            if last:
                out = eng.decode_step_greedy(part, list(range(i, i + len(part))), [slot] * len(part), logits_from=len(part) - 1)
            else:
                eng.forward(part, list(range(i, i + len(part))), [slot] * len(part), logits_from=len(part) - 1)
        ms = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rep > 0:                                  # rep 0 captures the graphs
            times.append(float(t[0]))
        eng.free_slot(slot)
        assert out is not None
    times.sort()
    p50 = times[len(times) // 2]
    return {"prompt_tokens": len(prompt), "chunk_rows": args.prefill_chunk, "p50_ms": round(p50, 3), "min_ms": round(times[0], 3),
            "prefill_tokens_per_s": round(len(prompt) * 1000.0 / p50, 1)}


def bench_ttft_long(args, path, rank, world, barrier, dist):
    """TTFT of one long prompt evaluated as ONE batch (an engine with max_rows = prompt length): the matmuls run as the
    two-phase GEMM on resident fp16-expanded weights (gemm_wh_tc.cu), attention on the tensor-core prefill kernel.
    Same clock as bench_ttft: host timer around H2D of the descriptor ... D2H of the first token, max over ranks."""
    import numpy as np
    import torch
    from dllama_b200.models.llama import LlamaEngine
    n = args.long_prompt
    eng = LlamaEngine(path, rank=rank, world=world, max_rows=n, n_slots=2, max_seq_len=n + 64, use_pdl=not args.no_pdl,
                      use_graphs=not args.no_graphs)
    prompt = np.random.default_rng(1).integers(1, eng.h.vocab_size, n).tolist()
    times = []
    for rep in range(4):
        slot = eng.alloc_slot()
        barrier()
        t0 = time.perf_counter()
        out = eng.decode_step_greedy(prompt, list(range(n)), [slot] * n, logits_from=n - 1)
        ms = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rep > 0:
            times.append(float(t[0]))
        eng.free_slot(slot)
        assert out is not None
    times.sort()
    p50 = times[len(times) // 2]
    flops = 2.0 * n * (eng.weight_bytes() - eng.wcls.numel()) / 9216 * 16384       # 128 x 128 weights per 9216-byte tile
    res = {"prompt_tokens": n, "chunk_rows": n, "p50_ms": round(p50, 3), "min_ms": round(times[0], 3),
           "prefill_tokens_per_s": round(n * 1000.0 / p50, 1), "weights": eng.wh_mode, "fp16_weight_copy_bytes": eng.wh_bytes(),
           "matmul_tflops_per_gpu_at_p50": round(flops / p50 / 1e9, 1)}
    eng.close()
    return res


def bench_users(eng, args, barrier, world, dist):
    """Aggregate tok/s of U concurrent users decoding in one batch (continuous-batching steady state)."""
    import torch
    U = args.users
    rng = __import__("numpy").random.default_rng(1)
    slots, toks, poss = [], [], []
    for u in range(U):
        s = eng.alloc_slot()
        p = rng.integers(1, eng.h.vocab_size, 32 + u).tolist()
        lg = eng.prefill(s, p, 0)
        slots.append(s)
        toks.append(int(torch.argmax(lg[0])))
        poss.append(len(p))
    K_, W = args.steps, args.warmup
    for s, p in zip(slots, poss):
        eng.ensure_capacity(s, p + K_ + W + 1)
    eng.set_batch(toks, poss, slots)
    eng.decode_greedy_device(W, U)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    eng.decode_greedy_device(K_, U)
    ev1.record()
    barrier()
    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0]) / K_
    return {"users": U, "ms_per_step": round(ms, 4), "aggregate_tokens_per_s": round(U * 1000.0 / ms, 1)}


def run_nccl_cublas(args):
    """B2: NCCL all-reduce + Q40 dequant + cuBLAS (models/baseline_nccl.py), same metric and config."""
    import torch
    import torch.distributed as dist
    from dllama_b200.models.baseline_nccl import NcclCublasEngine
    from dllama_b200.parallel.symm import init_distributed
    rank, world, _ = init_distributed()
    path = ensure_model(args, rank)
    if world > 1:
        dist.barrier()
    K_, W = args.steps, args.warmup
    eng = NcclCublasEngine(path, rank=rank, world=world, max_seq_len=max(512, args.prompt_len + K_ + W + 64),
                           cache_weights=bool(args.b2_cache))
    rng = __import__("numpy").random.default_rng(0)
    prompt = rng.integers(1, eng.h.vocab_size, args.prompt_len).tolist()
    lg = eng.forward(prompt, 0)
    tok, pos = int(lg[-1].argmax()), len(prompt)
    use_graph = bool(args.b2_graph)
    if use_graph:
        # one CUDA graph per step, greedy pick on the device: the honest form of "NCCL + cuBLAS"
        eng.start_graph_decode(tok, pos)
        for _ in range(W):
            eng.graph_step()
    else:
        for _ in range(W):
            tok = int(eng.forward([tok], pos)[0].argmax())
            pos += 1
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(K_):
        if use_graph:
            eng.graph_step()
        else:
            tok = int(eng.forward([tok], pos)[0].argmax())
            pos += 1
    ev1.record()
    torch.cuda.synchronize()
    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0]) / K_
    if rank == 0:
        print(json.dumps({"metric": METRIC, "value": round(1000.0 / ms, 2), "unit": "tokens/s", "n_gpus": world,
                          "steps": K_, "warmup": W, "ms_per_step": round(ms, 4), "higher_is_better": True,
                          "scaling": "strong", "vs_baseline": None, "dtype": "bf16 cuBLAS on dequantised Q40 (%s), NCCL all-reduce, %s" % ("bf16 copies cached in HBM" if args.b2_cache else "dequantised on every call", "one CUDA graph per step, device argmax" if use_graph else "eager, host argmax"),
                          "data": "synthetic", "impl": "nccl_cublas",
                          "config": {"model": args.model, "global_batch": 1, "parallelism": f"tp{world}"}}), flush=True)
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_reference(args):
    from baseline.reference_arm import run_reference_arm
    print(json.dumps(run_reference_arm(args, METRIC)), flush=True)


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries (NCCL's version banner, torchrun notices) also
    write to fd 1, so fd 1 is pointed at stderr for the whole run and the JSON line goes to the saved fd."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(saved, "w", buffering=1)


if __name__ == "__main__":
    a = parse_args()
    _claim_stdout()
    if a.impl == "reference":
        try:
            run_reference(a)
        except Exception as e:      # never fail the driver: report why instead
            if int(os.environ.get("RANK", "0")) == 0:
                print(json.dumps({"impl": "reference", "unavailable": f"{type(e).__name__}: {e}"[:300]}), flush=True)
        sys.exit(0)
    if a.impl == "nccl_cublas":
        run_nccl_cublas(a)
    else:
        run_b200(a)
