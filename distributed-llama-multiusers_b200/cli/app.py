"""Application runtime shared by ``dllama`` and ``dllama-api``: argument parsing, engine
construction, the root/worker bootstrap, and the three CLI modes.

Surface parity with the reference (src/app.cpp:33-136 flags; src/dllama.cpp:11-239 modes):

    dllama inference --model M --tokenizer T --prompt P --steps N [options]
    dllama chat      --model M --tokenizer T [options]
    dllama worker    --port 9999 [--gpu-index i]
    dllama-api       --model M --tokenizer T [--port 9990] [options]

Options: --buffer-float-type {f32,f16,q40,q80}, --workers host:port ..., --port, --nthreads,
--steps, --temperature, --topp, --seed, --chat-template {llama2,llama3,deepSeek3}, --max-seq-len,
--gpu-index, --gpu-segments a:b, --net-turbo {0,1}; unknown options raise (src/app.cpp:131-133).
New here: --max-batch (rows per forward / concurrent users), --n-slots, --device {auto,cuda,cpu},
--api-backend {native,python} (dllama-api front end: the C++ epoll server or the Python one).

Multi-GPU: every node is one process on one GPU of the same NVSwitch box.  Two equivalent ways to
start them: (a) the reference's way -- ``dllama worker --port P`` per extra GPU, then the root with
``--workers 127.0.0.1:P ...``: the root dials each worker and hands it rank / world size / the
torch.distributed rendezvous address / the run's arguments (the analogue of NnRootConfigWriter,
src/nn/nn-network.cpp:617-683; weights are NOT shipped: each rank reads its own slice of the `.m`);
(b) ``torchrun --nproc-per-node N dllama ...``: ranks > 0 become workers automatically.
"""
from __future__ import annotations

import json
import os
import socket
import struct
import sys
import time
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

from ..ops import quant
from ..text import native as T


@dataclass
class AppArgs:
    mode: Optional[str] = None
    help: bool = False
    n_batches: int = 32
    n_threads: int = 0          # CPU backend threads per node; 0 = one per available core share
    model_path: Optional[str] = None
    tokenizer_path: Optional[str] = None
    prompt: Optional[str] = None
    sync_type: int = quant.F32
    workers: List[Tuple[str, int]] = field(default_factory=list)
    port: int = 9990
    temperature: float = 0.8
    topp: float = 0.9
    steps: int = 0
    seed: int = 0
    chat_template_type: int = T.CHAT_UNKNOWN
    max_seq_len: int = 0
    net_turbo: bool = True
    gpu_index: int = -1
    gpu_segment_from: int = -1
    gpu_segment_to: int = -1
    max_batch: int = 0          # 0 = chosen from the entry point and the device (see AppArgs.parse)
    n_slots: int = 0
    device: str = "auto"
    api_backend: str = "native"
    benchmark: bool = False

    @staticmethod
    def parse(argv: List[str], require_mode: bool) -> "AppArgs":
        a = AppArgs(seed=int(time.time()))
        i = 1
        if require_mode and len(argv) > 1:
            a.mode = argv[1]
            i = 2
        if any(x in ("--usage", "--help", "-h") for x in argv):
            a.help = True
            return a
        while i + 1 < len(argv):
            name, value = argv[i], argv[i + 1]
            if name == "--model": a.model_path = value
            elif name == "--tokenizer": a.tokenizer_path = value
            elif name == "--prompt": a.prompt = value
            elif name == "--buffer-float-type": a.sync_type = quant.parse_float_type(value)
            elif name == "--workers":
                j = i + 1
                while j < len(argv) and not argv[j].startswith("-"):
                    j += 1
                for v in argv[i + 1:j]:
                    if ":" not in v:
                        raise ValueError(f"Invalid worker address: {v}")
                    host, port = v.rsplit(":", 1)
                    a.workers.append((host, int(port)))
                i += (j - i - 1) - 1
            elif name == "--port": a.port = int(value)
            elif name == "--nthreads": a.n_threads = int(value)
            elif name == "--steps": a.steps = int(value)
            elif name == "--temperature": a.temperature = float(value)
            elif name == "--topp": a.topp = float(value)
            elif name == "--seed": a.seed = int(value)
            elif name == "--chat-template":
                if value not in T.CHAT_NAMES:
                    raise ValueError(f"Invalid chat template type: {value}")
                a.chat_template_type = T.CHAT_NAMES[value]
            elif name == "--max-seq-len": a.max_seq_len = int(value)
            elif name == "--gpu-index": a.gpu_index = int(value)
            elif name == "--gpu-segments":
                if ":" not in value:
                    raise ValueError("GPU segments expected in the format <from>:<to>")
                f, t = value.split(":", 1)
                a.gpu_segment_from, a.gpu_segment_to = int(f), int(t)
                # the reference splits a layer's ops between CPU and Vulkan segments; here every op of a rank runs on
                # its one GPU (or all on the host backend), so there is nothing to segment
                print(f"💡 --gpu-segments {value} accepted and ignored: all ops of a rank run on one device", flush=True)
            elif name == "--net-turbo": a.net_turbo = int(value) == 1
            elif name == "--max-batch": a.max_batch = int(value)
            elif name == "--n-slots": a.n_slots = int(value)
            elif name == "--device": a.device = value
            elif name == "--api-backend":
                if value not in ("native", "python"):
                    raise ValueError(f"Invalid API backend: {value}")
                a.api_backend = value
            else:
                raise ValueError(f"Unknown option: {name}")
            i += 2
        if a.max_batch <= 0:
            # rows of one forward pass.  On a GPU a large value costs memory only (activations, and from 768 rows the
            # fp16 copy of the weights), and lets simultaneous prompts share one pass: dllama-api (require_mode False)
            # 1024, the CLI 256; the host backends stay at 32
            a.max_batch = (256 if require_mode else 1024) if use_cuda_device(a.device) else 32
        return a


USAGE = """Usage:
  dllama {inference|chat|worker} [options]       dllama-api [options]
Options:
  --model <path.m> --tokenizer <path.t> --prompt <text> --steps <n>
  --buffer-float-type {f32,f16,q40,q80}   --max-seq-len <n>   --temperature <f> --topp <f> --seed <n>
  --workers <host:port> ...   --port <n>   --chat-template {llama2,llama3,deepSeek3}
  --max-batch <rows>  --n-slots <sequences>  --gpu-index <i>  --device {auto,cuda,cpu,cpu-oracle}
  --api-backend {native,python}   (dllama-api: C++ epoll front end, or the Python one)
  --nthreads <n>  threads of the CPU backend (no GPU); --gpu-segments / --net-turbo are accepted for compatibility."""


# --------------------------------------------------------------------------------------------------
# root <-> worker bootstrap (TCP, once at start-up)
# --------------------------------------------------------------------------------------------------
_ACK = 23571114


def _send_msg(sock: socket.socket, obj):
    data = json.dumps(obj).encode()
    sock.sendall(struct.pack("<II", _ACK, len(data)) + data)


def _recv_msg(sock: socket.socket):
    hdr = b""
    while len(hdr) < 8:
        chunk = sock.recv(8 - len(hdr))
        if not chunk:
            raise ConnectionError("peer closed during bootstrap")
        hdr += chunk
    magic, n = struct.unpack("<II", hdr)
    if magic != _ACK:
        raise ConnectionError("bad bootstrap magic")
    data = b""
    while len(data) < n:
        chunk = sock.recv(n - len(data))
        if not chunk:
            raise ConnectionError("peer closed during bootstrap")
        data += chunk
    return json.loads(data.decode())


def dial_workers(args: AppArgs, master_port: int) -> int:
    """Root side: hand every worker its rank and the rendezvous; returns world size."""
    world = 1 + len(args.workers)
    for i, (host, port) in enumerate(args.workers):
        with socket.create_connection((host, port), timeout=60) as s:
            _send_msg(s, {"rank": i + 1, "world": world, "master_addr": "127.0.0.1", "master_port": master_port,
                          "model": os.path.abspath(args.model_path), "max_seq_len": args.max_seq_len,
                          "max_batch": args.max_batch, "n_slots": args.n_slots, "sync_type": args.sync_type,
                          "ctl": f"{os.getpid()}"})
            ack = _recv_msg(s)
            if not ack.get("ok"):
                raise RuntimeError(f"worker {host}:{port} refused: {ack}")
        print(f"⭕ Socket[{i}]: connecting to {host}:{port} worker", flush=True)
    return world


def serve_worker_once(port: int) -> dict:
    """Worker side: wait for the root's bootstrap message (reference: NnNetwork::serve)."""
    srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
    srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
    srv.bind(("0.0.0.0", port))
    srv.listen(1)
    print(f"Listening on 0.0.0.0:{port}...", flush=True)
    conn, addr = srv.accept()
    with conn:
        cfg = _recv_msg(conn)
        _send_msg(conn, {"ok": True})
    srv.close()
    print(f"⭕ The root node has connected ({addr[0]}), rank {cfg['rank']}/{cfg['world']}", flush=True)
    return cfg


# --------------------------------------------------------------------------------------------------
# engine construction
# --------------------------------------------------------------------------------------------------
def use_cuda_device(device: str) -> bool:
    import torch
    return device == "cuda" or (device == "auto" and torch.cuda.is_available())


def build_engine(model_path: str, rank: int, world: int, max_batch: int, n_slots: int, max_seq_len: int,
                 sync_type: int, device: str = "auto", gpu_index: int = -1, workers=(), n_threads: int = 0):
    import torch
    from ..formats.mfile import read_header
    h = read_header(model_path, max_seq_len)
    if world > h.n_kv_heads:
        raise ValueError(f"This version does not support more nodes than the number of KV heads in the model "
                         f"({world} > {h.n_kv_heads})")
    if h.weight_type == quant.Q40 and sync_type != quant.Q80:
        raise ValueError("This version supports only Q40 weights with Q80 sync type")
    use_cuda = use_cuda_device(device)
    seq_len = h.seq_len
    slots = n_slots or max(2, min(max_batch, 64))
    if use_cuda and h.weight_type == quant.Q40:
        from ..models.llama import LlamaEngine
        if gpu_index >= 0:
            torch.cuda.set_device(gpu_index)
        # bucket_rows: the CLI and the API server run arbitrary batch shapes (prompt chunks, changing user
        # counts); padding them to powers of two bounds the number of CUDA graphs (same flag on every rank)
        return LlamaEngine(model_path, rank=rank, world=world, max_rows=max_batch, n_slots=slots,
                           max_seq_len=seq_len, sync_q80=True, bucket_rows=True)
    if use_cuda:
        # f32 / f16 weights with an f32 buffer: the library path (cuBLAS + NCCL), same engine surface
        from ..models.f32_engine import F32Engine
        if gpu_index >= 0:
            torch.cuda.set_device(gpu_index)
        return F32Engine(model_path, rank=rank, world=world, max_rows=max_batch, n_slots=slots,
                         max_seq_len=min(seq_len, 8192))
    if device == "cpu-oracle":
        # the PyTorch restatement of the reference numerics (tests, debugging); single node
        from ..models.cpu_engine import CpuEngine
        return CpuEngine(model_path, max_rows=max_batch, n_slots=slots, max_seq_len=min(seq_len, 4096))
    # no GPU: the native host backend; `--workers` are `dllama worker` processes reached over TCP, which get their
    # configuration and weight slices from this root (reference: src/app.cpp:233-312)
    from ..models.native_cpu_engine import NativeCpuEngine
    return NativeCpuEngine(model_path, workers=list(workers), n_threads=n_threads, max_rows=max_batch, n_slots=slots,
                           max_seq_len=min(seq_len, 8192) if max_seq_len <= 0 else seq_len, sync_type=sync_type)


def setup_distributed(args: AppArgs):
    """Root: returns (rank, world, control channel or None)."""
    from ..parallel.control import ControlChannel
    from ..parallel.symm import init_distributed
    env_world = int(os.environ.get("WORLD_SIZE", "1"))
    if env_world > 1:                                  # launched under torchrun
        rank, world, _ = init_distributed()
        ctl = ControlChannel(os.environ.get("MASTER_PORT", "0"), rank, world, args.max_batch)
        return rank, world, ctl
    if args.workers and not use_cuda_device(args.device):
        return 0, 1, None                               # the native CPU engine dials its workers itself
    if args.workers:
        master_port = 20000 + (os.getpid() % 20000)
        world = dial_workers(args, master_port)
        os.environ.update(RANK="0", WORLD_SIZE=str(world), LOCAL_RANK="0" if args.gpu_index < 0 else str(args.gpu_index),
                          MASTER_ADDR="127.0.0.1", MASTER_PORT=str(master_port))
        rank, world, _ = init_distributed()
        ctl = ControlChannel(str(os.getpid()), 0, world, args.max_batch)
        return 0, world, ctl
    return 0, 1, None


def run_worker_app(args: AppArgs):
    """`dllama worker`: serve -> join the rendezvous -> load own weight slice -> replay batches;
    on root disconnect go back to listening (reference: runWorkerApp, src/app.cpp:405-464)."""
    if not use_cuda_device(args.device):
        # no GPU: a worker of the native host backend; everything it needs arrives over the socket
        from ..models.native_cpu_engine import serve_worker
        while True:
            try:
                serve_worker(args.port, args.n_threads)
            except RuntimeError as e:               # a bad handshake or a root that failed during set-up
                print(f"🚨 {e}", flush=True)
                time.sleep(0.5)
            print("🚁 Root disconnected, waiting for a new one", flush=True)
    from ..parallel.control import ControlChannel
    from ..parallel.symm import init_distributed
    from ..serving.batch_engine import worker_loop
    import torch.distributed as dist
    while True:
        cfg = serve_worker_once(args.port)
        os.environ.update(RANK=str(cfg["rank"]), WORLD_SIZE=str(cfg["world"]),
                          LOCAL_RANK=str(args.gpu_index if args.gpu_index >= 0 else cfg["rank"]),
                          MASTER_ADDR=cfg["master_addr"], MASTER_PORT=str(cfg["master_port"]))
        rank, world, _ = init_distributed()
        engine = build_engine(cfg["model"], rank, world, cfg["max_batch"], cfg["n_slots"], cfg["max_seq_len"],
                              cfg["sync_type"], "cuda", -1)
        ctl = ControlChannel(cfg["ctl"], rank, world, cfg["max_batch"])
        print("🚁 Worker is ready", flush=True)
        worker_loop(engine, ctl)
        engine.close()
        ctl.close()
        if dist.is_initialized():
            dist.destroy_process_group()
        print("🚁 Root disconnected, waiting for a new one", flush=True)


class InferenceContext:
    """What the mode handlers get (reference: AppInferenceContext, src/app.hpp:84-94)."""

    def __init__(self, args: AppArgs):
        self.args = args
        if args.model_path is None:
            raise ValueError("Model is required")
        if args.tokenizer_path is None:
            raise ValueError("Tokenizer is required")
        self.rank, self.world, self.control = setup_distributed(args)
        self.tokenizer = T.Tokenizer(args.tokenizer_path)
        print(f"📄 BosId: {self.tokenizer.bos_id} ({self.tokenizer.piece(self.tokenizer.bos_id).decode('utf-8', 'replace')})")
        print("📄 EosId: " + " ".join(f"{t} ({self.tokenizer.piece(t).decode('utf-8', 'replace')})"
                                      for t in self.tokenizer.eos_token_ids))
        print(f"📄 RegularVocabSize: {self.tokenizer.regular_vocab_size}")
        print(f"📄 SpecialVocabSize: {self.tokenizer.special_vocab_size}")
        self.engine = build_engine(args.model_path, self.rank, self.world, args.max_batch, args.n_slots,
                                   args.max_seq_len, args.sync_type, args.device, args.gpu_index, args.workers,
                                   args.n_threads)
        self.header = self.engine.h
        print(self.header.describe(), flush=True)
        try:
            from ..models.memory import describe, required_memory
            print(describe(required_memory(self.header, max(1, getattr(self.engine, "world", 1)), self.engine.max_rows,
                                           self.engine.n_slots, self.engine.seq_len,
                                           "cuda" if self.engine.device.type == "cuda" else "cpu")), flush=True)
        except Exception as e:                 # the estimate is informational: never let it stop a run
            print(f"📀 RequiredMemory: n/a ({e})", flush=True)
        if self.tokenizer.vocab_size != self.header.vocab_size:
            print(f"⚠️  tokenizer vocab ({self.tokenizer.vocab_size}) != model vocab ({self.header.vocab_size})")
        self.sampler = T.Sampler(self.header.vocab_size, args.temperature, args.topp, args.seed)

    def close(self):
        from ..parallel.control import OP_STOP
        if self.control is not None and self.rank == 0:
            self.control.send(OP_STOP)
            self.control.close(unlink=True)
        self.engine.close()
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                dist.destroy_process_group()
        except Exception:
            pass


def run_inference_app(args: AppArgs, handler):
    """(reference: runInferenceApp, src/app.cpp:233-312 -- minus the never-joined thread, D5)"""
    ctx = InferenceContext(args)
    try:
        if ctx.rank != 0:
            from ..serving.batch_engine import worker_loop
            worker_loop(ctx.engine, ctx.control)
        else:
            handler(ctx)
    finally:
        ctx.close()
