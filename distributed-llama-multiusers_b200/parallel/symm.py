"""Symmetric (peer-mapped) device buffers and the process-group bootstrap.

One process per GPU.  ``torch.distributed`` (NCCL on GPUs, gloo on CPU) is used only for the
rendezvous -- exchanging CUDA IPC handles -- and as the in-repo *baseline* collective path.  The
data plane of the product is loads/stores issued from inside the kernels to the peer pointers
held by ``SymmetricBuffer``.

Replaces the reference's mesh bootstrap (NnNetwork::connect / serve, src/nn/nn-network.cpp:264-348):
instead of a full mesh of TCP sockets every rank ends up with a table of peer *pointers*.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops.lib import check, cuda_lib


def init_distributed(backend: Optional[str] = None) -> tuple:
    """Initialise torch.distributed from the torchrun environment; returns (rank, world, local_rank)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    if torch.cuda.is_available():
        torch.cuda.set_device(local_rank % torch.cuda.device_count())
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kwargs = {}
        if backend == "nccl":
            kwargs["device_id"] = torch.device("cuda", torch.cuda.current_device())
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
    return rank, world, local_rank


def validate_tp(world: int, n_kv_heads: int, dim: int, hidden_dim: int, n_heads: int, vocab: int, q40: bool = True):
    """The reference's slicing rules (src/app.cpp:237-240, src/nn/nn-core.cpp:198-266,291):
    N = 2^n, N <= nKvHeads, every sliced dimension divisible; the Q40T tiling additionally needs
    the per-rank K extents to be multiples of 128."""
    if world < 1 or (world & (world - 1)) != 0:
        raise ValueError(f"number of nodes must be a power of two, got {world}")
    if world > n_kv_heads:
        raise ValueError(f"This version does not support more nodes than the number of KV heads in the model "
                         f"({world} > {n_kv_heads})")
    for name, v in (("dim", dim), ("hiddenDim", hidden_dim), ("nHeads", n_heads), ("nKvHeads", n_kv_heads),
                    ("vocabSize", vocab)):
        if v % world != 0:
            raise ValueError(f"{name}={v} is not divisible by nNodes={world}")
    if q40:
        if (dim // world) % 128 != 0 or (hidden_dim // world) % 128 != 0 or dim % 128 != 0:
            raise ValueError("Q40 tiling needs dim, dim/nNodes and hiddenDim/nNodes to be multiples of 128")


class SymmetricBuffer:
    """A zero-initialised device allocation of the same size on every rank, with every peer's copy
    mapped into this process (CUDA IPC over NVLink).  ``ptrs[r]`` is rank r's copy as seen from
    here; ``ptrs[rank]`` is the local allocation."""

    def __init__(self, nbytes: int, rank: int = 0, world: int = 1, group=None):
        self.nbytes = int(nbytes)
        self.rank, self.world = rank, world
        L = cuda_lib()
        local = C.c_void_p()
        handle = (C.c_uint8 * 64)()
        check(L.dl_ipc_alloc(C.byref(local), self.nbytes, handle if world > 1 else None), "dl_ipc_alloc")
        self.local_ptr = local.value
        self.ptrs: List[int] = [0] * world
        self.ptrs[rank] = self.local_ptr
        self._opened: List[int] = []
        if world > 1:
            handles: List[Optional[bytes]] = [None] * world
            dist.all_gather_object(handles, bytes(handle), group=group)
            for r in range(world):
                if r == rank:
                    continue
                p = C.c_void_p()
                buf = (C.c_uint8 * 64).from_buffer_copy(handles[r])
                check(L.dl_ipc_open(buf, C.byref(p)), f"dl_ipc_open(rank {r})")
                self.ptrs[r] = p.value
                self._opened.append(p.value)
            dist.barrier(group=group)

    def tensor(self, offset: int, shape, dtype: torch.dtype) -> torch.Tensor:
        """A torch view of the LOCAL copy (no ownership; keep this object alive)."""
        numel = 1
        for s in shape:
            numel *= s
        nbytes = numel * torch.empty(0, dtype=dtype).element_size()
        assert offset + nbytes <= self.nbytes
        return _cuda_view(self.local_ptr + offset, shape, dtype)

    def close(self):
        L = cuda_lib()
        for p in self._opened:
            L.dl_ipc_close(p)
        self._opened = []
        if self.local_ptr:
            L.dl_ipc_free(self.local_ptr)
            self.local_ptr = 0


class _CudaArrayInterface:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"data": (ptr, False), "shape": tuple(shape), "typestr": typestr,
                                         "version": 2, "strides": None}


_TYPESTR = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1", torch.float16: "<f2",
            torch.int8: "|i1"}


def _cuda_view(ptr: int, shape, dtype: torch.dtype) -> torch.Tensor:
    if dtype == torch.uint32:
        return torch.as_tensor(_CudaArrayInterface(ptr, shape, "<i4"), device="cuda")
    return torch.as_tensor(_CudaArrayInterface(ptr, shape, _TYPESTR[dtype]), device="cuda")
