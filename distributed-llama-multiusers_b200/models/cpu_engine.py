"""CPU engine with the same interface as ``LlamaEngine`` (slots, batched rows with per-row
positions), built on the PyTorch oracle.  It is what lets the scheduler / batch engine / HTTP API /
CLI run and be tested on a machine without a GPU, and it is the "plumbing, no GPU" configuration of
BASELINE.json.  Slow by design; numerics are the reference's (see models/oracle.py).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import torch

from ..formats.mfile import ModelFile
from .oracle import OracleLlama


class _Pages:
    """No paging on CPU: every slot owns a dense cache.  Same surface as PageAllocator."""

    def __init__(self, n_slots: int):
        self.free_slots = list(range(n_slots - 1, -1, -1))

    def alloc_slot(self) -> int:
        if not self.free_slots:
            raise RuntimeError("no free sequence slot")
        return self.free_slots.pop()

    def free_slot(self, slot: int):
        self.free_slots.append(slot)

    def release_pages(self, slot: int):
        pass

    def ensure(self, slot: int, n_tokens: int):
        pass


class CpuEngine:
    def __init__(self, model_path: str, max_rows: int = 32, n_slots: int = 4, max_seq_len: int = 512,
                 n_nodes: int = 1, **_unused):
        self.mf = ModelFile(model_path, max_seq_len=max_seq_len)
        self.h = self.mf.header
        self.max_rows, self.n_slots = max_rows, n_slots
        self.seq_len = min(self.h.seq_len, max_seq_len)
        self.device = torch.device("cpu")
        self.rank, self.world = 0, 1
        self._model = OracleLlama(self.mf, n_nodes=n_nodes, max_seq_len=self.seq_len)
        # per-slot caches (the oracle object is re-pointed at the active slot's cache)
        L, kv = self.h.n_layers, self.h.kv_dim
        self._k = [[torch.zeros(self.seq_len, kv) for _ in range(L)] for _ in range(n_slots)]
        self._v = [[torch.zeros(self.seq_len, kv) for _ in range(L)] for _ in range(n_slots)]
        self.pages = _Pages(n_slots)
        self.logits = torch.zeros(max_rows, self.h.vocab_size)

    def alloc_slot(self) -> int:
        return self.pages.alloc_slot()

    def free_slot(self, slot: int):
        self.pages.free_slot(slot)

    def ensure_capacity(self, slot: int, n_tokens: int):
        if n_tokens > self.seq_len:
            raise RuntimeError(f"sequence of {n_tokens} tokens exceeds max_seq_len")

    def forward(self, tokens: Sequence[int], positions: Sequence[int], slots: Sequence[int], logits_from: int = 0,
                greedy: bool = False) -> torch.Tensor:
        M = len(tokens)
        assert 1 <= M <= self.max_rows
        i = 0
        while i < M:
            # run maximal runs of consecutive positions of one slot as one chunk
            j = i + 1
            while j < M and slots[j] == slots[i] and positions[j] == positions[j - 1] + 1:
                j += 1
            s = slots[i]
            self._model.k_cache, self._model.v_cache = self._k[s], self._v[s]
            lg = self._model.forward(list(tokens[i:j]), int(positions[i]))
            self.logits[i:j] = lg
            i = j
        return self.logits[logits_from:M]

    def prefill(self, slot: int, tokens: Sequence[int], start_pos: int = 0, chunk: Optional[int] = None):
        chunk = chunk or self.max_rows
        out = None
        for i in range(0, len(tokens), chunk):
            part = list(tokens[i:i + chunk])
            pos = list(range(start_pos + i, start_pos + i + len(part)))
            out = self.forward(part, pos, [slot] * len(part), logits_from=len(part) - 1)
        return out

    def decode_step_greedy(self, tokens, positions, slots, logits_from: int = 0) -> List[int]:
        lg = self.forward(tokens, positions, slots, logits_from)
        return [int(t) for t in torch.argmax(lg, dim=-1)]

    def decode_greedy_host(self, tokens, positions, slots) -> List[int]:
        return self.decode_step_greedy(tokens, positions, slots, 0)

    def close(self):
        self.mf.close()
