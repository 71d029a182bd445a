"""Control plane between the root rank and the worker ranks of one node.

The reference broadcasts an 8-byte ``LlmControlPacket{position, batchSize}`` over its sockets before
every forward (src/app.cpp:198-229) and the workers derive row positions as ``position+i`` -- which
is what makes per-user positions impossible there (defect D2).  Here the packet carries the whole
batch descriptor (per-row token, position and KV slot, the first row that needs logits, the greedy
flag, and the slots to release), travels through a shared-memory mailbox (all ranks of a TP group
live on one NVSwitch node, so no sockets are needed on the hot path) and costs a few microseconds.

Protocol: single writer (rank 0), N-1 readers.  The writer fills ``payload`` then bumps ``seq``;
reader r copies the payload when ``seq`` changes and then publishes ``ack[r] = seq``.  The writer
waits for all acks of packet k before overwriting it with packet k+1.
"""
from __future__ import annotations

import mmap
import os
import time
from typing import List, Optional, Sequence

import numpy as np

OP_STOP, OP_FORWARD, OP_RESET, OP_WARMUP = 0, 1, 2, 3
_HDR = 8            # int32 words: op, M, logits_from, greedy, n_released, reserved...


class ControlChannel:
    def __init__(self, name: str, rank: int, world: int, max_rows: int, create: Optional[bool] = None):
        self.rank, self.world, self.max_rows = rank, world, max_rows
        self.n_words = _HDR + 3 * max_rows + max_rows          # header | tokens | pos | slots | released
        self.path = os.path.join("/dev/shm", f"dllama_b200_ctl_{name}")
        nbytes = 64 + 8 * 64 + 4 * self.n_words
        create = (rank == 0) if create is None else create
        self.send_timeout_s = float(os.environ.get("DLLAMA_CTL_TIMEOUT_S", "120"))
        if create:
            # a file of a crashed run with the same name must never be mapped by this run's workers: write the new
            # mailbox under a temporary name and rename it into place (workers that still map the old inode keep a
            # dead file; new workers check the run nonce below)
            tmp = self.path + f".{os.getpid()}.tmp"
            with open(tmp, "wb") as f:
                f.write(b"\0" * nbytes)
                f.seek(8)
                f.write(np.int64(self._nonce()).tobytes())
            os.replace(tmp, self.path)
        else:
            t0 = time.time()
            while True:
                try:
                    if os.path.getsize(self.path) >= nbytes:
                        with open(self.path, "rb") as f:
                            f.seek(8)
                            if np.frombuffer(f.read(8), dtype=np.int64)[0] == self._nonce():
                                break
                except OSError:
                    pass
                if time.time() - t0 > 60:
                    raise TimeoutError(f"control mailbox {self.path} of this run did not appear")
                time.sleep(0.01)
        self._f = open(self.path, "r+b")
        self._mm = mmap.mmap(self._f.fileno(), nbytes)
        self.seq = np.frombuffer(self._mm, dtype=np.int64, count=1, offset=0)
        self.fail = np.frombuffer(self._mm, dtype=np.int64, count=1, offset=16)      # set by a worker whose step failed
        self.ack = np.frombuffer(self._mm, dtype=np.int64, count=64, offset=64)
        self.payload = np.frombuffer(self._mm, dtype=np.int32, count=self.n_words, offset=64 + 8 * 64)
        self._seen = 0

    def _nonce(self) -> int:
        """Identifies the run.  The mailbox name already does in worker mode (it carries the root's pid); under
        torchrun the name is only the rendezvous port, so the launcher's pid (the common parent of all ranks) is added."""
        key = os.path.basename(self.path) + ":" + os.environ.get("DLLAMA_RUN_NONCE", "")
        if "TORCHELASTIC_RUN_ID" in os.environ:
            key += ":" + os.environ["TORCHELASTIC_RUN_ID"] + ":" + str(os.getppid())
        h = 1469598103934665603
        for ch in key.encode():
            h = ((h ^ ch) * 1099511628211) & 0x7FFFFFFFFFFFFFFF
        return h or 1

    def report_failure(self, msg: str):
        """Worker side: tell the root that this rank cannot go on (the root's next send raises)."""
        import sys
        sys.stderr.write(f"[control] rank {self.rank} failed: {msg}\n")
        self.fail[0] = self.rank + 1

    # ---- root -------------------------------------------------------------------------------
    def send(self, op: int, tokens: Sequence[int] = (), positions: Sequence[int] = (), slots: Sequence[int] = (),
             logits_from: int = 0, greedy: bool = False, released: Sequence[int] = ()):
        k = int(self.seq[0])
        t0 = time.time()
        for r in range(1, self.world):                      # previous packet consumed by everyone?
            spins = 0
            while int(self.ack[r]) < k:
                spins += 1
                if spins & 0xFFFF == 0:
                    if int(self.fail[0]) != 0:
                        raise RuntimeError(f"worker rank {int(self.fail[0]) - 1} reported a failed step")
                    if time.time() - t0 > self.send_timeout_s:
                        raise TimeoutError(f"worker rank {r} did not take the previous batch within {self.send_timeout_s:.0f} s")
        if int(self.fail[0]) != 0:
            raise RuntimeError(f"worker rank {int(self.fail[0]) - 1} reported a failed step")
        M, R = len(tokens), self.max_rows
        p = self.payload
        p[0], p[1], p[2], p[3], p[4] = op, M, logits_from, int(greedy), len(released)
        p[_HDR:_HDR + M] = tokens
        p[_HDR + R:_HDR + R + M] = positions
        p[_HDR + 2 * R:_HDR + 2 * R + M] = slots
        p[_HDR + 3 * R:_HDR + 3 * R + len(released)] = released
        self.seq[0] = k + 1                                 # publish (x86 TSO: payload stores are ordered before)

    # ---- workers ------------------------------------------------------------------------------
    def recv(self, timeout_s: Optional[float] = None):
        """Blocks (spinning, then yielding) for the next packet; returns a dict."""
        t0 = time.time()
        spins = 0
        while int(self.seq[0]) == self._seen:
            spins += 1
            if spins > 20000:
                time.sleep(0.0002)                          # idle: stop burning the core
                if timeout_s is not None and time.time() - t0 > timeout_s:
                    return None
        k = int(self.seq[0])
        p, R = self.payload, self.max_rows
        M, nrel = int(p[1]), int(p[4])
        pkt = {"op": int(p[0]), "M": M, "logits_from": int(p[2]), "greedy": bool(p[3]),
               "tokens": p[_HDR:_HDR + M].tolist(), "positions": p[_HDR + R:_HDR + R + M].tolist(),
               "slots": p[_HDR + 2 * R:_HDR + 2 * R + M].tolist(),
               "released": p[_HDR + 3 * R:_HDR + 3 * R + nrel].tolist()}
        self._seen = k
        self.ack[self.rank] = k
        return pkt

    def close(self, unlink: bool = False):
        self.seq = self.ack = self.payload = None
        try:
            self._mm.close()
        except BufferError:
            pass
        self._f.close()
        if unlink:
            try:
                os.unlink(self.path)
            except OSError:
                pass
