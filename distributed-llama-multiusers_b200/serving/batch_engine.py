"""Continuous-batching generation loop: many concurrent requests share every forward pass.

This is the *intended* behaviour of the fork's ``inference_loop`` (src/app.cpp:314-402) without its
defects: each request keeps its own prompt (prefilled in chunks), position,
KV slot, sampler (temperature / top-p / seed per request), streaming UTF-8 decoder and stop
detector; a finished request leaves the batch and a waiting one joins at the next step.  Batch
formation lives in the C++ scheduler (csrc/host/scheduler.cpp); this module drives the GPU engine,
samples, detokenises and hands text deltas to whoever submitted the request.
"""
from __future__ import annotations

import itertools
import os
import queue
import threading
import time
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence

import numpy as np
import torch

from ..parallel.control import OP_FORWARD, OP_STOP, OP_WARMUP, ControlChannel
from ..text import native as T


PAGE_TOKENS = 32          # tokens per KV page (ops.kernels.PAGE_SIZE)


@dataclass
class GenerationRequest:
    prompt_tokens: List[int]
    max_tokens: int = 128
    temperature: float = 0.8
    top_p: float = 0.9
    seed: int = 0
    stop: Sequence[str] = ()
    # filled by the engine
    id: int = -1
    out: "queue.Queue" = field(default_factory=queue.Queue)      # str deltas, then None
    text: str = ""
    tokens: List[int] = field(default_factory=list)
    finish_reason: Optional[str] = None
    error: Optional[str] = None                 # set when finish_reason == "error"
    t_submit: float = 0.0
    t_first_token: Optional[float] = None
    t_done: Optional[float] = None
    _sampler: Optional[T.Sampler] = None
    _decoder: Optional[T.StreamDecoder] = None
    _stops: Optional[T.EosDetector] = None

    def result(self, timeout: Optional[float] = None) -> str:
        """Block until the request finished; returns the generated text."""
        while True:
            d = self.out.get(timeout=timeout)
            if d is None:
                return self.text


class BatchEngine:
    def __init__(self, engine, tokenizer: T.Tokenizer, max_batch: Optional[int] = None,
                 prefill_chunk: Optional[int] = None, control: Optional[ControlChannel] = None):
        self.engine = engine
        self.tok = tokenizer
        self.max_rows = engine.max_rows
        self.sched = T.Scheduler(self.max_rows, engine.n_slots, prefill_chunk or self.max_rows, engine.seq_len)
        if hasattr(engine, "usable_pages"):                   # paged KV: admit by free pages, not only by free slot
            self.sched.set_pages(engine.usable_pages(), PAGE_TOKENS)
        self.control = control
        self.reqs = {}
        self._ids = itertools.count(1)
        self._lock = threading.Lock()
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self.n_steps = 0
        self.n_generated = 0
        self.n_rows = 0
        self.n_errors = 0
        self._last_ids, self._last_sampling = [], []
        # seconds spent forming batches / in the engine (incl. sampling) / detokenising + bookkeeping
        self.t_sched = self.t_engine = self.t_emit = 0.0
        eos_pieces = [tokenizer.piece(t).decode("utf-8", "replace") for t in tokenizer.eos_token_ids]
        self._eos_pieces = eos_pieces
        self._max_stop = max([len(p) for p in eos_pieces] + [1])
        # sampling: on a GPU engine the rows are sampled by `sample_kernel` (only token ids cross PCIe);
        # the host sampler (reference-exact arithmetic) serves CPU engines and DLLAMA_HOST_SAMPLER=1
        dev = getattr(engine, "device", torch.device("cpu"))
        self._device_sampler = dev.type == "cuda" and os.environ.get("DLLAMA_HOST_SAMPLER", "0") != "1"
        if self._device_sampler:
            from ..ops import kernels as _K
            self._K = _K
            R = self.max_rows
            self._h_params = torch.zeros(3, R, dtype=torch.float32).pin_memory()      # temperature | topp | coin
            self._d_params = torch.zeros(3, R, dtype=torch.float32, device=dev)
            self._d_tok = torch.zeros(R, dtype=torch.int32, device=dev)
            self._h_tok = torch.zeros(R, dtype=torch.int32).pin_memory()
            self._logits_host = None
        else:
            self._logits_host = torch.empty(self.max_rows, engine.h.vocab_size, dtype=torch.float32)
            if torch.cuda.is_available():
                self._logits_host = self._logits_host.pin_memory()

    # ---- client side (any thread) ------------------------------------------------------------------
    def submit(self, req: GenerationRequest) -> GenerationRequest:
        req.id = next(self._ids)
        req.t_submit = time.time()
        req._sampler = T.Sampler(self.engine.h.vocab_size, req.temperature, req.top_p, req.seed or req.id)
        req._decoder = self.tok.new_decoder()
        stops = list(self._eos_pieces) + list(req.stop)
        stop_ids = list(self.tok.eos_token_ids) + [-1] * len(req.stop)
        pad = max(self._max_stop, max([len(s) for s in req.stop] + [1]))
        req._stops = T.EosDetector(stop_ids, stops, pad, pad)
        with self._lock:
            self.reqs[req.id] = req
        rc = self.sched.submit(req.id, req.prompt_tokens, req.max_tokens)
        if rc != 0:
            with self._lock:
                del self.reqs[req.id]
            raise ValueError({-1: "prompt does not fit the context window", -2: "duplicate request id",
                              -3: "empty prompt"}.get(rc, f"scheduler error {rc}"))
        return req

    def cancel(self, req: GenerationRequest):
        self.sched.cancel(req.id)

    # ---- engine loop (one thread) --------------------------------------------------------------------
    def warmup(self):
        """Pre-capture the engine's CUDA graphs on every rank (call before start(); no-op for engines
        without bucketed shapes)."""
        if getattr(self.engine, "bucket_rows", False) and getattr(self.engine, "use_graphs", False):
            if self.control is not None:
                self.control.send(OP_WARMUP)
            self.engine.warmup_graphs()

    def start(self):
        self._thread = threading.Thread(target=self.run, name="dllama-batch-engine", daemon=True)
        self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=30)
        if self.control is not None:
            self.control.send(OP_STOP)

    def run(self):
        if self.engine.device.type == "cuda":
            torch.cuda.set_device(self.engine.device)
        while not self._stop.is_set():
            if not self.sched.wait(20):
                continue
            try:
                self.step()
            except Exception as e:                      # one bad batch must not take the server down
                self.n_errors += 1
                print(f"🚨 engine step failed: {type(e).__name__}: {e}", flush=True)
                self._fail_batch(e)

    def step(self) -> int:
        """One scheduler step = one forward pass.  Returns the number of rows run."""
        t_a = time.perf_counter()
        ids, toks, pos, slots, lf, sampling = self.sched.next_batch()
        released = self.sched.pop_released()
        for s in released:
            self.engine.pages.release_pages(s)
        self._last_ids, self._last_sampling = ids, sampling
        if not ids:
            self._finish_cancelled()
            return 0
        with self._lock:
            rows = [self.reqs.get(i) for i in sampling]
        greedy = all(r is not None and r.temperature == 0.0 for r in rows) and len(rows) > 0
        if self.control is not None:
            # workers replay exactly the root's launch: a pure prompt step computes the last row's logits only
            for s_, p_ in zip(slots, pos):          # a page shortage must surface before the workers are told to run
                self.engine.ensure_capacity(int(s_), int(p_) + 1)
            self.control.send(OP_FORWARD, toks, pos, slots, lf if sampling else len(toks) - 1, greedy, released)
        eng = self.engine
        n_s = len(sampling)
        t_b = time.perf_counter()
        self.t_sched += t_b - t_a
        if n_s == 0:
            eng.forward(toks, pos, slots, logits_from=len(toks) - 1)       # pure prefill chunk
            self.n_steps += 1
            self.n_rows += len(ids)
            self.t_engine += time.perf_counter() - t_b
            return len(ids)
        if greedy:
            sampled = eng.decode_step_greedy(toks, pos, slots, lf)[:n_s]
        elif self._device_sampler:
            logits = eng.forward(toks, pos, slots, logits_from=lf)
            hp = self._h_params.numpy()
            for i, r in enumerate(rows):
                if r is None:
                    hp[0, i], hp[1, i], hp[2, i] = 0.0, 0.0, 0.0
                else:
                    hp[0, i], hp[1, i] = r.temperature, r.top_p
                    hp[2, i] = r._sampler.coin() if r.temperature != 0.0 else 0.0
            self._d_params.copy_(self._h_params, non_blocking=True)
            self._K.sample_rows(logits[:n_s], self._d_params[0], self._d_params[1], self._d_params[2], self._d_tok)
            self._h_tok.copy_(self._d_tok, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            sampled = self._h_tok[:n_s].tolist()
        else:
            logits = eng.forward(toks, pos, slots, logits_from=lf)
            self._logits_host[:n_s].copy_(logits[:n_s], non_blocking=True)
            if logits.is_cuda:
                torch.cuda.current_stream().synchronize()
            lh = self._logits_host.numpy()
            sampled = [r._sampler.sample(lh[i]) if r is not None else 0 for i, r in enumerate(rows)]
        t_c = time.perf_counter()
        self.t_engine += t_c - t_b
        now = time.time()
        stop_flags = []
        for r, t in zip(rows, sampled):
            stop_flags.append(self._emit(r, t, now) if r is not None else True)
        finished = self.sched.commit(sampled, stop_flags)
        for r, fin in zip(rows, finished):
            if r is not None and fin:
                self._finalize(r, "stop" if r.finish_reason == "stop" else "length")
        self.n_steps += 1
        self.n_rows += len(ids)
        self.n_generated += n_s
        self.t_emit += time.perf_counter() - t_c
        return len(ids)

    # ---- per-token text handling -----------------------------------------------------------------------
    def _emit(self, r: GenerationRequest, token: int, now: float) -> bool:
        """Feed a sampled token through decoder + stop detector; returns True if generation stops."""
        if r.t_first_token is None:
            r.t_first_token = now
        r.tokens.append(token)
        piece = r._decoder.push(token)
        kind = r._stops.append(token, piece)
        if kind == T.EOS_NO or kind == T.EOS_YES:
            delta = r._stops.get_delta()
            if delta:
                r.text += delta
                r.out.put(delta)
            r._stops.reset()
        if kind == T.EOS_YES:
            r.finish_reason = "stop"
            return True
        return False

    def _finalize(self, r: GenerationRequest, reason: str):
        if r.finish_reason is None:
            # flush text held back while a stop string might have been forming
            delta = r._stops.get_delta()
            if delta:
                r.text += delta
                r.out.put(delta)
        r.finish_reason = reason
        r.t_done = time.time()
        r.out.put(None)
        with self._lock:
            self.reqs.pop(r.id, None)
        self.sched.forget(r.id)

    def _fail_batch(self, exc: Exception):
        """A forward raised (e.g. the KV page pool is exhausted): the scheduler has already advanced the
        requests of that batch, so they cannot be retried -- end them with finish_reason "error", give their
        slots and pages back, keep serving everybody else (reference: the API loop swallows per-request
        exceptions, src/dllama-api.cpp:362-364)."""
        ids = list(dict.fromkeys(self._last_ids))
        n_s = len(self._last_sampling)
        if n_s:
            try:
                self.sched.commit([0] * n_s, [True] * n_s)          # finishes the sampling rows' requests
            except Exception:
                pass
        for rid in ids:
            self.sched.cancel(rid)
            with self._lock:
                r = self.reqs.get(rid)
            if r is not None:
                r.error = f"{type(exc).__name__}: {exc}"
                self._finalize(r, "error")
        for s in self.sched.pop_released():
            self.engine.pages.release_pages(s)

    def _finish_cancelled(self):
        with self._lock:
            pending = list(self.reqs.values())
        # requests cancelled while waiting never produce a commit: close their streams
        st = self.sched.stats()
        if st["waiting"] == 0 and st["running"] == 0:
            for r in pending:
                self._finalize(r, "cancelled")


def worker_loop(engine, control: ControlChannel):
    """Worker ranks: replay the root's batches (reference: runWorkerApp loop, src/app.cpp:405-464).  A step that
    fails on a worker is reported to the root through the mailbox (the root's next send raises) instead of leaving the
    root's kernels spinning on a rank that is gone."""
    torch.cuda.set_device(engine.device)
    while True:
        pkt = control.recv()
        if pkt is None or pkt["op"] == OP_STOP:
            return
        try:
            for s in pkt["released"]:
                engine.pages.release_pages(s)
            if pkt["op"] == OP_WARMUP:
                engine.warmup_graphs()
                continue
            if pkt["op"] == OP_FORWARD and pkt["M"] > 0:
                M = engine.set_batch(pkt["tokens"], pkt["positions"], pkt["slots"])
                engine._run_batch(M, pkt["logits_from"], pkt["greedy"])
        except Exception as exc:        # noqa: BLE001 -- whatever it is, the root has to hear about it
            control.report_failure(f"{type(exc).__name__}: {exc}")
            raise
