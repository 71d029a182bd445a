"""`dllama-api` on the native front end (csrc/host/server.cpp).

The C++ server owns the sockets, HTTP/JSON/SSE, chat templates, tokenisation, the scheduler, the per-request
sampler state, detokenisation and stop strings.  This module is the engine thread: it asks the server for the
next batch, runs the forward pass (and the device sampler) on the engine, and hands the sampled token ids back.
Per step it makes three library calls and no per-token Python work, and no other Python thread exists, so the
GPU launch path never waits for the interpreter lock (the Python front end in serving/api.py shares it with one
handler thread per connection: measured 3.9 k tok/s over HTTP for 32 users on an engine that does 9.9 k).

Reference surface: src/dllama-api.cpp:338-365 (routes), src/app.cpp:314-402 (the fork's batching loop).
"""
from __future__ import annotations

import ctypes as C
import os
import threading
import time
from typing import Optional

import torch

from ..parallel.control import OP_FORWARD, OP_STOP, OP_WARMUP, ControlChannel
from ..text import native as T
from . import api_types as A


class NativeApiServer:
    def __init__(self, engine, tokenizer: T.Tokenizer, model_name: str, defaults: A.InferenceParams,
                 chat_template_type: int = T.CHAT_UNKNOWN, control: Optional[ControlChannel] = None,
                 prefill_chunk: Optional[int] = None, host: str = "0.0.0.0", port: int = 9990,
                 max_default_tokens: int = 128, verbose: bool = True):
        self.engine = engine
        self.tok = tokenizer
        self.control = control
        self.host, self.port = host, port
        self.max_rows = engine.max_rows
        self._L = T.host_lib()
        eos_piece = tokenizer.piece(tokenizer.eos_token_ids[-1]) if tokenizer.eos_token_ids else b""
        # template detection (and its errors) as in the Python front end
        tmpl = T.ChatTemplateGenerator(chat_template_type, tokenizer.chat_template, eos_piece.decode("utf-8", "replace"))
        R = self.max_rows
        self._h = self._L.dl_srv_new(tokenizer._h, model_name.encode("utf-8"), tmpl.type, eos_piece, engine.h.vocab_size,
                                     R, engine.n_slots, prefill_chunk or R, engine.seq_len,
                                     float(defaults.temperature), float(defaults.top_p),
                                     int(defaults.seed) & 0xFFFFFFFFFFFFFFFF, max_default_tokens, int(verbose))
        if not self._h:
            raise RuntimeError("dl_srv_new failed")
        if hasattr(engine, "usable_pages"):                   # paged KV: admit by free pages, not only by free slot
            self._L.dl_srv_set_pages(self._h, engine.usable_pages(), 32)
        dev = getattr(engine, "device", torch.device("cpu"))
        self._cuda = dev.type == "cuda"
        self._device_sampler = self._cuda and os.environ.get("DLLAMA_HOST_SAMPLER", "0") != "1"
        pin = (lambda t: t.pin_memory()) if torch.cuda.is_available() else (lambda t: t)
        # batch descriptor + sampling parameters are written by the server straight into (pinned) host tensors
        self._desc = torch.zeros(3, R, dtype=torch.int32)                       # tokens | positions | slots
        self._desc_np = self._desc.numpy()                                       # same memory, cheaper slicing
        self._h_params = pin(torch.zeros(3, R, dtype=torch.float32))            # temperature | topp | coin
        self._h_tok = pin(torch.zeros(R, dtype=torch.int32))
        self._released = (C.c_int * 256)()
        self._ptrs = tuple(t.data_ptr() for t in (self._desc[0], self._desc[1], self._desc[2], self._h_params[0],
                                                  self._h_params[1], self._h_params[2]))
        self._lf, self._ns, self._greedy, self._nrel = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
        if self._device_sampler:
            from ..ops import kernels as _K
            self._K = _K
            self._d_params = torch.zeros(3, R, dtype=torch.float32, device=dev)
            self._d_tok = torch.zeros(R, dtype=torch.int32, device=dev)
        else:
            self._logits_host = pin(torch.empty(R, engine.h.vocab_size, dtype=torch.float32))
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self.n_errors = 0
        # DLLAMA_STEP_LOG=1: (start time, rows, first logits row, sampled rows, seconds) of every step, for load tests
        self.step_log = [] if os.environ.get("DLLAMA_STEP_LOG") == "1" else None

    # ---- lifecycle -------------------------------------------------------------------------------------
    def warmup(self):
        if getattr(self.engine, "bucket_rows", False) and getattr(self.engine, "use_graphs", False):
            if self.control is not None:
                self.control.send(OP_WARMUP)
            self.engine.warmup_graphs()

    def listen(self) -> int:
        port = self._L.dl_srv_listen(self._h, self.host.encode(), self.port)
        if port < 0:
            raise OSError(-port, f"cannot listen on {self.host}:{self.port}: {os.strerror(-port)}")
        self.port = port
        return port

    def start(self):
        """Listen and run the engine loop on a background thread (tests, embedding in another program)."""
        self.listen()
        self._thread = threading.Thread(target=self.run, name="dllama-engine", daemon=True)
        self._thread.start()

    def serve_forever(self):
        self.listen()
        print(f"Server URL: http://127.0.0.1:{self.port}/v1/", flush=True)
        self.run()

    def shutdown(self):
        self._stop.set()
        if self._thread is not None and self._thread is not threading.current_thread():
            self._thread.join(timeout=30)
        if self.control is not None:
            self.control.send(OP_STOP)
        if self._h:
            self._L.dl_srv_shutdown(self._h)

    def close(self):
        self.shutdown()
        if self._h:
            self._L.dl_srv_free(self._h)
            self._h = None

    def stats(self) -> dict:
        out = (C.c_longlong * 10)()
        self._L.dl_srv_stats(self._h, out)
        keys = ("requests", "errors", "steps", "generated", "rows", "waiting", "running", "free_slots", "connections",
                "engine_errors")
        return dict(zip(keys, [int(v) for v in out]))

    # ---- engine loop -----------------------------------------------------------------------------------
    def run(self):
        if self._cuda:
            torch.cuda.set_device(self.engine.device)
        L, h = self._L, self._h
        while not self._stop.is_set():
            if not L.dl_srv_wait(h, 20):
                continue
            try:
                self.step()
            except Exception as e:                      # one bad batch must not take the server down
                self.n_errors += 1
                print(f"🚨 engine step failed: {type(e).__name__}: {e}", flush=True)
                # (the slots of the failed requests come back through the next dl_srv_next_batch)
                L.dl_srv_fail_batch(h, f"{type(e).__name__}: {e}".encode("utf-8", "replace"))

    def step(self) -> int:
        """One scheduler step = one forward pass.  Returns the number of rows run."""
        L, h, eng, R = self._L, self._h, self.engine, self.max_rows
        p = self._ptrs
        t_step = time.perf_counter() if self.step_log is not None else 0.0
        n = L.dl_srv_next_batch(h, p[0], p[1], p[2], R, C.byref(self._lf), C.byref(self._ns), p[3], p[4], p[5],
                                int(self._device_sampler), C.byref(self._greedy), C.addressof(self._released), 256,
                                C.byref(self._nrel))
        lf, n_s, greedy, n_rel = self._lf.value, self._ns.value, bool(self._greedy.value), self._nrel.value
        released = [self._released[i] for i in range(n_rel)]
        for s in released:
            eng.pages.release_pages(s)
        if n <= 0:
            return 0
        dn = self._desc_np
        toks, pos, slots = dn[0, :n].tolist(), dn[1, :n].tolist(), dn[2, :n].tolist()
        if self.control is not None:
            # workers replay exactly the root's launch: a pure prompt step computes the last row's logits only
            for s_, p_ in zip(slots, pos):          # a page shortage must surface before the workers are told to run
                self.engine.ensure_capacity(int(s_), int(p_) + 1)
            self.control.send(OP_FORWARD, toks, pos, slots, lf if n_s else n - 1, greedy, released)
        if n_s == 0:
            eng.forward(toks, pos, slots, logits_from=n - 1)                  # pure prompt chunk
            rc = L.dl_srv_commit(h, 0, None)
        elif greedy:
            sampled = eng.decode_step_greedy(toks, pos, slots, lf)[:n_s]
            arr = (C.c_int * n_s)(*sampled)
            rc = L.dl_srv_commit(h, n_s, C.addressof(arr))
        elif self._device_sampler:
            logits = eng.forward(toks, pos, slots, logits_from=lf)
            self._d_params.copy_(self._h_params, non_blocking=True)
            self._K.sample_rows(logits[:n_s], self._d_params[0], self._d_params[1], self._d_params[2], self._d_tok)
            self._h_tok.copy_(self._d_tok, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            rc = L.dl_srv_commit(h, n_s, self._h_tok.data_ptr())
        else:
            logits = eng.forward(toks, pos, slots, logits_from=lf)
            lh = self._logits_host
            lh[:n_s].copy_(logits[:n_s], non_blocking=True)
            if logits.is_cuda:
                torch.cuda.current_stream().synchronize()
            L.dl_srv_sample_host(h, lh.data_ptr(), lh.stride(0), self._h_tok.data_ptr())
            rc = L.dl_srv_commit(h, n_s, self._h_tok.data_ptr())
        if rc < 0:
            raise RuntimeError(f"dl_srv_commit failed ({rc})")
        if self.step_log is not None:
            self.step_log.append((t_step, n, lf, n_s, time.perf_counter() - t_step))
        return n
