"""ctypes binding of libdllama_b200_host.so: tokenizer, sampler, chat templates, stop detector and
the continuous-batching scheduler (csrc/host/*.cpp)."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

from ..ops.lib import HOST_LIB_PATH, NativeLibraryError

_host = None


def host_lib():
    global _host
    if _host is not None:
        return _host
    if not os.path.exists(HOST_LIB_PATH):
        raise NativeLibraryError(f"{HOST_LIB_PATH} is missing: run `python build.py`")
    L = C.CDLL(HOST_LIB_PATH)
    vp, i32, f32, u64, cp, ll = C.c_void_p, C.c_int, C.c_float, C.c_ulonglong, C.c_char_p, C.c_longlong
    ip = C.POINTER(C.c_int)
    sig = {
        "dl_tok_load": (vp, [cp, cp, i32]), "dl_tok_free": (None, [vp]), "dl_tok_vocab_size": (i32, [vp]),
        "dl_tok_bos": (i32, [vp]), "dl_tok_n_eos": (i32, [vp]), "dl_tok_eos": (i32, [vp, i32]),
        "dl_tok_n_regular": (i32, [vp]), "dl_tok_max_len": (i32, [vp]), "dl_tok_chat_template": (cp, [vp]),
        "dl_tok_piece": (vp, [vp, i32, ip]), "dl_tok_score": (f32, [vp, i32]), "dl_tok_is_eos": (i32, [vp, i32]),
        "dl_tok_encode": (i32, [vp, cp, i32, i32, i32, ip, i32]),
        "dl_dec_new": (vp, [vp]), "dl_dec_free": (None, [vp]), "dl_dec_reset": (None, [vp]),
        "dl_dec_push": (i32, [vp, i32, cp, i32]),
        "dl_sampler_new": (vp, [i32, f32, f32, u64]), "dl_sampler_free": (None, [vp]),
        "dl_sampler_set_temp": (None, [vp, f32]), "dl_sampler_set_topp": (None, [vp, f32]),
        "dl_sampler_set_seed": (None, [vp, u64]), "dl_sampler_sample": (i32, [vp, vp]), "dl_sampler_coin": (C.c_float, [vp]),
        "dl_rng_u32": (C.c_uint, [C.POINTER(u64)]), "dl_rng_f32": (f32, [C.POINTER(u64)]),
        "dl_softmax": (None, [vp, i32]),
        "dl_chat_detect": (i32, [cp]),
        "dl_chat_render": (i32, [i32, cp, i32, C.POINTER(cp), C.POINTER(cp), i32, cp, i32, ip]),
        "dl_stop_new": (vp, [i32, ip, C.POINTER(cp), i32, i32]), "dl_stop_free": (None, [vp]),
        "dl_stop_reset": (None, [vp]), "dl_stop_append": (i32, [vp, i32, cp, i32]),
        "dl_stop_delta": (vp, [vp, ip]),
        "dl_sched_new": (vp, [i32, i32, i32, i32]), "dl_sched_free": (None, [vp]),
        "dl_sched_set_pages": (None, [vp, i32, i32]), "dl_srv_set_pages": (None, [vp, i32, i32]),
        "dl_sched_submit": (i32, [vp, ll, ip, i32, i32]), "dl_sched_cancel": (i32, [vp, ll]),
        "dl_sched_wait": (i32, [vp, i32]),
        "dl_sched_next_batch": (i32, [vp, C.POINTER(ll), ip, ip, ip, i32, ip]),
        "dl_sched_sampling_rows": (i32, [vp, C.POINTER(ll), i32]),
        "dl_sched_commit": (i32, [vp, i32, ip, ip, ip]),
        "dl_sched_pop_released": (i32, [vp, ip, i32]), "dl_sched_forget": (i32, [vp, ll]),
        "dl_sched_stats": (None, [vp, ip, ip, ip]), "dl_sched_state": (i32, [vp, ll]),
        # native API server (csrc/host/server.cpp)
        "dl_srv_new": (vp, [vp, cp, i32, cp, i32, i32, i32, i32, i32, f32, f32, u64, i32, i32]),
        "dl_srv_listen": (i32, [vp, cp, i32]), "dl_srv_shutdown": (None, [vp]), "dl_srv_free": (None, [vp]),
        "dl_srv_wait": (i32, [vp, i32]),
        "dl_srv_next_batch": (i32, [vp, vp, vp, vp, i32, ip, ip, vp, vp, vp, i32, ip, vp, i32, ip]),
        "dl_srv_sample_host": (i32, [vp, vp, ll, vp]), "dl_srv_commit": (i32, [vp, i32, vp]),
        "dl_srv_fail_batch": (i32, [vp, cp]), "dl_srv_port": (i32, [vp]),
        "dl_srv_stats": (None, [vp, C.POINTER(ll)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    # The per-token helpers run for a microsecond; releasing the GIL around each of them (what CDLL does)
    # makes the engine thread queue behind the HTTP handler threads five times per generated token
    # (measured: ~63 us per token with 32 streaming clients, ~12 us of it actual work).  A second handle
    # of the same library that keeps the GIL serves those calls.
    global _host_fast
    F = C.PyDLL(HOST_LIB_PATH)
    for name in _FAST_CALLS:
        res, args = sig[name]
        fn = getattr(F, name)
        fn.restype = res
        fn.argtypes = args
    _host_fast = F
    _host = L
    return L


_FAST_CALLS = ("dl_dec_push", "dl_dec_reset", "dl_stop_append", "dl_stop_delta", "dl_stop_reset", "dl_sampler_coin",
               "dl_tok_is_eos")
_host_fast = None


def host_fast():
    """Handle for the microsecond-scale per-token calls (keeps the GIL, see host_lib)."""
    if _host_fast is None:
        host_lib()
    return _host_fast


# --------------------------------------------------------------------------------------------------
class Tokenizer:
    """`.t` tokenizer (reference: class Tokenizer, src/tokenizer.hpp:34-67)."""

    def __init__(self, path: str):
        L = host_lib()
        err = C.create_string_buffer(256)
        self._h = L.dl_tok_load(path.encode(), err, 256)
        if not self._h:
            raise RuntimeError(err.value.decode() or "Failed to open tokenizer file")
        self.vocab_size = L.dl_tok_vocab_size(self._h)
        self.bos_id = L.dl_tok_bos(self._h)
        self.eos_token_ids = [L.dl_tok_eos(self._h, i) for i in range(L.dl_tok_n_eos(self._h))]
        self.regular_vocab_size = L.dl_tok_n_regular(self._h)
        self.special_vocab_size = self.vocab_size - self.regular_vocab_size
        t = L.dl_tok_chat_template(self._h)
        self.chat_template: Optional[str] = t.decode("utf-8", "replace") if t else None
        self._dec = L.dl_dec_new(self._h)
        self._buf = C.create_string_buffer(65536)

    def __del__(self):
        try:
            L = host_lib()
            if getattr(self, "_dec", None):
                L.dl_dec_free(self._dec)
            if getattr(self, "_h", None):
                L.dl_tok_free(self._h)
        except Exception:
            pass

    def piece(self, token: int) -> bytes:
        n = C.c_int(0)
        p = host_lib().dl_tok_piece(self._h, token, C.byref(n))
        return C.string_at(p, n.value) if p else b""

    def is_eos(self, token: int) -> bool:
        return token in self.eos_token_ids

    def encode(self, text, add_bos: bool = True, add_special_tokens: bool = True) -> List[int]:
        raw = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        cap = len(raw) + 2
        out = (C.c_int * cap)()
        n = host_lib().dl_tok_encode(self._h, raw, len(raw), int(add_bos), int(add_special_tokens), out, cap)
        if n == -2:
            raise ValueError("text contains bytes that match no token")
        if n < 0:
            raise RuntimeError("token buffer too small")
        return list(out[:n])

    def new_decoder(self) -> "StreamDecoder":
        return StreamDecoder(self)

    # stateful single-stream interface, like the reference's decode()/resetDecoder()
    def reset_decoder(self):
        host_lib().dl_dec_reset(self._dec)

    def decode(self, token: int) -> Optional[str]:
        n = host_lib().dl_dec_push(self._dec, token, self._buf, len(self._buf))
        if n <= 0:
            return None
        return C.string_at(self._buf, n).decode("utf-8", "replace")     # (.raw would copy the whole 64 KB buffer per token)

    def decode_all(self, tokens: Sequence[int]) -> str:
        d = self.new_decoder()
        return "".join(filter(None, (d.push(t) for t in tokens)))


class StreamDecoder:
    """Per-request streaming detokenizer with UTF-8 reassembly (reference: Tokenizer::decode +
    detokUtf8, src/tokenizer.cpp:214-299) -- one instance per user, unlike the reference's single
    shared buffer."""

    def __init__(self, tok: Tokenizer):
        self._tok = tok
        self._h = host_lib().dl_dec_new(tok._h)
        self._buf = C.create_string_buffer(65536)

    def __del__(self):
        try:
            host_lib().dl_dec_free(self._h)
        except Exception:
            pass

    def reset(self):
        host_lib().dl_dec_reset(self._h)

    def push_bytes(self, token: int) -> bytes:
        n = host_fast().dl_dec_push(self._h, token, self._buf, len(self._buf))
        return C.string_at(self._buf, n)          # (.raw would copy the whole 64 KB buffer per token)

    def push(self, token: int) -> Optional[str]:
        b = self.push_bytes(token)
        return b.decode("utf-8", "replace") if b else None


# --------------------------------------------------------------------------------------------------
class Sampler:
    """temperature -> softmax -> top-p / multinomial, argmax at temperature 0, xorshift* RNG
    (reference: class Sampler, src/tokenizer.cpp:382-510).  One instance per request."""

    def __init__(self, vocab_size: int, temperature: float, topp: float, seed: int):
        self.vocab_size = vocab_size
        self._h = host_lib().dl_sampler_new(vocab_size, temperature, topp, seed & 0xFFFFFFFFFFFFFFFF)

    def __del__(self):
        try:
            host_lib().dl_sampler_free(self._h)
        except Exception:
            pass

    def set_temp(self, t: float):
        host_lib().dl_sampler_set_temp(self._h, t)

    def set_topp(self, p: float):
        host_lib().dl_sampler_set_topp(self._h, p)

    def set_seed(self, seed: int):
        host_lib().dl_sampler_set_seed(self._h, seed & 0xFFFFFFFFFFFFFFFF)

    def coin(self) -> float:
        """Next uniform [0, 1) draw of this request's RNG stream (fed to the device sampler)."""
        return float(host_fast().dl_sampler_coin(self._h))

    def sample(self, logits: np.ndarray) -> int:
        """``logits``: contiguous float32 [vocab]; mutated in place like the reference."""
        assert logits.dtype == np.float32 and logits.flags["C_CONTIGUOUS"] and logits.size >= self.vocab_size
        return int(host_lib().dl_sampler_sample(self._h, logits.ctypes.data))


def rng_u32(state: int) -> Tuple[int, int]:
    s = C.c_ulonglong(state)
    v = host_lib().dl_rng_u32(C.byref(s))
    return int(v), int(s.value)


def softmax(x: np.ndarray) -> np.ndarray:
    y = np.ascontiguousarray(x, dtype=np.float32).copy()
    host_lib().dl_softmax(y.ctypes.data, y.size)
    return y


# --------------------------------------------------------------------------------------------------
CHAT_UNKNOWN, CHAT_LLAMA2, CHAT_LLAMA3, CHAT_DEEPSEEK3 = 0, 1, 2, 3
CHAT_NAMES = {"llama2": CHAT_LLAMA2, "llama3": CHAT_LLAMA3, "deepSeek3": CHAT_DEEPSEEK3}


class ChatTemplateGenerator:
    """(reference: ChatTemplateGenerator, src/tokenizer.cpp:538-612)"""

    def __init__(self, template_type: int, chat_template: Optional[str], eos: str):
        if template_type == CHAT_UNKNOWN:
            if chat_template is None:
                raise ValueError("The tokenizer does not include chat template")
            template_type = host_lib().dl_chat_detect(chat_template.encode("utf-8"))
            if template_type == CHAT_UNKNOWN:
                raise ValueError("Not supported chat template")
        self.type = template_type
        self.eos = eos

    def generate(self, items: Sequence[Tuple[str, str]], append_generation_prompt: bool = True):
        """items: (role, message) pairs -> (content, public_prompt or None)."""
        n = len(items)
        roles = (C.c_char_p * n)(*[r.encode("utf-8") for r, _ in items])
        msgs = (C.c_char_p * n)(*[m.encode("utf-8") for _, m in items])
        pl = C.c_int(0)
        cap = sum(len(r) + len(m.encode("utf-8")) for r, m in items) + 128 * (n + 2) + len(self.eos) * (n + 1)
        buf = C.create_string_buffer(cap)
        ln = host_lib().dl_chat_render(self.type, self.eos.encode("utf-8"), n, roles, msgs,
                                       int(append_generation_prompt), buf, cap, C.byref(pl))
        assert ln < cap
        content = buf.raw[:ln].decode("utf-8")
        public = content.encode("utf-8")[ln - pl.value:].decode("utf-8") if pl.value > 0 else None
        return content, public


EOS_MAYBE, EOS_YES, EOS_NO = 0, 1, 2


class EosDetector:
    """Incremental stop-token / stop-string matcher (reference: EosDetector, src/tokenizer.cpp:614-699)."""

    def __init__(self, tokens: Sequence[int], pieces: Sequence[str], padding_left: int, padding_right: int):
        n = len(tokens)
        self._tokens = (C.c_int * n)(*tokens)
        self._pieces = (C.c_char_p * n)(*[p.encode("utf-8") if isinstance(p, str) else p for p in pieces])
        self._h = host_lib().dl_stop_new(n, self._tokens, self._pieces, padding_left, padding_right)

    def __del__(self):
        try:
            host_lib().dl_stop_free(self._h)
        except Exception:
            pass

    def append(self, token_id: int, piece) -> int:
        if piece is None:
            return host_fast().dl_stop_append(self._h, token_id, None, 0)
        b = piece.encode("utf-8") if isinstance(piece, str) else piece
        return host_fast().dl_stop_append(self._h, token_id, b, len(b))

    def get_delta(self) -> Optional[str]:
        n = C.c_int(0)
        p = host_fast().dl_stop_delta(self._h, C.byref(n))
        if not p or n.value == 0:
            return None
        return C.string_at(p, n.value).decode("utf-8", "replace")

    def reset(self):
        host_fast().dl_stop_reset(self._h)


# --------------------------------------------------------------------------------------------------
class Scheduler:
    """Continuous-batching scheduler (csrc/host/scheduler.cpp)."""

    def __init__(self, max_rows: int, n_slots: int, prefill_chunk: int, max_seq_len: int):
        self.max_rows = max_rows
        self._h = host_lib().dl_sched_new(max_rows, n_slots, prefill_chunk, max_seq_len)
        R = max_rows
        self._ids = (C.c_longlong * R)()
        self._tok = (C.c_int * R)()
        self._pos = (C.c_int * R)()
        self._slot = (C.c_int * R)()

    def set_pages(self, n_pages: int, page_size: int = 32):
        """KV page pool of the engine: a request is admitted only if the pages of its whole life (prompt + max_new
        tokens) fit beside the reservations of the running requests, so a step can never run out of pages."""
        host_lib().dl_sched_set_pages(self._h, n_pages, page_size)

    def __del__(self):
        try:
            host_lib().dl_sched_free(self._h)
        except Exception:
            pass

    def submit(self, req_id: int, tokens: Sequence[int], max_new_tokens: int) -> int:
        arr = (C.c_int * len(tokens))(*tokens)
        return host_lib().dl_sched_submit(self._h, req_id, arr, len(tokens), max_new_tokens)

    def cancel(self, req_id: int) -> int:
        return host_lib().dl_sched_cancel(self._h, req_id)

    def wait(self, timeout_ms: int) -> bool:
        return bool(host_lib().dl_sched_wait(self._h, timeout_ms))

    def next_batch(self):
        """-> (req_ids, tokens, positions, slots, logits_from, sampling_req_ids)"""
        lf = C.c_int(0)
        n = host_lib().dl_sched_next_batch(self._h, self._ids, self._tok, self._pos, self._slot, self.max_rows,
                                           C.byref(lf))
        ns = host_lib().dl_sched_sampling_rows(self._h, self._ids_s(), self.max_rows)
        return (list(self._ids[:n]), list(self._tok[:n]), list(self._pos[:n]), list(self._slot[:n]), lf.value,
                list(self._sids[:ns]))

    def _ids_s(self):
        if not hasattr(self, "_sids"):
            self._sids = (C.c_longlong * self.max_rows)()
        return self._sids

    def commit(self, sampled: Sequence[int], stop: Sequence[bool]) -> List[bool]:
        n = len(sampled)
        a = (C.c_int * n)(*sampled)
        b = (C.c_int * n)(*[1 if s else 0 for s in stop])
        fin = (C.c_int * n)()
        rc = host_lib().dl_sched_commit(self._h, n, a, b, fin)
        if rc != 0:
            raise RuntimeError("commit does not match the last batch")
        return [bool(x) for x in fin]

    def pop_released(self) -> List[int]:
        buf = (C.c_int * 256)()
        n = host_lib().dl_sched_pop_released(self._h, buf, 256)
        return list(buf[:n])

    def forget(self, req_id: int):
        host_lib().dl_sched_forget(self._h, req_id)

    def stats(self):
        a, b, c = C.c_int(0), C.c_int(0), C.c_int(0)
        host_lib().dl_sched_stats(self._h, C.byref(a), C.byref(b), C.byref(c))
        return {"waiting": a.value, "running": b.value, "free_slots": c.value}
