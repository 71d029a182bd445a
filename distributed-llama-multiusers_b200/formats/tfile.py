"""`.t` tokenizer file writer + pure-Python reader, and a synthetic tokenizer generator.

Layout (reference: converter/tokenizer-writer.py:3-55, src/tokenizer.cpp:65-135), little endian:

    int32 magic = 0x567124
    int32 headerSize = 8 + 8*nKv
    nKv x (int32 key, int32 value)   keys: version=0 (must be 1), vocab_size=1, max_token_length=2,
                                     bos_id=3, chat_template=7 (byte length), n_eos_tokens=9
    [chat template bytes] [nEos x int32] vocab x { float32 score; uint32 len; bytes }

The runtime treats ids < bos_id as the regular (mergeable) vocabulary and ids >= bos_id as special
tokens (src/tokenizer.cpp:137-139).
"""
from __future__ import annotations

import struct
from typing import List, Optional, Sequence, Tuple

MAGIC = 0x567124
KEYS = {"version": 0, "vocab_size": 1, "max_token_length": 2, "bos_id": 3, "chat_template": 7, "n_eos_tokens": 9}

LLAMA3_CHAT_TEMPLATE = (
    "{% set loop_messages = messages %}{% for message in loop_messages %}{% set content = "
    "'<|start_header_id|>' + message['role'] + '<|end_header_id|>\n\n'+ message['content'] | trim + "
    "'<|eot_id|>' %}{% if loop.index0 == 0 %}{% set content = bos_token + content %}{% endif %}{{ content }}"
    "{% endfor %}{% if add_generation_prompt %}{{ '<|start_header_id|>assistant<|end_header_id|>\n\n' }}"
    "{% endif %}")


def write_tokenizer(f, tokens: Sequence[bytes], scores: Sequence[float], chat_template: Optional[bytes],
                    bos_id: int, eos_tokens: Sequence[int]) -> None:
    assert len(tokens) == len(scores)
    params = {"bos_id": bos_id, "version": 1, "vocab_size": len(tokens),
              "max_token_length": max(len(t) for t in tokens)}
    if chat_template:
        params["chat_template"] = len(chat_template)
    params["n_eos_tokens"] = len(eos_tokens)
    data = b"".join(struct.pack("<ii", KEYS[k], v) for k, v in params.items())
    f.write(struct.pack("<ii", MAGIC, 8 + len(data)))
    f.write(data)
    if chat_template:
        f.write(chat_template)
    for e in eos_tokens:
        f.write(struct.pack("<i", e))
    for t, s in zip(tokens, scores):
        assert len(t) > 0
        f.write(struct.pack("<fI", s, len(t)))
        f.write(t)


def read_tokenizer(path: str) -> dict:
    """Pure-Python reader (tests / tooling; the runtime uses the C++ loader)."""
    with open(path, "rb") as f:
        magic, header_size = struct.unpack("<ii", f.read(8))
        if magic != MAGIC:
            raise ValueError("Invalid tokenizer file")
        kv = struct.unpack("<%di" % ((header_size - 8) // 4), f.read(header_size - 8))
        p = {kv[i]: kv[i + 1] for i in range(0, len(kv), 2)}
        if p.get(0) != 1:
            raise ValueError("Old tokenizer version, please regenerate your tokenizer")
        tmpl = f.read(p[7]) if p.get(7, 0) > 0 else None
        eos = [struct.unpack("<i", f.read(4))[0] for _ in range(p.get(9, 0))]
        tokens, scores = [], []
        for _ in range(p[1]):
            s, n = struct.unpack("<fI", f.read(8))
            tokens.append(f.read(n))
            scores.append(s)
    return {"vocab_size": p[1], "max_token_length": p[2], "bos_id": p[3], "chat_template": tmpl,
            "eos_tokens": eos, "tokens": tokens, "scores": scores}


def synthetic_vocab(vocab_size: int, n_special: int = 8) -> Tuple[List[bytes], List[float], int, List[int]]:
    """A byte-level BPE-shaped vocabulary for random-init models: 256 byte tokens, then merged
    pairs/words with descending scores, then special tokens (bos first, like Llama 3)."""
    assert vocab_size >= 256 + n_special + 2
    tokens: List[bytes] = [bytes([b]) for b in range(256)]
    # byte 0 cannot travel through C strings; keep it but it is never produced by text
    words = [b" the", b" a", b"he", b"in", b"er", b"th", b" t", b"an", b"on", b"re", b" s", b"at", b"en", b"ll",
             b"lo", b"hel", b"hello", b" w", b"or", b"wor", b"ld", b"world", b" world", b"ing", b" of", b" to",
             b" and", b"is", b" is", b"it", b"\n\n", b"  ", b"ou", b"yo", b" you", b"!!", b"??", b"es", b"ed"]
    n_regular = vocab_size - n_special
    i = 0
    seen = set(tokens)
    while len(tokens) < n_regular:
        if i < len(words):
            w = words[i]
        else:
            # deterministic filler: two-byte then three-byte lowercase combos
            j = i - len(words)
            a, b, c = j % 26, (j // 26) % 26, (j // 676) % 26
            w = bytes([97 + a, 97 + b]) if j < 676 else bytes([97 + a, 97 + b, 97 + c, 48 + (j // 17576) % 10])
        i += 1
        if w in seen:
            continue
        seen.add(w)
        tokens.append(w)
    scores = [-float(k) for k in range(len(tokens))]
    bos_id = len(tokens)
    specials = [b"<|begin_of_text|>", b"<|end_of_text|>", b"<|start_header_id|>", b"<|end_header_id|>",
                b"<|eot_id|>", b"<|pad|>", b"<|reserved_0|>", b"<|reserved_1|>"]
    while len(specials) < n_special:
        specials.append(b"<|reserved_%d|>" % len(specials))
    specials = specials[:n_special]
    tokens += specials
    scores += [-float(k) for k in range(bos_id, bos_id + n_special)]
    eos = [bos_id + 1, bos_id + 4] if n_special >= 5 else [bos_id + 1]
    return tokens, scores, bos_id, eos


def write_synthetic_tokenizer(path: str, vocab_size: int, chat_template: Optional[str] = LLAMA3_CHAT_TEMPLATE):
    tokens, scores, bos, eos = synthetic_vocab(vocab_size)
    with open(path, "wb") as f:
        write_tokenizer(f, tokens, scores, chat_template.encode("utf-8") if chat_template else None, bos, eos)
    return bos, eos
