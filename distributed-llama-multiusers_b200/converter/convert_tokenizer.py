"""Tokenizer converters -> `.t` files.

    python -m dllama_b200.converter.convert_tokenizer hf     <folder with tokenizer_config.json> <name>
    python -m dllama_b200.converter.convert_tokenizer llama3 <tokenizer.model (tiktoken ranks)>
    python -m dllama_b200.converter.convert_tokenizer llama2 <folder with tokenizer.model>

Behavioural parity with the reference's three scripts:
* hf (convert-tokenizer-hf.py:12-120): fast tokenizers are read from ``tokenizer.json`` through the
  inverse of GPT-2's byte<->unicode table, score = -id; ``LlamaTokenizer`` goes through sentencepiece
  ('▁' -> ' ', ``<0xNN>`` byte pieces); bos/eos come from the tokenizer or ``config.json``; the chat
  template is copied from ``tokenizer_config.json``.
* llama3 (convert-tokenizer-llama3.py:13-76): base64 rank file + the 256 special tokens, bos 128000,
  eos [128001, 128009], built-in Llama-3 chat template.
* llama2 (convert-tokenizer-llama2.py): sentencepiece model + the Llama-2 chat template.
"""
from __future__ import annotations

import base64
import json
import os
import sys
from typing import List, Optional, Tuple

from ..formats import tfile

LLAMA2_CHAT_TEMPLATE = (
    "{% if messages[0]['role'] == 'system' %}{% set loop_messages = messages[1:] %}{% set system_message = "
    "messages[0]['content'] %}{% else %}{% set loop_messages = messages %}{% set system_message = false %}"
    "{% endif %}{% for message in loop_messages %}{% if (message['role'] == 'user') != (loop.index0 % 2 == 0) %}"
    "{{ raise_exception('Conversation roles must alternate user/assistant/user/assistant/...') }}{% endif %}"
    "{% if loop.index0 == 0 and system_message != false %}{% set content = '<<SYS>>\\n' + system_message + "
    "'\\n<</SYS>>\\n\\n' + message['content'] %}{% else %}{% set content = message['content'] %}{% endif %}"
    "{% if message['role'] == 'user' %}{{ bos_token + '[INST] ' + content.strip() + ' [/INST]' }}"
    "{% elif message['role'] == 'assistant' %}{{ ' '  + content.strip() + ' ' + eos_token }}{% endif %}{% endfor %}")


def unicode_to_bytes() -> dict:
    """Inverse of GPT-2's bytes_to_unicode table."""
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(ord("¡"), ord("¬") + 1)) + list(range(ord("®"), ord("ÿ") + 1))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return {chr(c): b for c, b in zip(cs, bs)}


def tokens_from_tokenizer_json(path: str) -> Tuple[List[bytes], List[float]]:
    """vocab + added tokens of a `tokenizer.json`, ordered by id, decoded to raw bytes."""
    with open(path, encoding="utf-8") as f:
        tj = json.load(f)
    vocab = dict(tj["model"]["vocab"])
    for added in tj.get("added_tokens", []):
        vocab[added["content"]] = added["id"]
    size = max(vocab.values()) + 1
    by_id: List[Optional[str]] = [None] * size
    for s, i in vocab.items():
        by_id[i] = s
    utb = unicode_to_bytes()
    tokens, scores = [], []
    for i, s in enumerate(by_id):
        s = s if s is not None else f"<|unused_{i}|>"
        raw = bytearray()
        for ch in s:
            if ch in utb:
                raw.append(utb[ch])
            else:
                raw += ch.encode("utf-8")
        tokens.append(bytes(raw))
        scores.append(-float(i))
    return tokens, scores


def tokens_from_sentencepiece(model_path: str, byte_pieces: bool) -> Tuple[List[bytes], List[float], int, int]:
    from sentencepiece import SentencePieceProcessor
    sp = SentencePieceProcessor(model_file=model_path)
    tokens, scores = [], []
    for i in range(sp.vocab_size()):
        t = sp.id_to_piece(i).replace("▁", " ")
        if byte_pieces and len(t) == 6 and t.startswith("<0x") and t.endswith(">"):
            b = bytes(bytearray.fromhex(t[3:-1]))
        else:
            b = t.encode("utf-8")
        tokens.append(b)
        scores.append(sp.get_score(i))
    return tokens, scores, sp.bos_id(), sp.eos_id()


def convert_hf(folder: str, name: str) -> str:
    with open(os.path.join(folder, "tokenizer_config.json"), encoding="utf-8") as f:
        tc = json.load(f)
    cls = tc.get("tokenizer_class")
    bos_id = eos_ids = None
    if cls in ("PreTrainedTokenizerFast", "LlamaTokenizerFast", "PreTrainedTokenizer"):
        tokens, scores = tokens_from_tokenizer_json(os.path.join(folder, "tokenizer.json"))
        index = {t: i for i, t in enumerate(tokens)}

        def tid(v):
            if isinstance(v, dict):
                v = v.get("content")
            return index.get(v.encode("utf-8")) if isinstance(v, str) else None
        bos_id = tid(tc.get("bos_token"))
        e = tid(tc.get("eos_token"))
        eos_ids = [e] if e is not None else None
    elif cls == "LlamaTokenizer":
        tokens, scores, bos_id, e = tokens_from_sentencepiece(os.path.join(folder, "tokenizer.model"), True)
        eos_ids = [e]
    else:
        raise ValueError(f"Tokenizer {cls} is not supported")
    if bos_id is None or eos_ids is None:
        with open(os.path.join(folder, "config.json")) as f:
            cfg = json.load(f)
        if bos_id is None:
            bos_id = cfg["bos_token_id"]
        if eos_ids is None:
            e = cfg["eos_token_id"]
            eos_ids = e if isinstance(e, list) else [e]
    if bos_id is None or not eos_ids:
        raise ValueError("Cannot resolve bosId or eosIds")
    tmpl = tc.get("chat_template")
    out = f"dllama_tokenizer_{name}.t"
    with open(out, "wb") as f:
        tfile.write_tokenizer(f, tokens, scores, tmpl.encode("utf-8") if isinstance(tmpl, str) else None, bos_id, eos_ids)
    return out


def convert_llama3(model_path: str) -> str:
    n_special = 256
    special = ["<|begin_of_text|>", "<|end_of_text|>", "<|reserved_special_token_0|>", "<|reserved_special_token_1|>",
               "<|reserved_special_token_2|>", "<|reserved_special_token_3|>", "<|start_header_id|>",
               "<|end_header_id|>", "<|reserved_special_token_4|>", "<|eot_id|>"]
    special += [f"<|reserved_special_token_{i}|>" for i in range(5, n_special - 5)]
    tokens, scores = [], []
    with open(model_path) as f:
        for line in f:
            if not line.strip():
                continue
            b64, rank = line.split(" ")
            tokens.append(base64.b64decode(b64))
            scores.append(-float(rank))
    n = len(tokens)
    for i, s in enumerate(special):
        tokens.append(s.encode("utf-8"))
        scores.append(-float(n + i))
    out = "dllama_tokenizer_llama3.t"
    with open(out, "wb") as f:
        tfile.write_tokenizer(f, tokens, scores, tfile.LLAMA3_CHAT_TEMPLATE.encode("utf-8"), 128000, [128001, 128009])
    return out


def convert_llama2(folder: str) -> str:
    tokens, scores, bos, eos = tokens_from_sentencepiece(os.path.join(folder, "tokenizer.model"), False)
    out = "dllama_tokenizer_llama2.t"
    with open(out, "wb") as f:
        tfile.write_tokenizer(f, tokens, scores, LLAMA2_CHAT_TEMPLATE.encode("utf-8"), bos, [eos])
    return out


def main(argv=None) -> int:
    argv = sys.argv if argv is None else argv
    if len(argv) < 3 or argv[1] not in ("hf", "llama3", "llama2") or (argv[1] == "hf" and len(argv) < 4):
        print(__doc__)
        return 1
    out = convert_hf(argv[2], argv[3]) if argv[1] == "hf" else \
        convert_llama3(argv[2]) if argv[1] == "llama3" else convert_llama2(argv[2])
    print(f"✅ Created {out}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
