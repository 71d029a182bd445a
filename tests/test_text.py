"""Host text runtime (C++ via ctypes): tokenizer, streaming decoder, sampler, chat templates, stop
detector.  Goldens follow the reference's src/tokenizer-test.cpp:122-303 and src/nn/nn-cpu-ops-test.cpp:187-199."""
import numpy as np
import pytest

from dllama_b200.formats import tfile
from dllama_b200.text import native as T

EOS_ID = 10000


@pytest.fixture(scope="module")
def tok(tmp_path_factory):
    p = str(tmp_path_factory.mktemp("tok") / "syn.t")
    tfile.write_synthetic_tokenizer(p, 1024)
    return T.Tokenizer(p)


def test_tfile_roundtrip(tmp_path):
    p = str(tmp_path / "a.t")
    bos, eos = tfile.write_synthetic_tokenizer(p, 600)
    d = tfile.read_tokenizer(p)
    assert d["vocab_size"] == 600 and d["bos_id"] == bos and d["eos_tokens"] == eos
    assert d["tokens"][65] == b"A" and d["tokens"][bos] == b"<|begin_of_text|>"
    t = T.Tokenizer(p)
    assert t.vocab_size == 600 and t.bos_id == bos and t.eos_token_ids == eos
    assert t.regular_vocab_size == bos and t.special_vocab_size == 600 - bos
    assert t.piece(65) == b"A"
    assert "<|start_header_id|>" in t.chat_template


def test_bad_tokenizer_file(tmp_path):
    p = tmp_path / "bad.t"
    p.write_bytes(b"\x00" * 64)
    with pytest.raises(RuntimeError):
        T.Tokenizer(str(p))


def test_encode_merges_by_score_and_handles_special(tok):
    ids = tok.encode("hello world", add_bos=True, add_special_tokens=True)
    assert ids[0] == tok.bos_id
    assert tok.decode_all(ids) == "hello world"
    # merges happened (fewer tokens than bytes) and every merged piece is a regular-vocab entry
    pieces = [tok.piece(i) for i in ids[1:]]
    assert len(pieces) < len("hello world") and b"".join(pieces) == b"hello world"
    assert pieces[0] == b"he"                   # best-scored (lowest id) pair merges first
    ids2 = tok.encode("<|eot_id|>hi<|eot_id|>", add_bos=False, add_special_tokens=True)
    assert ids2[0] == ids2[-1] == tok.bos_id + 4
    ids3 = tok.encode("<|eot_id|>", add_bos=False, add_special_tokens=False)
    assert all(i < tok.bos_id for i in ids3) and len(ids3) > 1


def _reference_encode(tok, text: bytes, add_bos=True):
    """Literal restatement of the reference algorithm (O(n^2) rescans) as an oracle."""
    vocab = {}
    for i in range(tok.regular_vocab_size):
        vocab.setdefault(tok.piece(i), i)
    scores = [-float(i) for i in range(tok.vocab_size)]
    toks = [tok.bos_id] if add_bos else []
    acc = b""
    for b in text:
        acc += bytes([b])
        if acc in vocab:
            toks.append(vocab[acc])
            acc = b""
    assert acc == b""
    while True:
        best, best_id, best_idx = -1e10, -1, -1
        for i in range(len(toks) - 1):
            m = tok.piece(toks[i]) + tok.piece(toks[i + 1])
            j = vocab.get(m, -1)
            if j != -1 and scores[j] > best:
                best, best_id, best_idx = scores[j], j, i
        if best_idx == -1:
            break
        toks[best_idx:best_idx + 2] = [best_id]
    return toks


@pytest.mark.parametrize("text", ["the theory of the thing", "aaaaabbbbb  hello\n\nworld!!??", "zażółć gęślą jaźń 😃",
                                  "in the inn an ant ate the tent", ""])
def test_encode_equals_reference_algorithm(tok, text):
    got = tok.encode(text, add_bos=True, add_special_tokens=False)
    assert got == _reference_encode(tok, text.encode("utf-8"))
    assert tok.decode_all(got) == text


def test_streaming_decoder_reassembles_utf8(tok):
    emoji = "😃".encode("utf-8")                 # 4 bytes -> 4 byte-level tokens
    d = tok.new_decoder()
    outs = [d.push(b) for b in emoji]           # byte tokens have id == byte value
    assert outs[:3] == [None, None, None] and outs[3] == "😃"
    # emoji interrupted by EOS: the pending bytes are flushed raw (reference: decode() on eos)
    d.reset()
    assert d.push(emoji[0]) is None
    assert d.push_bytes(tok.eos_token_ids[0]) == emoji[:1]
    # stream recovery: a broken sequence becomes U+FFFD and decoding continues
    d.reset()
    assert d.push(emoji[0]) is None
    out = d.push(ord("x"))
    assert out == "�x"
    assert d.push(tok.bos_id) is None


def test_sampler_argmax_and_determinism():
    rng = np.random.default_rng(0)
    logits = rng.standard_normal(1000).astype(np.float32)
    s0 = T.Sampler(1000, 0.0, 0.9, 123)
    assert s0.sample(logits.copy()) == int(np.argmax(logits))
    a = T.Sampler(1000, 0.8, 0.9, 12345)
    b = T.Sampler(1000, 0.8, 0.9, 12345)
    seq_a = [a.sample(logits.copy()) for _ in range(20)]
    seq_b = [b.sample(logits.copy()) for _ in range(20)]
    assert seq_a == seq_b and len(set(seq_a)) > 1
    c = T.Sampler(1000, 0.8, 0.9, 999)
    assert [c.sample(logits.copy()) for _ in range(20)] != seq_a


def test_sampler_topp_restricts_support_and_matches_distribution():
    logits = np.log(np.array([0.5, 0.3, 0.15, 0.04, 0.01], dtype=np.float32))
    s = T.Sampler(5, 1.0, 0.9, 7)
    draws = [s.sample(logits.copy()) for _ in range(4000)]
    assert set(draws) <= {0, 1, 2}              # nucleus at p=0.9 is {0.5, 0.3, 0.15}
    freq = np.bincount(draws, minlength=5) / len(draws)
    np.testing.assert_allclose(freq[:3], np.array([0.5, 0.3, 0.15]) / 0.95, atol=0.03)
    s2 = T.Sampler(5, 1.0, 1.0, 7)              # topp >= 1 -> plain multinomial
    draws2 = [s2.sample(logits.copy()) for _ in range(4000)]
    assert set(draws2) == {0, 1, 2, 3, 4}


def test_xorshift_rng_matches_definition():
    state = 88172645463325252
    for _ in range(5):
        s = state
        s ^= s >> 12
        s ^= (s << 25) & 0xFFFFFFFFFFFFFFFF
        s ^= s >> 27
        want = ((s * 0x2545F4914F6CDD1D) & 0xFFFFFFFFFFFFFFFF) >> 32
        got, state = T.rng_u32(state)
        assert got == want and state == s


def test_softmax_golden():
    # reference golden (src/nn/nn-cpu-ops-test.cpp:187-199)
    x = np.array([0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8], dtype=np.float32)
    want = np.exp(x) / np.exp(x).sum()
    np.testing.assert_allclose(T.softmax(x), want, rtol=1e-6)


def test_chat_template_detection_and_rendering():
    llama3 = tfile.LLAMA3_CHAT_TEMPLATE
    g = T.ChatTemplateGenerator(T.CHAT_UNKNOWN, llama3, "<eos>")
    assert g.type == T.CHAT_LLAMA3
    content, public = g.generate([("system", "be brief"), ("user", "hi")], True)
    assert content == ("<|start_header_id|>system<|end_header_id|>\n\nbe brief<eos>"
                       "<|start_header_id|>user<|end_header_id|>\n\nhi<eos>"
                       "<|start_header_id|>assistant<|end_header_id|>\n\n")
    assert public is None
    g2 = T.ChatTemplateGenerator(T.CHAT_UNKNOWN, "... [INST] ...", "</s>")
    assert g2.type == T.CHAT_LLAMA2
    c2, _ = g2.generate([("system", "S"), ("user", "U"), ("assistant", "A"), ("user", "U2")], True)
    assert c2 == "[INST] <<SYS>>\nS\n<</SYS>>\n\nU [/INST]</s>A</s>[INST] U2 [/INST]</s>"
    g3 = T.ChatTemplateGenerator(T.CHAT_UNKNOWN, "x <｜Assistant｜> y", "<eos>")
    assert g3.type == T.CHAT_DEEPSEEK3
    c3, public3 = g3.generate([("system", "S"), ("user", "U")], True)
    assert c3 == "S<｜User｜>U<｜Assistant｜><think>\n" and public3 == "<think>\n"
    with pytest.raises(ValueError):
        T.ChatTemplateGenerator(T.CHAT_UNKNOWN, "no known markers", "<eos>")
    with pytest.raises(ValueError):
        T.ChatTemplateGenerator(T.CHAT_UNKNOWN, None, "<eos>")


def test_eos_detector_with_padding():
    d = T.EosDetector([EOS_ID, EOS_ID + 1], ["<eos>", "<stop>"], 1, 1)
    assert d.append(1, "<") == T.EOS_MAYBE
    assert d.append(2, "eo") == T.EOS_MAYBE
    assert d.append(3, "s>") == T.EOS_YES
    assert d.get_delta() is None
    d.reset()
    assert d.append(1, "<") == T.EOS_MAYBE
    assert d.append(2, "stop") == T.EOS_MAYBE
    assert d.append(3, "> ") == T.EOS_YES
    assert d.get_delta() is None
    d.reset()
    assert d.append(1, " ") == T.EOS_NO
    assert d.get_delta() == " "
    d.reset()
    assert d.append(1, "!<") == T.EOS_MAYBE
    assert d.append(2, "eos") == T.EOS_MAYBE
    assert d.append(3, "> ") == T.EOS_YES
    assert d.get_delta() == "!"
    d.reset()
    assert d.append(1, "<eo") == T.EOS_MAYBE
    assert d.append(2, "s>XY") == T.EOS_NO
    assert d.get_delta() == "<eos>XY"
    d.reset()
    assert d.append(1, "<eo") == T.EOS_MAYBE
    assert d.append(EOS_ID, None) == T.EOS_YES
    assert d.get_delta() == "<eo"
    d.reset()
    assert d.append(EOS_ID, None) == T.EOS_YES
    assert d.get_delta() is None
    d.reset()
    assert d.append(1, "x") == T.EOS_NO
    assert d.get_delta() == "x"
    d.reset()
    assert d.append(2, None) == T.EOS_NO
    assert d.get_delta() is None


def test_eos_detector_long_and_no_padding():
    d = T.EosDetector([EOS_ID], ["|end|"], 5, 5)
    assert d.append(1, "lipsum") == T.EOS_NO and d.get_delta() == "lipsum"
    d.reset()
    assert d.append(1, "lorem") == T.EOS_NO and d.get_delta() == "lorem"
    d.reset()
    assert d.append(1, "lorem|") == T.EOS_MAYBE
    assert d.append(2, "enQ") == T.EOS_NO and d.get_delta() == "lorem|enQ"
    n = T.EosDetector([EOS_ID], ["<eos>"], 0, 0)
    assert n.append(1, "<") == T.EOS_MAYBE and n.append(2, "eo") == T.EOS_MAYBE and n.append(3, "s>") == T.EOS_YES
    assert n.get_delta() is None
    n.reset()
    assert n.append(1, " <") == T.EOS_NO and n.get_delta() == " <"
    n.reset()
    assert n.append(1, "<eos") == T.EOS_MAYBE
    assert n.append(2, "> ") == T.EOS_NO and n.get_delta() == "<eos> "
    n.reset()
    assert n.append(EOS_ID, None) == T.EOS_YES and n.get_delta() is None
    n.reset()
    assert n.append(EOS_ID, "😃") == T.EOS_YES and n.get_delta() == "😃"


def test_encode_decode_roundtrip_property(tmp_path):
    """Any UTF-8 text survives encode -> streaming decode (token by token, multi-byte characters split across
    tokens included), and the native server's tokenizer path (same C++ encoder) agrees on the token count."""
    pytest.importorskip("hypothesis")
    from hypothesis import given, settings, strategies as st
    p = str(tmp_path / "prop.t")
    tfile.write_synthetic_tokenizer(p, 2048)
    tok = T.Tokenizer(p)

    @settings(max_examples=200, deadline=None)
    @given(st.text(max_size=60))
    def check(text):
        if "\x00" in text:
            text = text.replace("\x00", " ")
        ids = tok.encode(text, add_bos=True, add_special_tokens=False)
        assert ids[0] == tok.bos_id and all(0 <= i < tok.vocab_size for i in ids)
        assert tok.decode_all(ids) == text
        d = tok.new_decoder()
        pieces = [d.push(i) for i in ids]
        assert "".join(x for x in pieces if x) == text

    check()
