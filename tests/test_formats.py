"""File formats, codecs and converters (CPU)."""
import json
import os
import struct

import numpy as np
import pytest
import torch

from dllama_b200.formats import mfile, tfile
from dllama_b200.ops import quant

# hex golden of the reference's Q40 writer (converter/writer-test.py:13): torch.manual_seed(1); randn(32,16)
Q40_GOLDEN = ('7e346345a692b89665b2c5790537876e598aaa366d988876a898b8d788a98868ce660c66f6b3a88cba5ce9a871987ba9cc5bcaaa'
              '760c1eb556a4455b747b6b9504968828ef2a8d7c1db5c6be3764799e66db6d8e76463126a30e4333cad7a4f645947c6cf97f9de0'
              '86d468c8d535a6ba7dc799d3d0c657bab6799468cad8bb349eb7d7635c7c798998696bb38e4085a9eb34444ba96a7f8ba7b2b42d'
              '746a96cf9660aeb4499d8708ad5c7b9a7558947645f3bbb6b0346a656887ad9a86059baac5c596ab781c703569bb8a4356a4bd58'
              'cb78736ba09759bb0e34a6274e827b957d7a67dfa86846955660d234b6d9d78a378094a8a8708a7a774ae92f8a36b8c999a9b77a'
              '7d958a69747c807963941235379886d69a7a8767b3a6a4ac71999760')


def test_q40_writer_hex_golden():
    torch.manual_seed(1)
    t = torch.randn(32, 16)
    assert quant.quantize_q40(t.numpy()).tobytes().hex() == Q40_GOLDEN


def test_q40_q80_roundtrip_tolerances():
    rng = np.random.default_rng(0)
    x = rng.standard_normal(4096).astype(np.float32)
    # the reference accepts 0.13 (Q40) and 0.01 (Q80) absolute error on N(0,1) data (nn-cpu-ops-test.cpp)
    assert np.abs(quant.dequantize_q40(quant.quantize_q40(x)) - x).max() < 0.35
    assert np.abs(quant.dequantize_q40(quant.quantize_q40(x)) - x).mean() < 0.13
    assert np.abs(quant.dequantize_q80(quant.quantize_q80(x)) - x).max() < 0.02
    z = np.zeros(64, dtype=np.float32)
    assert (quant.dequantize_q80(quant.quantize_q80(z)) == 0).all() and (quant.dequantize_q40(quant.quantize_q40(z)) == 0).all()
    # half-to-even rounding of Q80 (x86 reference + python writer)
    blk = np.zeros(32, dtype=np.float32)
    blk[0], blk[1], blk[2] = 127.0, 0.5, 1.5
    q = quant.quantize_q80(blk)[0, 2:].view(np.int8)
    assert q[0] == 127 and q[1] == 0 and q[2] == 2
    assert quant.tensor_bytes(quant.Q40, 64) == 36 and quant.tensor_bytes(quant.Q80, 64) == 68


def test_model_file_roundtrip_and_errors(tmp_path):
    p = str(tmp_path / "m.m")
    h = mfile.write_random_model(p, "tiny", quant.Q40, exact=True, seed=2)
    assert (h.dim, h.n_layers, h.n_heads, h.n_kv_heads, h.vocab_size) == (256, 2, 8, 4, 1024)
    assert h.head_size == 32 and h.kv_dim == 128 and h.weight_type == quant.Q40 and h.header_size == 8 + 8 * 14
    mf = mfile.ModelFile(p, max_seq_len=100)
    assert mf.header.seq_len == 100 and mf.header.orig_seq_len == 512
    assert [e.name for e in mf.entries[:4]] == ["embedding", "wq", "wk", "wv"]
    assert mf.entries[-1].name == "wcls" and mf.entries[-1].offset + mf.entries[-1].nbytes == os.path.getsize(p)
    wq = mf.f32("wq", 0)
    assert wq.shape == (256, 256) and np.isfinite(wq).all()
    np.testing.assert_array_equal(quant.dequantize_q40(mf.raw_rows("wq", 0, 10, 2)).reshape(2, 256), wq[10:12])
    mf.close()
    with open(p, "r+b") as f:                       # truncated file
        f.truncate(os.path.getsize(p) - 7)
    with pytest.raises(ValueError, match="Missing bytes"):
        mfile.ModelFile(p)
    bad = tmp_path / "bad.m"
    bad.write_bytes(struct.pack("<ii", 0xABCD00, 8))
    with pytest.raises(ValueError, match="Old model format"):
        mfile.read_header(str(bad))
    bad.write_bytes(struct.pack("<iiii", mfile.MAGIC, 16, 99, 1))
    with pytest.raises(ValueError, match="Unsupported header key"):
        mfile.read_header(str(bad))
    bad.write_bytes(struct.pack("<iiii", mfile.MAGIC, 16, 2, 64))
    with pytest.raises(ValueError, match="weight type"):
        mfile.read_header(str(bad))


def test_llama31_rope_goldens():
    """RoPE table goldens of the reference's Vulkan op test (src/nn/nn-vulkan-test.cpp:467-482):
    theta 5e5, scaling 32/1/4/8192, head size 64."""
    from dllama_b200.models.oracle import apply_rope, rope_frequencies
    h = mfile.LlmHeader(dim=2048, n_heads=32, n_kv_heads=8, rope_theta=500000.0, rope_type=mfile.ROPE_LLAMA3_1,
                        rope_scaling_factor=32.0, rope_scaling_low_freq_factor=1.0,
                        rope_scaling_high_freq_factor=4.0, rope_scaling_orig_max_seq_len=8192)
    f = rope_frequencies(h)
    assert f.shape[0] == 32 and abs(float(f[0]) - 1.0) < 1e-7
    # high-frequency dims are untouched, the lowest ones are divided by the factor
    base = 500000.0 ** (-(2 * 31) / 64)
    assert abs(float(f[31]) - base / 32.0) / (base / 32.0) < 1e-4
    x = torch.ones(1, 64)
    y = apply_rope(x, torch.tensor([6]), f, 64)
    assert abs(float(y[0, 0]) - (np.cos(6.0) - np.sin(6.0))) < 1e-5
    assert abs(float(y[0, 1]) - (np.sin(6.0) + np.cos(6.0))) < 1e-5
    n0 = torch.linalg.norm(x)
    assert abs(float(torch.linalg.norm(y) - n0)) < 1e-4           # rotation preserves the norm


def test_hf_converter_matches_transformers(tmp_path):
    """HF safetensors -> .m (q/k row permutation, header mapping) checked end to end against
    transformers' own Llama implementation."""
    transformers = pytest.importorskip("transformers")
    from dllama_b200.converter import convert_hf
    from dllama_b200.models.oracle import OracleLlama
    cfg = transformers.LlamaConfig(hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=4,
                                   num_key_value_heads=2, vocab_size=320, max_position_embeddings=256,
                                   rope_theta=10000.0, rms_norm_eps=1e-5, hidden_act="silu", tie_word_embeddings=False)
    torch.manual_seed(0)
    model = transformers.LlamaForCausalLM(cfg).eval()
    src = tmp_path / "hf"
    model.save_pretrained(str(src), safe_serialization=True)
    out = str(tmp_path / "model.m")
    convert_hf.convert(str(src), quant.F32, out, log=lambda *_: None)
    mf = mfile.ModelFile(out)
    assert (mf.header.dim, mf.header.hidden_dim, mf.header.n_kv_heads, mf.header.vocab_size) == (128, 256, 2, 320)
    oracle = OracleLlama(mf, sync_q80=False, max_seq_len=64)
    ids = [1, 50, 7, 300, 12, 99]
    got = oracle.forward(ids, 0).numpy()
    with torch.no_grad():
        want = model(torch.tensor([ids])).logits[0].numpy()
    np.testing.assert_allclose(got, want, rtol=2e-3, atol=2e-4)
    # tied embeddings: lm_head falls back to embed_tokens
    sd = {k: v for k, v in model.state_dict().items() if k != "lm_head.weight"}
    from safetensors.torch import save_file
    tied = tmp_path / "tied"
    tied.mkdir()
    save_file({k: v.contiguous() for k, v in sd.items()}, str(tied / "model.safetensors"))
    (tied / "config.json").write_text((src / "config.json").read_text())
    out2 = str(tmp_path / "tied.m")
    convert_hf.convert(str(tied), quant.Q40, out2, log=lambda *_: None)
    mf2 = mfile.ModelFile(out2)
    emb = mf2.f32("embedding")
    assert np.abs(mf2.f32("wcls") - emb).max() < np.abs(emb).max() / 8 + 1e-6      # Q40 of the same matrix


def test_meta_converter(tmp_path):
    from dllama_b200.converter import convert_llama
    dim, hid, L, nh, vocab = 64, 128, 1, 4, 96
    torch.manual_seed(1)
    full = {"tok_embeddings.weight": torch.randn(vocab, dim), "norm.weight": torch.randn(dim),
            "output.weight": torch.randn(vocab, dim)}
    for l in range(L):
        p = f"layers.{l}."
        full.update({p + "attention.wq.weight": torch.randn(dim, dim), p + "attention.wk.weight": torch.randn(dim, dim),
                     p + "attention.wv.weight": torch.randn(dim, dim), p + "attention.wo.weight": torch.randn(dim, dim),
                     p + "feed_forward.w1.weight": torch.randn(hid, dim), p + "feed_forward.w2.weight": torch.randn(dim, hid),
                     p + "feed_forward.w3.weight": torch.randn(hid, dim), p + "attention_norm.weight": torch.randn(dim),
                     p + "ffn_norm.weight": torch.randn(dim)})
    src = tmp_path / "meta"
    src.mkdir()
    for s in range(2):                                   # two model-parallel shards
        shard = {}
        for k, v in full.items():
            if v.dim() == 1:
                shard[k] = v
            elif k == "tok_embeddings.weight" or k.endswith("wo.weight") or k.endswith("w2.weight"):
                shard[k] = v.chunk(2, dim=1)[s].clone()
            else:
                shard[k] = v.chunk(2, dim=0)[s].clone()
        torch.save(shard, str(src / f"consolidated.{s:02d}.pth"))
    (src / "params.json").write_text(json.dumps({"dim": dim, "n_layers": L, "n_heads": nh, "vocab_size": vocab,
                                                 "max_seq_len": 128, "rope_theta": 10000.0}))
    out = str(tmp_path / "meta.m")
    convert_llama.convert(str(src), out, quant.F32, log=lambda *_: None)
    mf = mfile.ModelFile(out)
    assert mf.header.hidden_dim == hid and mf.header.n_kv_heads == nh
    np.testing.assert_array_equal(mf.f32("embedding"), full["tok_embeddings.weight"].numpy())
    np.testing.assert_array_equal(mf.f32("w2", 0), full["layers.0.feed_forward.w2.weight"].numpy())
    np.testing.assert_array_equal(mf.f32("wq", 0), full["layers.0.attention.wq.weight"].numpy())
    np.testing.assert_array_equal(mf.f32("wcls"), full["output.weight"].numpy())


def test_tokenizer_converters(tmp_path, monkeypatch):
    import base64
    from dllama_b200.converter import convert_tokenizer as ct
    from dllama_b200.text.native import Tokenizer
    monkeypatch.chdir(tmp_path)
    # llama3-style tiktoken rank file
    ranks = [bytes([b]) for b in range(256)] + [b"he", b"ll", b"hello"]
    (tmp_path / "tokenizer.model").write_text("".join(f"{base64.b64encode(t).decode()} {i}\n" for i, t in enumerate(ranks)))
    out = ct.convert_llama3(str(tmp_path / "tokenizer.model"))
    d = tfile.read_tokenizer(out)
    assert d["vocab_size"] == len(ranks) + 256 and d["bos_id"] == 128000 and d["eos_tokens"] == [128001, 128009]
    assert d["tokens"][len(ranks)] == b"<|begin_of_text|>" and d["tokens"][len(ranks) + 9] == b"<|eot_id|>"
    # HF fast tokenizer (tokenizer.json) through the GPT-2 byte table
    btu = {v: k for k, v in ct.unicode_to_bytes().items()}
    enc = lambda b: "".join(btu[x] for x in b)
    vocab = {enc(bytes([b])): b for b in range(256)}
    for i, piece in enumerate((b" t", b" th", b" the")):           # BPE needs every intermediate merge
        vocab[enc(piece)] = 256 + i
    hf = tmp_path / "hf"
    hf.mkdir()
    (hf / "tokenizer.json").write_text(json.dumps({"model": {"vocab": vocab}, "added_tokens": [
        {"id": 259, "content": "<s>"}, {"id": 260, "content": "</s>"}]}), encoding="utf-8")
    (hf / "tokenizer_config.json").write_text(json.dumps({"tokenizer_class": "PreTrainedTokenizerFast", "bos_token": "<s>",
                                                           "eos_token": {"content": "</s>"}, "chat_template": "x [INST] y"}))
    out = ct.convert_hf(str(hf), "tst")
    t = Tokenizer(out)
    assert t.vocab_size == 261 and t.bos_id == 259 and t.eos_token_ids == [260] and t.piece(258) == b" the"
    assert t.piece(0x20) == b" " and t.chat_template == "x [INST] y"
    ids = t.encode("a the", add_bos=True)
    assert ids == [259, ord("a"), 258] and t.decode_all(ids) == "a the"


def test_required_memory_estimate_matches_file_sizes_and_fits_405b_on_8_gpus():
    """(reference: printNodeRequiredMemory) the per-node weight bytes of all nodes add up to the file's matrices, and
    the estimate answers the BASELINE question "does Llama-3.1-405B Q40 fit 8 x 180 GB"."""
    from dllama_b200.models.memory import B200_HBM_BYTES, required_memory
    for name in ("llama3.2-1b", "llama3.1-8b"):
        shape = mfile.MODEL_SHAPES[name]
        h = mfile.LlmHeader(dim=shape["dim"], hidden_dim=shape["hidden_dim"], n_layers=shape["n_layers"],
                            n_heads=shape["n_heads"], n_kv_heads=shape["n_kv_heads"], vocab_size=shape["vocab_size"],
                            seq_len=4096, weight_type=quant.Q40)
        h.header_size = 0
        file_matrices = mfile.expected_file_size(h) - h.vocab_size * h.dim * 4          # everything but the embedding
        for n in (1, 2, 8):
            m = required_memory(h, n, backend="cpu")
            replicated = (n - 1) * (2 * h.n_layers + 1) * h.dim * 4                       # norm weights live on every node
            assert n * m["weights"] == file_matrices + replicated
    s = mfile.MODEL_SHAPES["llama3.1-405b"]
    h = mfile.LlmHeader(dim=s["dim"], hidden_dim=s["hidden_dim"], n_layers=s["n_layers"], n_heads=s["n_heads"],
                        n_kv_heads=s["n_kv_heads"], vocab_size=s["vocab_size"], seq_len=8192, weight_type=quant.Q40)
    m8 = required_memory(h, 8, max_rows=256, n_slots=34, backend="cuda")
    assert 28e9 < m8["weights"] < 31e9 and m8["total"] < 0.5 * B200_HBM_BYTES             # ~30 GB of Q40T per rank
    assert required_memory(h, 1, backend="cuda")["total"] > B200_HBM_BYTES                # and it does not fit one GPU


def test_native_random_model_writer_produces_a_valid_equivalent_file(tmp_path):
    """csrc/host/model_writer.cpp (multi-threaded, for the 40-240 GB files): same layout, same size, same weight
    statistics as the NumPy generator, deterministic per seed, loadable by the engines."""
    a, b, c = str(tmp_path / "a.m"), str(tmp_path / "b.m"), str(tmp_path / "c.m")
    ha = mfile.write_random_model(a, "tiny-tp8", seed=7, native=True)
    mfile.write_random_model(b, "tiny-tp8", seed=7, native=True)
    mfile.write_random_model(c, "tiny-tp8", seed=7, native=False)
    assert os.path.getsize(a) == mfile.expected_file_size(ha) == os.path.getsize(c)
    assert open(a, "rb").read() == open(b, "rb").read()                       # same seed -> same bytes, whatever the thread timing
    ma, mc = mfile.ModelFile(a), mfile.ModelFile(c)
    assert ma.header.dim == mc.header.dim and ma.header.n_layers == mc.header.n_layers
    for name, layer in (("wq", 0), ("w2", 1), ("wcls", -1)):
        wa, wc = ma.f32(name, layer), mc.f32(name, layer)
        assert wa.shape == wc.shape and abs(wa.std() - wc.std()) < 0.05 * wc.std() and abs(wa.mean()) < 1e-3
    assert abs(ma.f32("embedding").std() - 1.0) < 0.02 and abs(ma.f32("rms_ffn", 1).mean() - 1.0) < 0.02
    ma.close()
    mc.close()
    from dllama_b200.models.native_cpu_engine import NativeCpuEngine
    eng = NativeCpuEngine(a, n_threads=2, max_rows=4, n_slots=1, max_seq_len=64)
    lg = eng.forward([1, 2, 3], [0, 1, 2], [0, 0, 0]).numpy()
    assert np.isfinite(lg).all() and lg.std() > 0
    eng.close()


def test_required_memory_counts_the_fp16_weight_copy():
    """Engines that can see >= 768-row batches keep an fp16 copy of the layer matrices (2 bytes per weight): the
    measured sizes were 13.96 GB for the 8B model on one GPU and 17.11 GB per rank for the 70B model on 8."""
    from dllama_b200.models.memory import required_memory
    for name, n, measured in (("llama3.1-8b", 1, 13958643712), ("llama3.1-8b", 2, 6979321856), ("llama3.3-70b", 8, 17112760320)):
        s = mfile.MODEL_SHAPES[name]
        h = mfile.LlmHeader(dim=s["dim"], hidden_dim=s["hidden_dim"], n_layers=s["n_layers"], n_heads=s["n_heads"],
                            n_kv_heads=s["n_kv_heads"], vocab_size=s["vocab_size"], seq_len=4096, weight_type=quant.Q40)
        assert required_memory(h, n, max_rows=4096, backend="cuda")["fp16_weight_copy"] == measured
        assert "fp16_weight_copy" not in required_memory(h, n, max_rows=256, backend="cuda")
