"""End-to-end engine vs the CPU oracle (which mirrors the reference's quantisation points)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _oracle(path, n_nodes=1):
    from dllama_b200.formats.mfile import ModelFile
    from dllama_b200.models.oracle import OracleLlama
    return OracleLlama(ModelFile(path), n_nodes=n_nodes, max_seq_len=512)


def _engine(path, **kw):
    from dllama_b200.models.llama import LlamaEngine
    kw.setdefault("max_seq_len", 512)
    return LlamaEngine(path, **kw)


def _rel_err(a, b):
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-9))


@pytest.mark.parametrize("use_graphs,use_pdl", [(False, False), (True, False), (True, True)])
def test_decode_logits_match_oracle(tiny_model_path, use_graphs, use_pdl):
    o = _oracle(tiny_model_path)
    e = _engine(tiny_model_path, max_rows=8, n_slots=2, use_graphs=use_graphs, use_pdl=use_pdl)
    prompt = [1, 17, 300, 42, 7]
    slot = e.alloc_slot()
    ref = o.forward(prompt, 0).numpy()
    got = e.prefill(slot, prompt, 0).float().cpu().numpy()          # logits of the last prompt token
    assert _rel_err(got[0], ref[-1]) < 2e-2
    tok, pos = int(np.argmax(ref[-1])), len(prompt)
    for _ in range(6):                                                # teacher-forced decode
        ref = o.forward([tok], pos).numpy()
        got = e.forward([tok], [pos], [slot]).float().cpu().numpy()
        assert _rel_err(got[0], ref[0]) < 2e-2
        tok, pos = int(np.argmax(ref[0])), pos + 1
    e.close()


def test_tensor_core_path_matches_oracle(tiny_model_path):
    """M > 8 rows run through the tcgen05 GEMM (prefill chunk of 20 rows, all logits)."""
    o = _oracle(tiny_model_path)
    e = _engine(tiny_model_path, max_rows=32, n_slots=2, use_graphs=False, use_pdl=False)
    rng = np.random.default_rng(0)
    prompt = rng.integers(0, 1024, 20).tolist()
    slot = e.alloc_slot()
    ref = o.forward(prompt, 0).numpy()
    got = e.forward(prompt, list(range(20)), [slot] * 20, logits_from=0).float().cpu().numpy()
    for t in range(20):
        assert _rel_err(got[t], ref[t]) < 3e-2, t
    # continue decoding on the GEMV path from the KV cache the GEMM path wrote
    tok = int(np.argmax(ref[-1]))
    ref1 = o.forward([tok], 20).numpy()
    got1 = e.forward([tok], [20], [slot]).float().cpu().numpy()
    assert _rel_err(got1[0], ref1[0]) < 3e-2
    e.close()


def test_wide_prefill_chunk_matches_oracle(tiny_model_path):
    """One 200-row forward: 256-row tcgen05 tiles + the tensor-core prefill attention (3 query tiles, causal
    inside the chunk), then a second chunk that starts at a non-zero position."""
    o = _oracle(tiny_model_path)
    e = _engine(tiny_model_path, max_rows=256, n_slots=2, use_graphs=True, use_pdl=True)
    rng = np.random.default_rng(5)
    prompt = rng.integers(1, 1024, 200).tolist()
    slot = e.alloc_slot()
    ref = o.forward(prompt, 0).numpy()
    got = e.forward(prompt, list(range(200)), [slot] * 200, logits_from=0).float().cpu().numpy()
    for t in (0, 1, 63, 64, 127, 128, 199):
        assert _rel_err(got[t], ref[t]) < 3e-2, t
    more = rng.integers(1, 1024, 40).tolist()
    ref2 = o.forward(more, 200).numpy()
    got2 = e.forward(more, list(range(200, 240)), [slot] * 40, logits_from=0).float().cpu().numpy()
    for t in (0, 17, 39):
        assert _rel_err(got2[t], ref2[t]) < 3e-2, t
    e.close()


@pytest.mark.parametrize("mode", ["resident", "scratch"])
def test_prompt_sized_batch_on_expanded_weights(tiny_tp_model_path, mode, monkeypatch):
    """One 500-row prompt forward through the two-phase GEMM (fp16-expanded weights, CTA pairs): same logits as the
    fused dequantising kernel (identical operand values, different summation order), the oracle at sampled rows, and
    a decode step that reads the KV cache the big batch wrote."""
    import dllama_b200.models.llama as LL
    monkeypatch.setattr(LL, "WH_MIN_ROWS", 256)
    monkeypatch.setattr(LL, "WH_MIN_ROWS_SCRATCH", 256)
    o = _oracle(tiny_tp_model_path)
    rng = np.random.default_rng(11)
    n = 500
    prompt = rng.integers(1, 2048, n).tolist()
    outs = {}
    for wh in (mode, "off"):
        e = _engine(tiny_tp_model_path, max_rows=512, n_slots=2, use_graphs=True, use_pdl=True, prefill_wh=wh)
        assert e.wh_mode == wh
        slot = e.alloc_slot()
        lg = e.forward(prompt, list(range(n)), [slot] * n, logits_from=n - 4).float().cpu().numpy()
        nxt = e.forward([5], [n], [slot]).float().cpu().numpy()
        outs[wh] = (lg, nxt)
        e.close()
    ref = o.forward(prompt, 0).numpy()
    ref1 = o.forward([5], n).numpy()
    errs = {wh: [_rel_err(outs[wh][0][i], ref[t]) for i, t in enumerate(range(n - 4, n))] + [_rel_err(outs[wh][1][0], ref1[0])]
            for wh in outs}
    cross = [_rel_err(outs[mode][0], outs["off"][0]), _rel_err(outs[mode][1], outs["off"][1])]
    print("errors vs oracle:", errs, "between the two kernels:", cross)
    assert max(errs[mode]) < 3e-2, errs
    # the two kernels see identical operand values and differ only in summation order; a Q80 rounding flip downstream
    # is worth up to 1/127 of a block's range, so they agree to about the same tolerance as with the oracle
    assert max(cross) < 2e-2, cross


def test_prompt_chunk_mixed_with_decode_rows(tiny_model_path):
    """A batch holding a 40-row prompt chunk of one sequence plus decode rows of two others: the chunk goes
    through the tensor-core prefill attention, the other rows through the per-row kernel (row_skip)."""
    e = _engine(tiny_model_path, max_rows=64, n_slots=4, use_graphs=True, use_pdl=True)
    rng = np.random.default_rng(9)
    pa, pb = rng.integers(1, 1024, 7).tolist(), rng.integers(1, 1024, 3).tolist()
    chunk = rng.integers(1, 1024, 40).tolist()
    sa, sb, sc = e.alloc_slot(), e.alloc_slot(), e.alloc_slot()
    e.prefill(sa, pa, 0)
    e.prefill(sb, pb, 0)
    toks = chunk + [11, 12]
    pos = list(range(40)) + [7, 3]
    slots = [sc] * 40 + [sa, sb]
    got = e.forward(toks, pos, slots, logits_from=39).float().cpu().numpy().copy()      # rows 39 (chunk end), 40, 41
    oc = _oracle(tiny_model_path)
    ref_c = oc.forward(chunk, 0).numpy()[-1]
    oa = _oracle(tiny_model_path)
    oa.forward(pa, 0)
    ref_a = oa.forward([11], 7).numpy()[0]
    ob = _oracle(tiny_model_path)
    ob.forward(pb, 0)
    ref_b = ob.forward([12], 3).numpy()[0]
    assert _rel_err(got[0], ref_c) < 3e-2 and _rel_err(got[1], ref_a) < 3e-2 and _rel_err(got[2], ref_b) < 3e-2
    e.close()


def test_multi_user_batch_isolated_kv(tiny_model_path):
    """Three users with different prompts and positions decode in ONE batch; each must equal its
    own single-user run (the reference's shared KV cache / shared position cannot do this)."""
    e = _engine(tiny_model_path, max_rows=8, n_slots=4, use_graphs=True, use_pdl=True)
    prompts = [[5, 6, 7], [100, 200, 300, 400, 500, 600], [9]]
    singles = []
    for p in prompts:
        o = _oracle(tiny_model_path)
        o.forward(p[:-1], 0) if len(p) > 1 else None
        singles.append(o.forward([p[-1]], len(p) - 1).numpy()[0])
    slots = [e.alloc_slot() for _ in prompts]
    for s, p in zip(slots, prompts):
        if len(p) > 1:
            e.prefill(s, p[:-1], 0)
    got = e.forward([p[-1] for p in prompts], [len(p) - 1 for p in prompts], slots).float().cpu().numpy()
    for i in range(len(prompts)):
        assert _rel_err(got[i], singles[i]) < 2e-2
    e.close()


def test_bucketed_rows_match_exact_shapes(tiny_model_path):
    """bucket_rows pads every batch to a power of two with rows of a hidden scratch sequence (serving:
    bounded number of CUDA graphs); results for the real rows must be unchanged."""
    ea = _engine(tiny_model_path, max_rows=32, n_slots=3, use_graphs=True, use_pdl=True)
    eb = _engine(tiny_model_path, max_rows=32, n_slots=3, use_graphs=True, use_pdl=True, bucket_rows=True)
    rng = np.random.default_rng(3)
    outs = []
    for e in (ea, eb):
        s1, s2 = e.alloc_slot(), e.alloc_slot()
        p1 = rng.integers(1, 1024, 11).tolist() if e is ea else p1
        p2 = rng.integers(1, 1024, 5).tolist() if e is ea else p2
        r = []
        r.append(e.prefill(s1, p1, 0).float().cpu().numpy().copy())                       # 11 rows -> 16, last-row logits
        r.append(e.prefill(s2, p2, 0).float().cpu().numpy().copy())                       # 5 rows -> 8 (GEMV path)
        # mixed batch: 3 prompt rows of a third sequence are not needed; two decode rows + all logits
        r.append(e.forward([7, 9, 11], [11, 5, 12], [s1, s2, s1], logits_from=0).float().cpu().numpy().copy())
        lg = e.forward([3, 4], [13, 6], [s1, s2], logits_from=0).float().cpu().numpy().copy()
        r.append(lg)
        # the greedy pick must agree with the engine's own logits for the real rows (not the padding rows)
        toks = e.decode_step_greedy([3, 4], [13, 6], [s1, s2], 0)
        assert toks == [int(t) for t in lg.argmax(axis=1)]
        r.append(e.forward([5], [14], [s1], logits_from=0).float().cpu().numpy().copy())
        outs.append(r)
        e.close()
    for a, b in zip(*outs):
        # a padded batch may cross the GEMV (exact int8 dp4a) / tcgen05 (fp16 operands) boundary and changes the
        # attention split count: same function, different rounding
        assert _rel_err(b, a) < 2e-2


def test_greedy_device_loop_matches_host_loop(tiny_model_path):
    e = _engine(tiny_model_path, max_rows=8, n_slots=2, use_graphs=True, use_pdl=True)
    prompt = [3, 1, 4, 1, 5, 9]
    slot = e.alloc_slot()
    logits = e.prefill(slot, prompt, 0)
    tok = int(torch.argmax(logits[0]))
    # host loop: H2D descriptor, forward, argmax, D2H token every step
    host = []
    t, pos = tok, len(prompt)
    for _ in range(8):
        t = e.decode_greedy_host([t], [pos], [slot])[0]
        host.append(t)
        pos += 1
    # device loop from the same state (KV rows are simply rewritten)
    e.ensure_capacity(slot, len(prompt) + 9)
    e.set_batch([tok], [len(prompt)], [slot])
    torch.cuda.synchronize()
    e.decode_greedy_device(8, 1)
    torch.cuda.synchronize()
    first = e.current_step() - 8
    dev = e.read_history(8, 1, first)[:, 0].tolist()
    assert dev == host
    e.close()


def test_nccl_cublas_baseline_engine_matches_oracle(tiny_model_path):
    """The B2 baseline (dequant + cuBLAS + PyTorch ops) computes the same function."""
    from dllama_b200.models.baseline_nccl import NcclCublasEngine
    o = _oracle(tiny_model_path)
    b = NcclCublasEngine(tiny_model_path, max_seq_len=512)
    prompt = [1, 17, 300, 42, 7]
    ref = o.forward(prompt, 0).numpy()
    got = b.forward(prompt, 0).float().cpu().numpy()
    assert _rel_err(got[-1], ref[-1]) < 5e-2          # bf16 operands
    ref1 = o.forward([11], 5).numpy()
    got1 = b.forward([11], 5).float().cpu().numpy()
    assert _rel_err(got1[0], ref1[0]) < 5e-2
    b.close()
