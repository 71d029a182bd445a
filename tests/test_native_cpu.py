"""Native CPU backend (csrc/host/cpu_backend.cpp): codecs and kernels against the NumPy twins and the reference's
published goldens, the engine against the PyTorch oracle, and root + workers over loopback TCP (workers without a
model file) against the single-node run -- the distributed test the reference does not have."""
import socket
import threading
import time

import numpy as np
import pytest
import torch

from dllama_b200.formats import mfile
from dllama_b200.models import native_cpu_engine as N
from dllama_b200.models.oracle import OracleLlama
from dllama_b200.ops import quant


def test_f16_conversion_matches_numpy_everywhere():
    L = N._lib()
    # every half value -> float -> half is the identity; float -> half rounds like NumPy (nearest even)
    for h in list(range(0, 0x7c01, 7)) + [0x0001, 0x03ff, 0x0400, 0x7bff, 0x7c00, 0x8001, 0xfbff]:
        f = L.dl_cpu_f16_to_f32(h)
        assert f == float(np.array([h], dtype=np.uint16).view(np.float16)[0])
        assert L.dl_cpu_f32_to_f16(f) == h
    rng = np.random.default_rng(0)
    xs = np.concatenate([rng.standard_normal(2000).astype(np.float32) * s for s in (1e-8, 1e-6, 1e-4, 1.0, 300.0, 7e4)])
    xs = np.concatenate([xs, np.array([0.0, -0.0, 65504.0, 65519.9, 65520.0, 1e9, 2.0 ** -24, 2.0 ** -25, 3 * 2.0 ** -26,
                                       6.1035e-05], dtype=np.float32)])
    with np.errstate(over="ignore"):
        want = xs.astype(np.float16).view(np.uint16)
    got = np.array([L.dl_cpu_f32_to_f16(float(x)) for x in xs], dtype=np.uint16)
    assert (got == want).all()


def test_q80_q40_codecs_match_numpy_twins_and_reference_tolerances():
    rng = np.random.default_rng(1)
    x = (rng.standard_normal(32 * 64) * 3).astype(np.float32)
    x[:32] = 0.0                                                   # an all-zero block (d = 0)
    for scalar in (False, True):
        q = N.quantize_q80(x, force_scalar=scalar)
        assert q.tobytes() == quant.quantize_q80(x).tobytes()      # bit-identical to the NumPy codec (half-even)
    y = N.dequantize_q80(q, x.size)
    assert np.abs(y - x).max() < 0.01 * 3 * 4                      # reference tolerance 0.01 at unit scale (nn-cpu-ops-test.cpp)
    assert np.allclose(y, quant.dequantize_q80(np.frombuffer(q.tobytes(), dtype=np.uint8)).reshape(-1))
    q4 = N.quantize_q40(x)
    y4 = N.dequantize_q40(q4, x.size)
    assert np.abs(y4 - x).max() < 0.13 * 3 * 4                     # reference tolerance 0.13 at unit scale
    py = np.frombuffer(quant.quantize_q40(x).tobytes(), dtype=np.uint8)
    assert np.allclose(N.dequantize_q40(py, x.size), quant.dequantize_q40(py).reshape(-1))


def test_q40_q80_matmul_kernels_agree_with_dequantised_f32():
    rng = np.random.default_rng(2)
    d, n, M = 40, 256, 11
    w = (rng.standard_normal((d, n)) * 0.05).astype(np.float32)
    x = rng.standard_normal((M, n)).astype(np.float32)
    wq = np.frombuffer(quant.quantize_q40(w).tobytes(), dtype=np.uint8)
    xq = N.quantize_q80(x)
    want = quant.dequantize_q80(np.frombuffer(xq.tobytes(), dtype=np.uint8)).reshape(M, n).astype(np.float64) @ \
        quant.dequantize_q40(wq).reshape(d, n).astype(np.float64).T
    ys = N.matmul_q40_q80(wq, d, n, xq, M, force_scalar=True)
    yv = N.matmul_q40_q80(wq, d, n, xq, M, force_scalar=False)
    assert np.abs(ys - want).max() < 1e-4 and np.abs(yv - want).max() < 1e-4
    assert np.abs(ys - yv).max() < 1e-5
    # batch invariance: a row's result does not depend on its batch mates (odd block counts included)
    for n2 in (256, 288):
        w2 = (rng.standard_normal((d, n2)) * 0.05).astype(np.float32)
        x2 = rng.standard_normal((M, n2)).astype(np.float32)
        wq2 = np.frombuffer(quant.quantize_q40(w2).tobytes(), dtype=np.uint8)
        xq2 = N.quantize_q80(x2)
        rb = n2 // 32 * 34
        rows = np.stack([N.matmul_q40_q80(wq2, d, n2, xq2[m * rb:(m + 1) * rb], 1)[0] for m in range(M)])
        assert np.array_equal(N.matmul_q40_q80(wq2, d, n2, xq2, M), rows)


def _free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _start_workers(n_workers, n_threads=1):
    ports = _free_ports(n_workers)
    errs = []

    def run(p):
        try:
            N.serve_worker(p, n_threads)
        except Exception as e:          # pragma: no cover - shown by the assertion below
            errs.append(e)

    threads = [threading.Thread(target=run, args=(p,), daemon=True) for p in ports]
    [t.start() for t in threads]
    time.sleep(0.3)
    return ports, threads, errs


def _rel(a, b):
    return float(np.abs(a - b).max() / max(1e-9, np.abs(b).max()))


def test_single_node_matches_the_oracle_with_batched_users(tiny_model_path):
    eng = N.NativeCpuEngine(tiny_model_path, n_threads=2, max_rows=16, n_slots=3, max_seq_len=128)
    o = OracleLlama(mfile.ModelFile(tiny_model_path), max_seq_len=128)
    p1, p2 = [5, 6, 7, 8, 9, 10, 11], [100, 200, 300]
    # two users' prompts packed into one batch (rows of a sequence in increasing position order, per-row slots)
    toks, pos, slots = p1 + p2, list(range(7)) + list(range(3)), [0] * 7 + [2] * 3
    lg = eng.forward(toks, pos, slots, logits_from=0).numpy().copy()
    w1 = o.forward(p1, 0).numpy()
    o.reset()
    w2 = o.forward(p2, 0).numpy()
    assert _rel(lg[:7], w1) < 2e-4 and _rel(lg[7:], w2) < 2e-4
    # one decode step for both users in the same forward
    t1, t2 = int(w1[-1].argmax()), int(w2[-1].argmax())
    lg = eng.forward([t1, t2], [7, 3], [0, 2], logits_from=0).numpy().copy()
    want2 = o.forward([t2], 3).numpy()
    o.reset()
    o.forward(p1, 0)
    want1 = o.forward([t1], 7).numpy()
    assert _rel(lg[0], want1[0]) < 2e-4 and _rel(lg[1], want2[0]) < 2e-4
    with pytest.raises(RuntimeError):
        eng.forward([1], [128], [0])                                # outside the context window
    with pytest.raises(RuntimeError):
        eng.forward([99999], [0], [0])                              # outside the vocabulary
    eng.close()


@pytest.mark.parametrize("n_nodes", [2, 4])
def test_root_and_workers_over_tcp_match_the_oracle_and_the_single_node(tiny_model_path, n_nodes):
    ports, threads, errs = _start_workers(n_nodes - 1)
    eng = N.NativeCpuEngine(tiny_model_path, workers=[("127.0.0.1", p) for p in ports], n_threads=1, max_rows=16, n_slots=2,
                            max_seq_len=128)
    one = N.NativeCpuEngine(tiny_model_path, n_threads=1, max_rows=16, n_slots=2, max_seq_len=128)
    o = OracleLlama(mfile.ModelFile(tiny_model_path), n_nodes=n_nodes, max_seq_len=128)     # per-node Q80 partial sums
    prompt = [5, 6, 7, 8, 9, 10, 11, 12, 13]
    lg = eng.forward(prompt, list(range(9)), [1] * 9).numpy().copy()
    assert _rel(lg, o.forward(prompt, 0).numpy()) < 2e-4
    st = eng.network_stats()
    h = eng.h
    # every node sends its Q80 slice of every row to every peer twice per layer (+ control, x and logits traffic)
    slice_bytes = h.dim // 32 * 34
    assert st["sent_bytes"] >= 2 * h.n_layers * 9 * slice_bytes * (n_nodes - 1) and st["recv_bytes"] > 0 and st["sync_us"] > 0
    # greedy continuation: distributed run == single node, token for token (what the reference shows manually)
    ref_lg = one.forward(prompt, list(range(9)), [1] * 9).numpy().copy()
    tok_d, tok_s = int(lg[-1].argmax()), int(ref_lg[-1].argmax())
    out_d, out_s = [], []
    for i in range(6):
        out_d.append(tok_d)
        out_s.append(tok_s)
        tok_d = eng.decode_step_greedy([tok_d], [9 + i], [1])[0]
        tok_s = one.decode_step_greedy([tok_s], [9 + i], [1])[0]
    assert out_d == out_s
    eng.close()                                                     # sends the stop packet
    one.close()
    for t in threads:
        t.join(timeout=30)
        assert not t.is_alive()
    assert not errs, errs


def test_f32_model_on_two_nodes(tmp_path):
    p = str(tmp_path / "tiny_f32.m")
    mfile.write_random_model(p, "tiny", weight_type=quant.F32, seed=5)
    ports, threads, errs = _start_workers(1)
    eng = N.NativeCpuEngine(p, workers=[("127.0.0.1", ports[0])], n_threads=2, max_rows=8, n_slots=1, max_seq_len=64)
    o = OracleLlama(mfile.ModelFile(p), n_nodes=2, max_seq_len=64)
    prompt = [3, 4, 5, 6]
    lg = eng.forward(prompt, [0, 1, 2, 3], [0] * 4).numpy().copy()
    assert _rel(lg, o.forward(prompt, 0).numpy()) < 1e-4
    eng.close()
    threads[0].join(timeout=30)
    assert not errs, errs


def test_invalid_node_counts_are_rejected(tiny_model_path):
    with pytest.raises(RuntimeError, match="KV heads|power of two"):
        N.NativeCpuEngine(tiny_model_path, workers=[("127.0.0.1", 1)] * 7, max_seq_len=64)      # tiny has 4 KV heads
    with pytest.raises(RuntimeError, match="power of two"):
        N.NativeCpuEngine(tiny_model_path, workers=[("127.0.0.1", 1)] * 2, max_seq_len=64)
    with pytest.raises((RuntimeError, FileNotFoundError)):
        N.NativeCpuEngine("/nonexistent.m")


# ---- op-level goldens of the reference's unit tests (src/nn/nn-cpu-ops-test.cpp, src/nn/nn-vulkan-test.cpp) -------
def _op_lib():
    import ctypes as C
    L = N._lib()
    L.dl_cpu_inv_rms.restype = C.c_float
    L.dl_cpu_inv_rms.argtypes = [C.c_void_p, C.c_int, C.c_float]
    L.dl_cpu_rms_norm.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float]
    L.dl_cpu_activation.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.dl_cpu_rope.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_float, C.c_float, C.c_float, C.c_int]
    L.dl_cpu_softmax.argtypes = [C.c_void_p, C.c_int]
    return L


def test_attention_softmax_kernel_golden_and_accuracy():
    """The vectorised softmax of the attention kernel: the reference's 8-value golden within its tolerance
    (nn-cpu-ops-test.cpp:187-198; its own values come from an approximate exp) and float64 accuracy for every tail
    length, large ranges and fully masked-out magnitudes."""
    L = _op_lib()
    y = (np.arange(8) / 8.0).astype(np.float32)
    L.dl_cpu_softmax(y.ctypes.data, 8)
    gold = [0.077399, 0.087780, 0.099500, 0.112761, 0.127778, 0.144793, 0.164072, 0.185917]
    assert np.abs(y - np.array(gold, dtype=np.float32)).max() < 0.001
    rng = np.random.default_rng(0)
    for n in (1, 2, 7, 8, 9, 15, 16, 17, 100, 1000, 2049):
        for scale in (0.1, 1.0, 10.0, 60.0):
            x = (rng.standard_normal(n) * scale).astype(np.float32)
            got = x.copy()
            L.dl_cpu_softmax(got.ctypes.data, n)
            x64 = x.astype(np.float64)
            ref = np.exp(x64 - x64.max())
            ref /= ref.sum()
            assert np.abs(got - ref).max() < 1e-6 * ref.max() + 1e-12 and abs(float(got.sum()) - 1.0) < 1e-5
    x = np.array([-200.0, 0.0, -100.0], dtype=np.float32)
    L.dl_cpu_softmax(x.ctypes.data, 3)
    assert x.tolist() == [0.0, 1.0, 0.0]


def test_inv_rms_rmsnorm_and_silu_goldens():
    L = _op_lib()
    x = np.array([0.1, 0.3, 0.2, 0.4, 0.6, 0.5, 0.0, 0.8], dtype=np.float32)
    assert abs(L.dl_cpu_inv_rms(x.ctypes.data, 8, 1e-5) - 1.0 / 0.4402) < 0.001            # nn-cpu-ops-test.cpp:102-118
    # the Vulkan op test's rmsnorm goldens: rows of 256 values i/256 scaled by (b+1)  (nn-vulkan-test.cpp:91-93)
    w = np.linspace(0.5, 1.5, 8, dtype=np.float32)
    y = np.empty(8, dtype=np.float32)
    L.dl_cpu_rms_norm(x.ctypes.data, w.ctypes.data, y.ctypes.data, 8, 1e-5)
    inv = 1.0 / np.sqrt((x.astype(np.float64) ** 2).mean() + 1e-5)
    assert np.allclose(y, w * (inv * x), rtol=1e-6)
    # rmsNorm on Q80-rounded input stays within the reference's 0.01 of the f32 result (nn-cpu-ops-test.cpp:121-137)
    rng = np.random.default_rng(3)
    xr, wr = rng.uniform(-1, 1, 256).astype(np.float32), rng.uniform(-1, 1, 256).astype(np.float32)
    xq = N.dequantize_q80(N.quantize_q80(xr), 256)
    y1, y2 = np.empty(256, dtype=np.float32), np.empty(256, dtype=np.float32)
    inv = L.dl_cpu_inv_rms(xr.ctypes.data, 256, 1e-5)
    L.dl_cpu_rms_norm(xr.ctypes.data, wr.ctypes.data, y1.ctypes.data, 256, 1e-5)
    y2[:] = wr * (inv * xq)
    assert np.abs(y1 - y2).max() < 0.01
    s = (np.arange(8) / 8.0).astype(np.float32)
    L.dl_cpu_activation(s.ctypes.data, 8, 1)
    want = [0.000000, 0.066401, 0.140544, 0.222250, 0.311233, 0.407116, 0.509461, 0.617802]   # nn-cpu-ops-test.cpp:200-216
    assert np.abs(s - np.array(want, dtype=np.float32)).max() < 0.001
    g = np.array([-1.0, 0.0, 1.0, 2.0], dtype=np.float32)
    L.dl_cpu_activation(g.ctypes.data, 4, 0)
    assert np.allclose(g, [-0.158808, 0.0, 0.841192, 1.954598], atol=1e-5)                      # tanh-GELU


def test_rope_goldens_of_the_reference_vulkan_test():
    """x = 1 everywhere, dim 2048 = 32 heads of 64, theta 5e5, Llama-3.1 scaling 32 / 1 / 4 / 8192, positions 6 and 31
    (nn-vulkan-test.cpp:426-482) -- for the host kernel and for the oracle's rotation used by the CUDA tests."""
    from dllama_b200.models.oracle import apply_rope, rope_frequencies
    L = _op_lib()
    gold = {6: {0: 1.239586, 1: 0.680755, 2: 0.077202, 3: -1.412105, 1988: -1.356766, 2022: 0.999923, 2023: 1.000077},
            31: {0: 1.318780, 1: 0.510705, 1078: 0.999985, 1079: 1.000015}}
    h = mfile.LlmHeader(dim=2048, n_heads=32, n_kv_heads=8, rope_theta=500000.0, rope_type=mfile.ROPE_LLAMA3_1,
                        rope_scaling_factor=32.0, rope_scaling_low_freq_factor=1.0, rope_scaling_high_freq_factor=4.0,
                        rope_scaling_orig_max_seq_len=8192)
    f = rope_frequencies(h)
    for pos, vals in gold.items():
        row = np.ones(2048, dtype=np.float32)
        L.dl_cpu_rope(row.ctypes.data, 2048, 64, pos, 500000.0, 2, 32.0, 1.0, 4.0, 8192)
        y = apply_rope(torch.ones(1, 2048), torch.tensor([pos]), f, 64)[0].numpy()
        for i, v in vals.items():
            assert abs(row[i] - v) < 5e-6, (pos, i, row[i], v)
            assert abs(y[i] - v) < 5e-6, (pos, i, y[i], v)


def test_long_greedy_continuation_is_reproducible_and_thread_count_invariant(tiny_tp_model_path):
    """The role of the reference's examples/macbeth.sh (a long fixed-seed continuation compared exactly), runnable
    anywhere: 4 nodes over loopback TCP, 160 tokens; the ids must not depend on the thread count (rows and output
    ranges are split across threads, the arithmetic per output value is fixed) nor change from run to run."""
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("long_context", os.path.join(root, "examples", "long_context.py"))
    lc = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(lc)
    a = lc.generate_cpu(tiny_tp_model_path, 160, 4, 1)
    b = lc.generate_cpu(tiny_tp_model_path, 160, 4, 2)
    assert len(a) == 160 and a == b and len(set(a)) > 8
