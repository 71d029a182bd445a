"""The f32-weights / f32-buffer path (reference: matmul_F32_F32_F32 + f32 merge-add): F32Engine vs the
oracle, single process and tensor-parallel over a 2-rank gloo group (CPU), plus the same on a GPU."""
import os
import socket
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _model(tmp):
    from dllama_b200.formats import mfile
    from dllama_b200.ops import quant
    path = os.path.join(tmp, "tiny_f32.m")
    if not os.path.exists(path):
        mfile.write_random_model(path, "tiny", weight_type=quant.F32, seed=3)
    return path


def _oracle(path):
    from dllama_b200.formats.mfile import ModelFile
    from dllama_b200.models.oracle import OracleLlama
    return OracleLlama(ModelFile(path), sync_q80=False, max_seq_len=256)


def _check_engine(e, o, tol):
    s = e.alloc_slot()
    prompt = [1, 5, 9, 200, 33]
    ref = o.forward(prompt, 0).numpy()
    got = e.forward(prompt, list(range(5)), [s] * 5).float().cpu().numpy()
    assert np.abs(got - ref).max() / np.abs(ref).max() < tol
    # two sequences at different positions in one batch
    s2 = e.alloc_slot()
    e.prefill(s2, [7, 8, 9], 0)
    got = e.forward([11, 12], [5, 3], [s, s2]).float().cpu().numpy()
    ref1 = o.forward([11], 5).numpy()
    assert np.abs(got[0] - ref1[0]).max() / np.abs(ref1).max() < tol
    o2 = _oracle(e.mf.path)
    o2.forward([7, 8, 9], 0)
    ref2 = o2.forward([12], 3).numpy()
    assert np.abs(got[1] - ref2[0]).max() / np.abs(ref2).max() < tol
    assert e.decode_step_greedy([11], [6], [s])[0] == int(np.argmax(o.forward([11], 6).numpy()[0]))


def test_f32_engine_matches_oracle_cpu(tmp_path):
    from dllama_b200.models.f32_engine import F32Engine
    path = _model(str(tmp_path))
    e = F32Engine(path, max_rows=16, n_slots=3, max_seq_len=256, device="cpu")
    _check_engine(e, _oracle(path), 1e-5)
    e.close()


def _tp_worker(rank, world, port, path, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from dllama_b200.models.f32_engine import F32Engine
        e = F32Engine(path, rank=rank, world=world, max_rows=16, n_slots=3, max_seq_len=256, device="cpu")
        s = e.alloc_slot()
        prompt = [1, 5, 9, 200, 33]
        got = e.forward(prompt, list(range(5)), [s] * 5).numpy().copy()
        tok = e.decode_step_greedy([11], [5], [s])[0]
        if rank == 0:
            o = _oracle(path)
            ref = o.forward(prompt, 0).numpy()
            ret["err"] = float(np.abs(got - ref).max() / np.abs(ref).max())
            ret["tok_ok"] = tok == int(np.argmax(o.forward([11], 5).numpy()[0]))
        toks = [None] * world
        dist.all_gather_object(toks, tok)
        if rank == 0:
            ret["agree"] = all(t == toks[0] for t in toks)
        e.close()
    finally:
        dist.destroy_process_group()


def test_f32_engine_tensor_parallel_gloo(tmp_path):
    """world_size 2 over gloo: per-rank weight slices + all-reduce of the wo / w2 partial sums must
    reproduce the single-node result (the host-side TP logic, no GPU needed)."""
    path = _model(str(tmp_path))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ret = mp.Manager().dict()
    mp.spawn(_tp_worker, args=(2, port, path, ret), nprocs=2, join=True)
    assert ret["err"] < 1e-5 and ret["tok_ok"] and ret["agree"]


@pytest.mark.gpu
def test_f32_engine_matches_oracle_gpu(tmp_path):
    from dllama_b200.models.f32_engine import F32Engine
    path = _model(str(tmp_path))
    e = F32Engine(path, max_rows=16, n_slots=3, max_seq_len=256)
    assert e.device.type == "cuda"
    _check_engine(e, _oracle(path), 1e-4)
    e.close()
