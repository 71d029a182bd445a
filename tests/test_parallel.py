"""Host-side distributed plumbing (no GPU): the shared-memory control channel between root and workers
(reference: LlmControlPacket broadcast, src/app.cpp:198-229), the TCP bootstrap of `dllama worker`
(reference: NnNetwork::serve / connect, src/nn/nn-network.cpp:264-348) and the tensor-parallel slicing rules
(reference: src/app.cpp:237-240, src/nn/nn-core.cpp:198-266)."""
import os
import socket
import threading
import time

import pytest

from dllama_b200.cli.app import AppArgs, dial_workers, serve_worker_once
from dllama_b200.parallel.control import OP_FORWARD, OP_STOP, OP_WARMUP, ControlChannel
from dllama_b200.parallel.symm import validate_tp


def test_control_channel_root_to_two_workers():
    name = f"test_{os.getpid()}_{int(time.time() * 1e6) % 1000000}"
    root = ControlChannel(name, 0, 3, max_rows=8)
    got = {1: [], 2: []}

    def worker(rank):
        ch = ControlChannel(name, rank, 3, max_rows=8)
        while True:
            pkt = ch.recv(timeout_s=20)
            assert pkt is not None, "worker timed out"
            got[rank].append(pkt)
            if pkt["op"] == OP_STOP:
                break
        ch.close()

    ts = [threading.Thread(target=worker, args=(r,)) for r in (1, 2)]
    [t.start() for t in ts]
    root.send(OP_WARMUP)
    root.send(OP_FORWARD, [5, 6, 7], [10, 11, 0], [2, 2, 1], logits_from=1, greedy=True, released=[3])
    # the writer must wait for both acks before it overwrites the mailbox with the next packet
    root.send(OP_FORWARD, [9], [12], [2], logits_from=0, greedy=False, released=[])
    root.send(OP_STOP)
    [t.join(timeout=30) for t in ts]
    root.close(unlink=True)
    for r in (1, 2):
        ops = [p["op"] for p in got[r]]
        assert ops == [OP_WARMUP, OP_FORWARD, OP_FORWARD, OP_STOP]
        a, b = got[r][1], got[r][2]
        assert a["M"] == 3 and a["tokens"] == [5, 6, 7] and a["positions"] == [10, 11, 0] and a["slots"] == [2, 2, 1]
        assert a["logits_from"] == 1 and a["greedy"] is True and a["released"] == [3]
        assert b["M"] == 1 and b["tokens"] == [9] and b["positions"] == [12] and b["greedy"] is False and b["released"] == []
    assert not os.path.exists(root.path)


def test_worker_bootstrap_over_tcp(tmp_path):
    """`dllama worker --port P` + root `--workers 127.0.0.1:P`: the root hands each worker its rank, the world
    size, the rendezvous address, the model path and the engine geometry."""
    ports = []
    for _ in range(2):
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            ports.append(s.getsockname()[1])
    cfgs = {}

    def worker(port):
        cfgs[port] = serve_worker_once(port)

    ts = [threading.Thread(target=worker, args=(p,)) for p in ports]
    [t.start() for t in ts]
    time.sleep(0.3)
    model = tmp_path / "m.m"
    model.write_bytes(b"x")
    args = AppArgs.parse(["dllama", "inference", "--model", str(model), "--tokenizer", "t.t", "--prompt", "x", "--steps", "1",
                          "--max-seq-len", "512", "--workers", f"127.0.0.1:{ports[0]}", f"127.0.0.1:{ports[1]}"], True)
    world = dial_workers(args, master_port=23456)
    [t.join(timeout=30) for t in ts]
    assert world == 3
    ranks = sorted(c["rank"] for c in cfgs.values())
    assert ranks == [1, 2]
    for c in cfgs.values():
        assert c["world"] == 3 and c["master_port"] == 23456 and c["master_addr"] == "127.0.0.1"
        assert c["model"] == str(model) and c["max_seq_len"] == 512 and c["max_batch"] == args.max_batch
        assert c["ctl"] == str(os.getpid())


def test_tensor_parallel_slicing_rules():
    ok = dict(n_kv_heads=8, dim=4096, hidden_dim=14336, n_heads=32, vocab=128256)
    for w in (1, 2, 4, 8):
        validate_tp(w, **ok)
    with pytest.raises(ValueError, match="power of two"):
        validate_tp(3, **ok)
    with pytest.raises(ValueError, match="KV heads"):
        validate_tp(16, **ok)
    with pytest.raises(ValueError):                       # Q40T tiles need 128-multiples of the per-rank K extents
        validate_tp(8, n_kv_heads=8, dim=512, hidden_dim=1024, n_heads=8, vocab=1024)
    validate_tp(8, n_kv_heads=8, dim=512, hidden_dim=1024, n_heads=8, vocab=1024, q40=False)
