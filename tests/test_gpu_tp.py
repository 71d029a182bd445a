"""Tensor-parallel correctness inside the GPU suite: scripts/tp_check.py under torchrun on 2, 4 and 8 ranks when the box
has that many GPUs (skipped otherwise).  Every rank runs the engine on its weight slice; rank 0 compares logits with the
CPU oracle run with n_nodes = world (per-node Q80 partial sums, like the reference's SYNC_NODE_SLICES + OP_MERGE_ADD,
src/nn/nn-network.cpp:537-569, src/llm.cpp:314-322), all ranks must agree on greedy tokens, with CUDA graphs on and off.
Both exchange implementations are covered: the launch sequence (peer stores + flags / LL words) and the persistent
decode kernel (Q80 LL words, merger CTAs, cross-rank greedy pick)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run_tp_check(world: int, mega: str, port: int, script: str = "tp_check.py", marker: str = "TP_CHECK PASS"):
    env = dict(os.environ)
    env["DLLAMA_MEGA"] = mega
    env["DLLAMA_MEGA_TIMEOUT_S"] = "20"
    env.pop("RANK", None)
    env.pop("WORLD_SIZE", None)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "scripts", script)]
    p = subprocess.run(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    tail = "\n".join(p.stdout.splitlines()[-40:])
    assert p.returncode == 0 and f"{marker} world={world}" in p.stdout, tail
    assert "FAIL" not in p.stdout, tail


@pytest.mark.parametrize("world,mega", [(2, "0"), (2, "1"), (4, "1"), (8, "0"), (8, "1")])
def test_tensor_parallel_matches_oracle(world, mega):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, this box has {torch.cuda.device_count()}")
    _run_tp_check(world, mega, 29600 + 10 * world + int(mega))


@pytest.mark.parametrize("world", [2, 4])
def test_f32_path_tensor_parallel(world):
    """The native f32 path (gemm_f32 + f32 peer-store exchange, no NCCL on the data path) on 2 and 4 ranks."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, this box has {torch.cuda.device_count()}")
    _run_tp_check(world, "auto", 29700 + world, script="tp_check_f32.py", marker="TP_CHECK_F32 PASS")
