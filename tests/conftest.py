import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # a fresh checkout has no built artefacts (*.so are git-ignored): build what is missing so the suite is
    # self-sufficient (host library: g++, seconds; CUDA library: nvcc cross-compiles without a GPU)
    import build as _build
    lib_dir = os.path.join(ROOT, "distributed-llama-multiusers_b200", "lib")
    if not os.path.exists(os.path.join(lib_dir, "libdllama_b200_host.so")):
        _build.build_host()
    if not os.path.exists(os.path.join(lib_dir, "libdllama_b200_cuda.so")):
        try:
            _build.build_cuda()
        except Exception as e:          # CPU-only tests do not need it
            print(f"[conftest] CUDA library not built: {e}")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def tiny_model_path(tmp_path_factory):
    from dllama_b200.formats import mfile
    p = str(tmp_path_factory.mktemp("models") / "tiny.m")
    mfile.write_random_model(p, "tiny", exact=True, seed=3)
    return p


@pytest.fixture(scope="session")
def tiny_tp_model_path(tmp_path_factory):
    from dllama_b200.formats import mfile
    p = str(tmp_path_factory.mktemp("models") / "tiny_tp8.m")
    mfile.write_random_model(p, "tiny-tp8", exact=True, seed=4)
    return p
