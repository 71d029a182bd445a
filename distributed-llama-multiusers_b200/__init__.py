"""b200-dllama: a Blackwell-native tensor-parallel Llama inference framework.

Same public surface as distributed-llama (MultiUsers fork): the ``dllama inference / chat /
worker`` CLI, the multi-user ``dllama-api`` OpenAI-style server, the ``.m`` model and ``.t``
tokenizer formats, Q40 weights with Q80 activations/sync (plus f32/f32), and the
``2^n <= n_kv_heads`` slicing rule -- re-designed for 8xB200 (sm_100a kernels, peer-memory
collectives fused into the GEMMs).  Import as ``dllama_b200``.
"""
__version__ = "0.1.0"

REPO_ROOT = __import__("os").path.dirname(__import__("os").path.dirname(
    __import__("os").path.abspath(__file__)))
