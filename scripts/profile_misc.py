#!/usr/bin/env python3
"""ncu target for the kernels that are not matmuls (1 GPU, 2-layer 8B-shaped model): one 4096-row prompt forward
(attention_prefill_tc_kernel, reduce_norm_quant_kernel at thousands of rows), decode steps at a 4096-token context
(attention_kernel, split-KV), and a 32-row sampling call (sample_kernel).

  python scripts/profile_misc.py                                       (plain run, must exit 0)
  ncu --set full --clock-control none --import-source on -k regex:'attention_prefill_tc|attention_kernel|reduce_norm_quant|sample_kernel' -c 14 -o out python scripts/profile_misc.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    from dllama_b200.formats import mfile
    from dllama_b200.models.llama import LlamaEngine
    from dllama_b200.ops import kernels as K, quant
    path = "/dev/shm/dllama_b200_llama3.1-8b-l2_q40.m"
    if not os.path.exists(path):
        mfile.write_random_model(path, "llama3.1-8b", quant.Q40, seed=1, n_layers=2)
    n = 4096
    eng = LlamaEngine(path, max_rows=n, n_slots=2, max_seq_len=n + 64, use_graphs=False, use_pdl=False)
    prompt = np.random.default_rng(1).integers(1, eng.h.vocab_size, n).tolist()
    slot = eng.alloc_slot()
    out = eng.decode_step_greedy(prompt, list(range(n)), [slot] * n, logits_from=n - 1)       # prompt forward
    tok = out[0]
    for i in range(2):                                                                         # decode at 4096 tokens of context
        tok = eng.decode_step_greedy([tok], [n + i], [slot])[0]
    logits = torch.randn(32, eng.h.vocab_size, device="cuda")
    temp = torch.full((32,), 0.8, device="cuda")
    topp = torch.full((32,), 0.9, device="cuda")
    coin = torch.rand(32, device="cuda")
    toks = torch.zeros(32, dtype=torch.int32, device="cuda")
    K.sample_rows(logits, temp, topp, coin, toks)
    torch.cuda.synchronize()
    print("ok", tok, toks[:4].tolist())
    eng.close()


if __name__ == "__main__":
    main()
