"""Every DLLAMA_* environment variable the framework reads, in one place: `python -m dllama_b200.utils.envvars`
(or `dllama --env-help`) prints the table.  Values are read where they are used; this module is the documentation
and the check that nothing undocumented creeps in (tests/test_cli.py scans the sources against it)."""
from __future__ import annotations

ENV_VARS = {
    # engine policy
    "DLLAMA_MEGA": ("auto", "persistent decode kernel for batches of <= 4 rows: auto = on for 8 ranks, 1 / 0 force it"),
    "DLLAMA_MEGA_CTAS": ("SM count", "grid of the persistent kernel (debugging)"),
    "DLLAMA_MEGA_TIMEOUT_S": ("20", "deadline of every spin loop in the persistent kernel; a hang becomes a trap with a code"),
    "DLLAMA_MEGA_PF": ("0", "persistent kernel: L2 prefetch distance in ring stages (measured: no gain)"),
    "DLLAMA_MEGA_DBG": ("0", "persistent kernel timing experiments (results invalid): 2 = no arithmetic, 4 = do not wait for weights"),
    "DLLAMA_GEMV_MAX_M": ("3", "largest batch that uses the dp4a GEMV path of the launch sequence"),
    "DLLAMA_LL": ("auto", "LL-word exchange of the launch sequence: auto = groups of <= 4 ranks, 1 / 0 force it"),
    "DLLAMA_WIRE_Q80": ("1", "partial sums of tensor-core batches travel as Q80 rows (0 = f32 rows)"),
    "DLLAMA_PREFILL_WH": ("auto", "fp16-expanded weights for prompt-sized batches: auto / resident / scratch / off"),
    "DLLAMA_WH_MIN_ROWS": ("768", "batch size from which the resident fp16 copy is used"),
    "DLLAMA_WH_MIN_ROWS_SCRATCH": ("2048", "same for the scratch mode (expansion per call)"),
    "DLLAMA_PREFILL_TC_MIN_ROWS": ("128", "batch size from which prompt chunks use the tcgen05 attention kernel"),
    # serving / tensor-parallel control
    "DLLAMA_HOST_SAMPLER": ("0", "1 = sample on the host (reference-exact sampler) instead of the device kernel"),
    "DLLAMA_MAX_WAITING": ("4096", "requests the native server queues before it answers 503"),
    "DLLAMA_CTL_TIMEOUT_S": ("120", "how long the root waits for a worker's acknowledgement of a batch"),
    "DLLAMA_RUN_NONCE": ("derived", "run identifier of the control mailbox (stale files of earlier runs are ignored)"),
    "DLLAMA_STEP_LOG": ("0", "1 = the native server keeps (start, rows, first logits row, sampled rows, seconds) per step"),
    # models / tools
    "DLLAMA_MODEL_DIR": ("/dev/shm", "where bench.py and the scripts keep random-init model files"),
    "DLLAMA_FAST_WRITER": ("0", "1 = write random-init .m files with the multi-threaded C++ writer (default above 16 GB)"),
    # host (CPU) backend
    "DLLAMA_BIND_ADDR": ("0.0.0.0", "IPv4 address `dllama worker --device cpu` listens on"),
    "DLLAMA_CPU_SCALAR": ("0", "1 = scalar kernels only"),
    "DLLAMA_CPU_NO_AVX512": ("0", "1 = no AVX-512 VNNI kernels (AVX2 kernels instead)"),
    "DLLAMA_CPU_SPIN": ("200000", "spin iterations of a pool thread between parallel regions before it sleeps (0 = sleep at once)"),
    "DLLAMA_CPU_SPIN_POLLS": ("20000", "non-blocking polls of the TCP exchange before it falls back to poll()"),
    "DLLAMA_CPU_PROFILE": ("0", "1 = per-op time table at exit"),
    "DLLAMA_CPU_DEBUG_OPS": ("0", "1 = print every op's output checksum"),
}


def table() -> str:
    w = max(len(k) for k in ENV_VARS)
    return "\n".join(f"{k:<{w}}  [{d}]  {h}" for k, (d, h) in ENV_VARS.items())


if __name__ == "__main__":
    print(table())
