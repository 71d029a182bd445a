"""ctypes binding of libdllama_b200_cuda.so (C ABI declared in include/dllama_b200.h).

The library is built in-tree by ``build.py`` and MUST be present on a GPU box: every op fails
loudly (``NativeLibraryError``) instead of silently falling back to PyTorch.
"""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA_LIB_PATH = os.path.join(_PKG, "lib", "libdllama_b200_cuda.so")
HOST_LIB_PATH = os.path.join(_PKG, "lib", "libdllama_b200_host.so")

DL_MAX_PEERS = 8
EPI_STORE, EPI_QKV_ROPE, EPI_SILU_MUL = 0, 1, 2


class NativeLibraryError(RuntimeError):
    pass


class DlEpilogue(C.Structure):
    """Mirror of ``DlEpilogue`` in include/dllama_b200.h (field order and types must match)."""
    _fields_ = [
        ("kind", C.c_int),
        ("dst", C.c_void_p * DL_MAX_PEERS),
        ("n_dst", C.c_int),
        ("ld", C.c_int),
        ("dst_off", C.c_longlong),
        ("q80_roundtrip", C.c_int),
        ("residual", C.c_void_p),
        ("residual_ld", C.c_int),
        ("flag", C.c_void_p * DL_MAX_PEERS),
        ("step_counter", C.c_void_p),
        ("flag_mul", C.c_uint32),
        ("flag_add", C.c_uint32),
        ("cta_counter", C.c_void_p),
        ("q_out", C.c_void_p),
        ("q_dim", C.c_int),
        ("kv_dim", C.c_int),
        ("head_size", C.c_int),
        ("rope_freqs", C.c_void_p),
        ("pos", C.c_void_p),
        ("slot", C.c_void_p),
        ("block_table", C.c_void_p),
        ("max_pages", C.c_int),
        ("page_size", C.c_int),
        ("n_kv_heads", C.c_int),
        ("k_cache", C.c_void_p),
        ("v_cache", C.c_void_p),
        ("kv_f32", C.c_int),
        ("xq_out", C.c_void_p),
        ("xh_out", C.c_void_p),
        ("out_mcap", C.c_int),
        ("act_gelu", C.c_int),
        ("ll", C.c_int),
    ]


class DlNormProlog(C.Structure):
    """Mirror of ``DlNormProlog`` in include/dllama_b200.h."""
    _fields_ = [
        ("x_in", C.c_void_p),
        ("x_out", C.c_void_p),
        ("partials", C.c_void_p),
        ("n_src", C.c_int),
        ("src_stride", C.c_longlong),
        ("flags", C.c_void_p),
        ("step_counter", C.c_void_p),
        ("flag_mul", C.c_uint32),
        ("flag_add", C.c_uint32),
        ("norm_w", C.c_void_p),
        ("eps", C.c_float),
        ("ll", C.c_int),
    ]


class DlDecodeLayer(C.Structure):
    """Mirror of ``DlDecodeLayer`` (per-layer device pointers of the persistent decode kernel)."""
    _fields_ = [("qkv", C.c_void_p), ("wo", C.c_void_p), ("w13", C.c_void_p), ("w2", C.c_void_p),
                ("rms_att", C.c_void_p), ("rms_ffn", C.c_void_p), ("k_cache", C.c_void_p), ("v_cache", C.c_void_p)]


class DlDecodeArgs(C.Structure):
    """Mirror of ``DlDecodeArgs`` in include/dllama_b200.h (field order and types must match)."""
    _fields_ = [
        ("n_layers", C.c_int), ("dim", C.c_int), ("q_dim", C.c_int), ("kv_dim", C.c_int), ("hid", C.c_int),
        ("vocab_loc", C.c_int), ("vocab", C.c_int), ("head_size", C.c_int), ("n_heads", C.c_int),
        ("n_kv_heads", C.c_int),
        ("rank", C.c_int), ("world", C.c_int),
        ("eps", C.c_float),
        ("act_gelu", C.c_int),
        ("q80_sync", C.c_int),
        ("layers", C.c_void_p),
        ("rms_final", C.c_void_p),
        ("wcls", C.c_void_p),
        ("embedding", C.c_void_p),
        ("rope_freqs", C.c_void_p),
        ("M", C.c_int), ("row0", C.c_int),
        ("tokens", C.c_void_p),
        ("pos", C.c_void_p),
        ("slot", C.c_void_p),
        ("block_table", C.c_void_p),
        ("max_pages", C.c_int),
        ("x", C.c_void_p * 2),
        ("q", C.c_void_p),
        ("h", C.c_void_p),
        ("att_o", C.c_void_p),
        ("att_ml", C.c_void_p),
        ("max_splits", C.c_int),
        ("att_splits", C.c_int),
        ("cta_best", C.c_void_p),
        ("grid_bar", C.c_void_p),
        ("step_counter", C.c_void_p),
        ("ll_att", C.c_void_p * DL_MAX_PEERS),
        ("ll_ffn", C.c_void_p * DL_MAX_PEERS),
        ("ll_tok", C.c_void_p * DL_MAX_PEERS),
        ("logits", C.c_void_p * DL_MAX_PEERS),
        ("flag_logits", C.c_void_p * DL_MAX_PEERS),
        ("greedy", C.c_int),
        ("tokens_out", C.c_void_p),
        ("advance", C.c_int),
        ("history", C.c_void_p),
        ("history_cap", C.c_int),
        ("timeout_ns", C.c_ulonglong),
        ("debug", C.c_void_p),
        ("trace", C.c_void_p),
        ("trace_cta", C.c_int),
        ("dbg", C.c_int),
        ("pf_dist", C.c_int),
    ]


_cuda = None


def cuda_lib():
    """Load (once) and return the CUDA library with argtypes set."""
    global _cuda
    if _cuda is not None:
        return _cuda
    if not os.path.exists(CUDA_LIB_PATH):
        raise NativeLibraryError(
            f"{CUDA_LIB_PATH} is missing: run `python build.py` (nvcc, sm_100a) before using GPU ops")
    L = C.CDLL(CUDA_LIB_PATH)
    vp, i32, i64, u32, f32 = C.c_void_p, C.c_int, C.c_longlong, C.c_uint32, C.c_float
    L.dl_last_error.restype = C.c_char_p
    L.dl_device_sm_count.restype = i32
    L.dl_q40t_bytes.restype = i64
    L.dl_q40t_bytes.argtypes = [i32, i32]
    L.dl_q40_repack.argtypes = [vp, i32, i32, vp, i32, i32, i32, vp]
    L.dl_q40t_dequant.argtypes = [vp, i32, i32, vp, vp]
    L.dl_reduce_norm_quant.argtypes = [vp, vp, vp, i32, i32, vp, vp, u32, u32, i32, vp, f32, i32, i32, i32,
                                       vp, vp, vp, i32, vp]
    L.dl_quant_q80.argtypes = [vp, i32, i32, i32, vp, vp, vp]
    L.dl_embedding.argtypes = [vp, vp, vp, i32, i32, vp, vp]
    L.dl_sample_rows.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp, vp]
    L.dl_gemm_f32.argtypes = [vp, i32, i32, i64, vp, i32, i32, vp, i32, i32, vp]
    L.dl_gemv_q40.argtypes = [vp, i32, i32, vp, i32, i32, i32, i32, C.POINTER(DlNormProlog), C.POINTER(DlEpilogue), i32, vp]
    L.dl_gemv_q40_smem_bytes.restype = i64
    L.dl_gemv_q40_smem_bytes.argtypes = [i32, i32, i32]
    L.dl_gemm_q40_tc.argtypes = [vp, i32, i32, vp, i32, i32, i32, C.POINTER(DlEpilogue), i32, vp]
    L.dl_wh_bytes.restype = i64
    L.dl_wh_bytes.argtypes = [i32, i32]
    L.dl_q40t_to_wh.argtypes = [vp, i32, i32, vp, vp]
    L.dl_gemm_wh_tc.argtypes = [vp, i32, i32, vp, i32, i32, C.POINTER(DlEpilogue), i32, i32, vp]
    L.dl_attention_scratch_bytes.restype = i64
    L.dl_attention_scratch_bytes.argtypes = [i32, i32, i32, i32]
    L.dl_attention.argtypes = [vp, i32, i32, i32, i32, vp, vp, i32, vp, vp, vp, i32, i32, i32, vp, vp, vp,
                               i32, vp, vp, i32, vp]
    L.dl_attention_prefill.argtypes = [vp, i32, i32, i32, i32, vp, vp, vp, vp, vp, i32, i32, vp, vp, vp, i32, vp, i32, vp]
    L.dl_attention_prefill_tc.argtypes = [vp, i32, i32, i32, i32, vp, vp, vp, vp, vp, i32, i32, vp, vp, vp, i32, vp, i32, i32, vp]
    L.dl_argmax.argtypes = [vp, i32, i32, i32, vp, vp, i32, vp, u32, u32, vp, vp, i32, vp, i32, vp]
    L.dl_argmax_scratch_bytes.restype = i64
    L.dl_argmax_scratch_bytes.argtypes = [i32]
    L.dl_wait_flags.argtypes = [vp, i32, vp, u32, u32, vp]
    L.dl_ipc_alloc.argtypes = [C.POINTER(vp), i64, vp]
    L.dl_ipc_open.argtypes = [vp, C.POINTER(vp)]
    L.dl_ipc_close.argtypes = [vp]
    L.dl_ipc_free.argtypes = [vp]
    L.dl_peer_barrier.argtypes = [C.POINTER(vp), vp, i32, i32, u32, vp]
    L.dl_f32_rope_kv.argtypes = [vp, i32, i32, i32, i32, i32, vp, vp, vp, vp, i32, vp, vp, vp, i32, vp]
    L.dl_f32_silu_mul.argtypes = [vp, vp, vp, i64, i32, vp]
    L.dl_f32_push.argtypes = [vp, i32, i32, i32, C.POINTER(vp), i32, i64, i32, C.POINTER(vp), vp, u32, u32, vp, vp]
    for name in ("dl_f32_rope_kv", "dl_f32_silu_mul", "dl_f32_push"):
        getattr(L, name).restype = i32
    L.dl_decode_mega_supported.argtypes = [C.POINTER(DlDecodeArgs), i32]
    L.dl_decode_mega_supported.restype = i32
    L.dl_decode_mega.argtypes = [C.POINTER(DlDecodeArgs), i32, vp]
    L.dl_decode_mega.restype = i32
    L.dl_decode_mega_plan.argtypes = [i32, i32, i32, i32, i32, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
    L.dl_decode_mega_plan.restype = i32
    for name in ("dl_q40_repack", "dl_q40t_dequant", "dl_reduce_norm_quant", "dl_quant_q80", "dl_embedding",
                 "dl_sample_rows", "dl_gemm_f32", "dl_gemv_q40", "dl_gemm_q40_tc", "dl_q40t_to_wh", "dl_gemm_wh_tc", "dl_attention", "dl_attention_prefill", "dl_attention_prefill_tc", "dl_argmax", "dl_ipc_alloc", "dl_ipc_open",
                 "dl_ipc_close", "dl_ipc_free", "dl_peer_barrier", "dl_wait_flags"):
        getattr(L, name).restype = i32
    _cuda = L
    return L


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = cuda_lib().dl_last_error().decode("utf-8", "replace")
        raise NativeLibraryError(f"{what} failed ({rc}): {msg}")


# kernels launched by this process (bench.py reports it as gpu_launches)
launch_count = 0


def count_launch(n: int = 1):
    global launch_count
    launch_count += n
