"""ctypes binding of include/b2_engine.h (the same stub INTEGRATION.md shows a maintainer)."""
from __future__ import annotations

import ctypes as C
import os

# B2_LIB_PATH: an alternative build of the same library (A/B measurements of one source change); never a fallback
LIB_PATH = os.environ.get("B2_LIB_PATH") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb2engine.so")
ABI_VERSION = 2
ARCH_QWEN3, ARCH_PHI3, ARCH_OPT = 0, 1, 2
NUM_KERNEL_CLASSES = 12


class B2Error(RuntimeError):
    pass


class b2_config(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("num_layers", C.c_int32), ("hidden", C.c_int32), ("num_q_heads", C.c_int32),
        ("num_kv_heads", C.c_int32), ("head_dim", C.c_int32), ("intermediate", C.c_int32), ("vocab", C.c_int32),
        ("tie_embeddings", C.c_int32), ("rms_eps", C.c_float), ("rope_theta", C.c_float), ("max_model_len", C.c_int32),
        ("block_size", C.c_int32), ("num_blocks", C.c_int32), ("gpu_mem_fraction", C.c_float), ("max_seqs", C.c_int32),
        ("max_batch_tokens", C.c_int32), ("use_cuda_graphs", C.c_int32), ("rope_inv_freq", C.POINTER(C.c_float)),
        ("arch", C.c_int32), ("rotary_dim", C.c_int32), ("max_positions", C.c_int32), ("sliding_window", C.c_int32),
    ]


class b2_sampling(C.Structure):
    _fields_ = [("temperature", C.c_float), ("top_k", C.c_int32), ("top_p", C.c_float), ("seed", C.c_uint64)]


_P = C.c_void_p
_I = C.c_int
_I32P = C.POINTER(C.c_int32)

# name -> (restype, argtypes); every symbol include/b2_engine.h declares
SIGNATURES = {
    "b2_last_error": (C.c_char_p, []),
    "b2_abi_version": (_I, []),
    "b2_engine_create": (_I, [C.POINTER(b2_config), C.c_char_p, _I, C.POINTER(_P)]),
    "b2_engine_load_tensor": (_I, [_P, C.c_char_p, _P, C.c_size_t]),
    "b2_engine_load_safetensors": (_I, [_P, C.c_char_p]),
    "b2_engine_finalize": (_I, [_P]),
    "b2_engine_free": (None, [_P]),
    "b2_engine_info": (_I, [_P, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "b2_seq_create": (_I, [_P, C.c_int32, _I32P]),
    "b2_seq_free": (_I, [_P, C.c_int32]),
    "b2_seq_len": (_I, [_P, C.c_int32]),
    "b2_seq_truncate": (_I, [_P, C.c_int32, C.c_int32]),
    "b2_prefill": (_I, [_P, _I32P, _I32P, C.c_int32, _I32P, _I32P]),
    "b2_decode_step": (_I, [_P, C.c_int32, _I32P, _I32P, _I32P]),
    "b2_decode_run": (_I, [_P, C.c_int32, _I32P, C.c_int32, _I32P, C.POINTER(C.c_float)]),
    "b2_get_logits": (_I, [_P, C.c_int32, C.POINTER(C.c_float)]),
    "b2_set_debug": (_I, [_P, C.c_int32]),
    "b2_get_hidden": (_I, [_P, C.c_int32, C.c_int32, C.POINTER(C.c_uint16)]),
    "b2_sample": (_I, [_P, C.c_int32, C.POINTER(b2_sampling), _I32P]),
    "b2_sample_rows": (_I, [_P, C.c_int32, _I32P, C.POINTER(b2_sampling), _I32P]),
    "b2_set_next_tokens": (_I, [_P, C.c_int32, _I32P, _I32P]),
    "b2_decode_profile": (_I, [_P, C.c_int32, _I32P, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_int64)]),
    "b2_attn_profile": (_I, [_P, C.c_int32, _I32P, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_int64)]),
    "b2_launch_count": (C.c_int64, [_P]),
    "b2_set_tuning": (_I, [_P, C.c_int32, C.c_int64]),
    "b2_prefix_cache_stats": (_I, [_P, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "b2k_add_rmsnorm": (_I, [_P, _P, _I, _I, _P, _P, _I, _I, C.c_float, _P]),
    "b2k_qknorm_rope_kv": (_I, [_P, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, C.c_float, _P]),
    "b2k_silu_mul": (_I, [_P, _I, _I, _P, _I, _I, _P]),
    "b2k_linear": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P]),
    "b2k_gateup_swiglu": (_I, [_P, _P, _P, _I, _I, _I, _P]),
    "b2k_qkv_rope": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, C.c_float, _P]),
    "b2k_linear_resid": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "b2k_linear_add": (_I, [_P, _P, _P, _I, _I, _I, _P]),
    "b2k_linear_multicast": (_I, [_P, _P, _P, _I, _I, _I, _P]),
    "b2k_gateup_swiglu_norm": (_I, [_P, _P, _P, _P, _I, _I, C.c_float, _P, _I, _I, _I, _P]),
    "b2k_qkv_rope_norm": (_I, [_P, _P, _P, _P, _I, _I, C.c_float, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, C.c_float, _P]),
    "b2k_debug_step_trace": (_I, [_P]),
    "b2k_debug_gemm_trace": (_I, [_P]),
    "b2k_debug_chain_trace": (_I, [_P]),
    "b2k_debug_fused_trace": (_I, [_P]),
    "b2k_lmhead_argmax": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "b2k_attn_decode": (_I, [_P, _P, _P, _P, _I, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "b2k_attn_decode_scratch_floats": (C.c_size_t, [_I, _I, _I, _I]),
    "b2k_attn_decode_noalloc": (_I, [_P, _P, _P, _P, _I, _P, _P, _I, _I, _I, _I, _I, _P]),
    "b2k_attn_prefill": (_I, [_P, _P, _P, _P, _I, _I32P, _I32P, _I32P, _I32P, _I, _P, _I, _I, _I, _I, _P]),
}

_lib = None


def load_library(path: str | None = None):
    """Loads libb2engine.so and types every entry point.  Fails loudly when it is not built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise B2Error(
            "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
            "There is no CPU fallback for this engine." % p)
    lib = C.CDLL(p)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    if lib.b2_abi_version() != ABI_VERSION:
        raise B2Error("ABI mismatch: library %d, binding %d" % (lib.b2_abi_version(), ABI_VERSION))
    if path is None:
        _lib = lib
    return lib


def check(rc: int, lib=None) -> None:
    if rc != 0:
        lib = lib or load_library()
        raise B2Error("b2 error %d: %s" % (rc, (lib.b2_last_error() or b"").decode("utf-8", "replace")))
