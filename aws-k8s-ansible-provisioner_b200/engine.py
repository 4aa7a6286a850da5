"""Host-side mirror of the engine's C ABI (include/b2_engine.h) for tests, bench.py and the
Python tooling.  Every method is one C call; there is no arithmetic here and no fallback.

The names follow the serving domain of the reference deployment (sequences, pages, prefill,
decode step; llm-d-deploy.yaml:118, llm-d-test.yaml:61-73), not a training framework's.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import capi
from .capi import B2Error, b2_config, b2_sampling, check, load_library

KERNEL_CLASSES = ["embed", "norm", "gemm_qkv", "rope_kv", "attn_decode", "chain", "gemm_o", "gemm_gateup", "act",
                  "gemm_down", "lmhead", "misc"]


@dataclass
class EngineConfig:
    num_layers: int = 28
    hidden: int = 1024
    num_q_heads: int = 16
    num_kv_heads: int = 8
    head_dim: int = 128
    intermediate: int = 3072
    vocab: int = 151936
    tie_embeddings: bool = True
    rms_eps: float = 1e-6
    rope_theta: float = 1e6
    max_model_len: int = 4096
    block_size: int = 16
    num_blocks: int = 0
    gpu_mem_fraction: float = 0.9
    max_seqs: int = 64
    max_batch_tokens: int = 8192
    use_cuda_graphs: bool = True
    arch: str = "qwen3"          # "qwen3" | "phi3" | "opt" (include/b2_engine.h, enum b2_arch)
    rotary_dim: int = 0          # 0 = head_dim
    max_positions: int = 0       # OPT: learned position table rows - 2 (0 = max_model_len)
    sliding_window: int = 0      # Phi-3: attend to the last N keys only (0 = all)

    def kv_bytes_per_token(self) -> int:
        return 2 * self.num_layers * self.num_kv_heads * self.head_dim * 2

    def weight_bytes(self) -> int:
        q, kv = self.num_q_heads * self.head_dim, self.num_kv_heads * self.head_dim
        per_layer = (q + 2 * kv) * self.hidden + self.hidden * q + 3 * self.intermediate * self.hidden + 2 * self.hidden + 2 * self.head_dim
        total = self.num_layers * per_layer + self.vocab * self.hidden * (1 if self.tie_embeddings else 2) + self.hidden
        return 2 * total


QWEN3_0_6B = EngineConfig()


def _i32(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32))


def _p32(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


class Engine:
    """One engine per GPU (request-level data parallelism: one replica per device, no collective)."""

    def __init__(self, cfg: EngineConfig, device: int = 0, safetensors_path: str | None = None, rope_inv_freq=None):
        self.lib = load_library()
        self.cfg = cfg
        c = b2_config()
        c.abi_version = capi.ABI_VERSION
        for f in ("num_layers", "hidden", "num_q_heads", "num_kv_heads", "head_dim", "intermediate", "vocab", "max_model_len",
                  "block_size", "num_blocks", "max_seqs", "max_batch_tokens"):
            setattr(c, f, int(getattr(cfg, f)))
        c.tie_embeddings = int(cfg.tie_embeddings)
        c.use_cuda_graphs = int(cfg.use_cuda_graphs)
        c.rms_eps = cfg.rms_eps
        c.rope_theta = cfg.rope_theta
        c.gpu_mem_fraction = cfg.gpu_mem_fraction
        c.arch = {"qwen3": capi.ARCH_QWEN3, "phi3": capi.ARCH_PHI3, "opt": capi.ARCH_OPT}[cfg.arch]
        c.rotary_dim = int(cfg.rotary_dim)
        c.max_positions = int(cfg.max_positions)
        c.sliding_window = int(cfg.sliding_window)
        self._inv = None
        if rope_inv_freq is not None:
            self._inv = np.ascontiguousarray(np.asarray(rope_inv_freq, dtype=np.float32))
            assert self._inv.shape == ((cfg.rotary_dim or cfg.head_dim) // 2,)
            c.rope_inv_freq = self._inv.ctypes.data_as(C.POINTER(C.c_float))
        h = C.c_void_p()
        path = safetensors_path.encode() if safetensors_path else None
        check(self.lib.b2_engine_create(C.byref(c), path, device, C.byref(h)), self.lib)
        self.h = h
        self.device = device

    # ---- weights -----------------------------------------------------------------------
    def load_safetensors(self, paths) -> None:
        """Loads one or more safetensors files (the shards of a checkpoint) and finalises the weights."""
        for p in ([paths] if isinstance(paths, str) else list(paths)):
            check(self.lib.b2_engine_load_safetensors(self.h, p.encode()), self.lib)
        check(self.lib.b2_engine_finalize(self.h), self.lib)

    def load_state_dict(self, sd) -> None:
        """Uploads an HF-named state dict of torch bf16 CPU tensors (no file round trip)."""
        import torch

        for name, t in sd.items():
            if "rotary_emb" in name:
                continue
            t = t.detach().to(torch.bfloat16).contiguous().cpu()
            check(self.lib.b2_engine_load_tensor(self.h, name.encode(), C.c_void_p(t.data_ptr()), t.numel() * 2), self.lib)
        check(self.lib.b2_engine_finalize(self.h), self.lib)

    # ---- sequences ----------------------------------------------------------------------
    def seq_create(self, reserve_tokens: int) -> int:
        sid = C.c_int32(-1)
        check(self.lib.b2_seq_create(self.h, int(reserve_tokens), C.byref(sid)), self.lib)
        return sid.value

    def seq_free(self, seq_id: int) -> None:
        check(self.lib.b2_seq_free(self.h, int(seq_id)), self.lib)

    def seq_len(self, seq_id: int) -> int:
        return self.lib.b2_seq_len(self.h, int(seq_id))

    def seq_truncate(self, seq_id: int, new_len: int) -> None:
        check(self.lib.b2_seq_truncate(self.h, int(seq_id), int(new_len)), self.lib)

    def info(self) -> dict:
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        check(self.lib.b2_engine_info(self.h, C.byref(a), C.byref(b), C.byref(c)), self.lib)
        return {"num_blocks": a.value, "free_blocks": b.value, "weight_bytes": c.value}

    # ---- forward passes ------------------------------------------------------------------
    def prefill(self, seq_ids, prompts) -> np.ndarray:
        """prompts: list of int sequences (ragged).  Returns the greedy next token per sequence."""
        ids = _i32(np.concatenate([np.asarray(p, dtype=np.int32) for p in prompts]))
        offs = _i32(np.cumsum([0] + [len(p) for p in prompts]))
        sids = _i32(seq_ids)
        out = np.zeros(len(prompts), dtype=np.int32)
        check(self.lib.b2_prefill(self.h, _p32(ids), _p32(offs), len(prompts), _p32(sids), _p32(out)), self.lib)
        return out

    def decode_step(self, seq_ids, in_tokens=None, want_output: bool = True):
        sids = _i32(seq_ids)
        n = len(sids)
        tin = _i32(in_tokens) if in_tokens is not None else None
        out = np.zeros(n, dtype=np.int32) if want_output else None
        check(self.lib.b2_decode_step(self.h, n, _p32(sids), _p32(tin) if tin is not None else None,
                                      _p32(out) if out is not None else None), self.lib)
        return out

    def decode_run(self, seq_ids, steps: int, want_tokens: bool = True):
        """`steps` decode steps with token feedback on the device.  Returns (tokens[steps][n], ms)."""
        sids = _i32(seq_ids)
        n = len(sids)
        out = np.zeros((steps, n), dtype=np.int32) if want_tokens else None
        ms = C.c_float(0)
        check(self.lib.b2_decode_run(self.h, n, _p32(sids), int(steps), _p32(out) if out is not None else None, C.byref(ms)),
              self.lib)
        return out, ms.value

    def decode_profile(self, seq_ids, steps: int) -> dict:
        sids = _i32(seq_ids)
        ms = (C.c_float * capi.NUM_KERNEL_CLASSES)()
        cnt = (C.c_int64 * capi.NUM_KERNEL_CLASSES)()
        check(self.lib.b2_decode_profile(self.h, len(sids), _p32(sids), int(steps), ms, cnt), self.lib)
        return {KERNEL_CLASSES[i]: {"ms": float(ms[i]), "launches": int(cnt[i])} for i in range(capi.NUM_KERNEL_CLASSES)}

    def attn_profile(self, seq_ids, reps: int = 8):
        """Decode-attention pass of every layer back to back, `reps` times, between one event pair.
        Returns (ms per launch, K/V bytes one launch must read)."""
        sids = _i32(seq_ids)
        ms = C.c_float(0)
        nbytes = C.c_int64(0)
        check(self.lib.b2_attn_profile(self.h, len(sids), _p32(sids), int(reps), C.byref(ms), C.byref(nbytes)), self.lib)
        return ms.value, int(nbytes.value)

    def prefix_cache_stats(self) -> dict:
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        check(self.lib.b2_prefix_cache_stats(self.h, C.byref(a), C.byref(b), C.byref(c)), self.lib)
        return {"queried_tokens": a.value, "hit_tokens": b.value, "cached_pages": c.value}

    def set_tuning(self, skip_mask: int | None = None, two_lane_min_batch: int | None = None, prefix_cache: bool | None = None) -> None:
        """Run-time measurement knobs (include/b2_engine.h, b2_set_tuning); drops the captured decode graphs."""
        if skip_mask is not None:
            check(self.lib.b2_set_tuning(self.h, 0, int(skip_mask)), self.lib)
        if two_lane_min_batch is not None:
            check(self.lib.b2_set_tuning(self.h, 1, int(two_lane_min_batch)), self.lib)
        if prefix_cache is not None:
            check(self.lib.b2_set_tuning(self.h, 2, 1 if prefix_cache else 0), self.lib)

    def launch_count(self) -> int:
        return int(self.lib.b2_launch_count(self.h))

    # ---- inspection ----------------------------------------------------------------------
    def set_debug(self, hidden: bool = False, logits: bool = False) -> None:
        check(self.lib.b2_set_debug(self.h, (1 if hidden else 0) | (2 if logits else 0)), self.lib)

    def get_logits(self, nseq: int) -> np.ndarray:
        out = np.zeros((nseq, self.cfg.vocab), dtype=np.float32)
        check(self.lib.b2_get_logits(self.h, nseq, out.ctypes.data_as(C.POINTER(C.c_float))), self.lib)
        return out

    def get_hidden(self, layer: int, ntokens: int) -> np.ndarray:
        """bf16 bit patterns [ntokens][hidden] as uint16."""
        out = np.zeros((ntokens, self.cfg.hidden), dtype=np.uint16)
        check(self.lib.b2_get_hidden(self.h, layer, ntokens, out.ctypes.data_as(C.POINTER(C.c_uint16))), self.lib)
        return out

    def sample(self, nseq: int, temperature: float = 1.0, top_k: int = 0, top_p: float = 1.0, seed: int = 0) -> np.ndarray:
        p = b2_sampling(temperature, top_k, top_p, seed)
        out = np.zeros(nseq, dtype=np.int32)
        check(self.lib.b2_sample(self.h, nseq, C.byref(p), _p32(out)), self.lib)
        return out

    def set_next_tokens(self, seq_ids, tokens) -> None:
        sids, toks = _i32(seq_ids), _i32(tokens)
        check(self.lib.b2_set_next_tokens(self.h, len(sids), _p32(sids), _p32(toks)), self.lib)

    # ---- lifetime ------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "h", None):
            self.lib.b2_engine_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def bf16_bits_to_f32(bits: np.ndarray) -> np.ndarray:
    return (bits.astype(np.uint32) << 16).view(np.float32)
