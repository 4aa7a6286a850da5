"""Shared helpers for the parity tests (test infrastructure)."""
from __future__ import annotations

import ctypes as C
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG_NAME = "aws-k8s-ansible-provisioner_b200"


def pkg():
    return importlib.import_module(PKG_NAME)


def bf16_ulp_diff(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Distance in bf16 ulps between two bf16 tensors (monotone integer mapping of the bit patterns)."""
    def key(t):
        i = t.contiguous().view(torch.int16).to(torch.int32) & 0xFFFF
        neg = i >= 0x8000
        return torch.where(neg, 0x8000 - i, i)
    return (key(a.cpu()) - key(b.cpu())).abs()


def assert_bf16_close(got: torch.Tensor, ref: torch.Tensor, max_ulp: int, max_mismatch_frac: float, what: str,
                      abs_floor: float = 0.0):
    """Parity bar for bf16 outputs: every element within `max_ulp` bf16 ulps of the oracle (or within
    abs_floor absolutely, for values near zero where an ulp is meaningless), and at most
    `max_mismatch_frac` of the elements differing at all."""
    got, ref = got.cpu(), ref.cpu()
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    assert torch.isfinite(got.float()).all(), what + ": non-finite output"
    d = bf16_ulp_diff(got, ref)
    absd = (got.float() - ref.float()).abs()
    bad = (d > max_ulp) & (absd > abs_floor)
    frac = (d > 0).float().mean().item()
    assert not bad.any(), "%s: %d elements beyond %d ulp (max %d ulp, max abs %.3g)" % (
        what, int(bad.sum()), max_ulp, int(d.max()), float(absd.max()))
    assert frac <= max_mismatch_frac, "%s: %.4f of elements differ (allowed %.4f)" % (what, frac, max_mismatch_frac)
    return frac


def rel_err(got: torch.Tensor, ref: torch.Tensor) -> float:
    """max |got-ref| / max |ref|  -- the 'within 1e-2 relative' metric of BASELINE.json's north_star."""
    g, r = got.float().cpu(), ref.float().cpu()
    return ((g - r).abs().max() / r.abs().max().clamp_min(1e-30)).item()


def rel_fro(got: torch.Tensor, ref: torch.Tensor) -> float:
    """||got-ref||_F / ||ref||_F."""
    g, r = got.float().cpu(), ref.float().cpu()
    return ((g - r).norm() / r.norm().clamp_min(1e-30)).item()


def ptr(t: torch.Tensor):
    return C.c_void_p(t.data_ptr())


def cuda_stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def engine_config_from_arch(p, arch, **kw):
    cfg = p.EngineConfig(num_layers=arch.num_layers, hidden=arch.hidden, num_q_heads=arch.num_q_heads,
                         num_kv_heads=arch.num_kv_heads, head_dim=arch.head_dim, intermediate=arch.intermediate,
                         vocab=arch.vocab, tie_embeddings=arch.tie_embeddings, rms_eps=arch.rms_eps,
                         rope_theta=arch.rope_theta, arch=getattr(arch, "family", "qwen3"),
                         rotary_dim=getattr(arch, "rotary_dim", 0) or 0,
                         max_positions=getattr(arch, "max_position", 0) if getattr(arch, "family", "") == "opt" else 0,
                         sliding_window=getattr(arch, "sliding_window", 0) or 0)
    for k, v in kw.items():
        setattr(cfg, k, v)
    return cfg


def kv_swizzle(pages: torch.Tensor) -> torch.Tensor:
    """Logical [..., 16 tokens, 128 dims] -> the stored page layout of include/b2_engine.h: the 16-byte
    chunk c of token row r sits at chunk position c ^ (r & 7).  The permutation is its own inverse."""
    shp = pages.shape
    x = pages.reshape(-1, 16, 16, 8)  # [slice][row][chunk][8 dims]
    r = torch.arange(16)[:, None]
    c = torch.arange(16)[None, :]
    src = c ^ (r & 7)                  # stored[r][c] = logical[r][c ^ (r & 7)]
    out = torch.gather(x, 2, src[None, :, :, None].expand(x.shape[0], 16, 16, 8))
    return out.reshape(shp)


def record_parity(key: str, value) -> None:
    """Keeps the numbers a parity test printed (leading-match counts, gaps at the first mismatch, errors) in a JSON
    file that outlives `pytest -q`: gpurun_out/parity.json on the GPU box (merged back by gpurun; copied to
    profiles/ when committed as evidence)."""
    import json
    d = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(d, exist_ok=True)
        path = os.path.join(d, "parity.json")
        cur = {}
        if os.path.exists(path):
            try:
                cur = json.load(open(path))
            except ValueError:
                cur = {}
        cur[key] = value
        json.dump(cur, open(path, "w"), indent=1, sort_keys=True)
    except OSError:
        pass


def bf16_ulp_of(x: float) -> float:
    """Size of one bf16 ulp at magnitude |x| (8 significand bits)."""
    import math
    x = abs(float(x))
    if x == 0.0:
        return 2.0 ** -133
    return 2.0 ** (math.floor(math.log2(x)) - 7)
