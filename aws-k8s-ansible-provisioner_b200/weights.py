"""Synthetic checkpoints for the benchmark and the server: there is no network for HF downloads
(the reference pulls Qwen/Qwen3-0.6B with an HF token, llm-d-deploy.yaml:121-132,184), so weights
are random-init bf16 with HF's default init statistics (N(0, 0.02) for Linear/Embedding, ones for
norms), under HF's tensor names."""
from __future__ import annotations

import json
import struct

import torch

from .engine import EngineConfig


def tensor_shapes(cfg: EngineConfig) -> dict[str, tuple[int, ...]]:
    q, kv = cfg.num_q_heads * cfg.head_dim, cfg.num_kv_heads * cfg.head_dim
    s: dict[str, tuple[int, ...]] = {"model.embed_tokens.weight": (cfg.vocab, cfg.hidden)}
    for i in range(cfg.num_layers):
        p = f"model.layers.{i}."
        s[p + "input_layernorm.weight"] = (cfg.hidden,)
        s[p + "self_attn.q_proj.weight"] = (q, cfg.hidden)
        s[p + "self_attn.k_proj.weight"] = (kv, cfg.hidden)
        s[p + "self_attn.v_proj.weight"] = (kv, cfg.hidden)
        s[p + "self_attn.o_proj.weight"] = (cfg.hidden, q)
        s[p + "self_attn.q_norm.weight"] = (cfg.head_dim,)
        s[p + "self_attn.k_norm.weight"] = (cfg.head_dim,)
        s[p + "post_attention_layernorm.weight"] = (cfg.hidden,)
        s[p + "mlp.gate_proj.weight"] = (cfg.intermediate, cfg.hidden)
        s[p + "mlp.up_proj.weight"] = (cfg.intermediate, cfg.hidden)
        s[p + "mlp.down_proj.weight"] = (cfg.hidden, cfg.intermediate)
    s["model.norm.weight"] = (cfg.hidden,)
    if not cfg.tie_embeddings:
        s["lm_head.weight"] = (cfg.vocab, cfg.hidden)
    return s


def random_state_dict(cfg: EngineConfig, seed: int = 0, std: float = 0.02) -> dict[str, torch.Tensor]:
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, shape in tensor_shapes(cfg).items():
        if "norm" in name:
            sd[name] = torch.ones(shape, dtype=torch.bfloat16)
        else:
            sd[name] = (torch.randn(shape, generator=g, dtype=torch.float32) * std).to(torch.bfloat16)
    return sd


def write_safetensors(sd: dict[str, torch.Tensor], path: str) -> None:
    """Minimal safetensors writer so the C++ server can mmap a synthetic checkpoint: every tensor is stored in its
    own dtype (bf16, fp16 or fp32 -- the three the engine's loader reads)."""
    names = {torch.bfloat16: "BF16", torch.float16: "F16", torch.float32: "F32"}
    header = {}
    off = 0
    for name, t in sd.items():
        n = t.numel() * t.element_size()
        header[name] = {"dtype": names[t.dtype], "shape": list(t.shape), "data_offsets": [off, off + n]}
        off += n
    hb = json.dumps(header, separators=(",", ":")).encode()
    hb += b" " * ((8 - len(hb) % 8) % 8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(hb)))
        f.write(hb)
        for t in sd.values():
            t = t.contiguous()
            f.write(t.view(torch.int16 if t.element_size() == 2 else torch.int32).numpy().tobytes())
