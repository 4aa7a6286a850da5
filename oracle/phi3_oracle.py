"""CPU oracle for the Phi-3 variant of the hot path -- TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this package; nothing the
product ships goes through it.

What is graded is HF transformers' Phi3ForCausalLM on CPU (transformers 5.5.0,
models/phi3/modeling_phi3.py, cited below as "HF:").  `RestatedPhi3` is a plain-torch restatement of
that forward, reusing the Qwen3 restatement's pieces where the two models share arithmetic
(RMSNorm, rotate_half, sdpa attention, residual structure) and stating the Phi-3 differences:

  * fused `qkv_proj` [q + k + v rows] and `gate_up_proj` [gate; up]            HF:...:53-64,224,237-241
  * no per-head QK-norm
  * rotary over the first `rotary_dim` dims of each head, the rest passes      HF:...:178-205
  * `up * silu(gate)` (operands in the other order than Qwen3; same roundings)  HF:...:60-64
  * untied lm_head, RMSNorm eps 1e-5, one query head per kv head (MHA)

Parity pin: tests/test_oracle.py checks the restatement bit-for-bit against HF on the tiny
configuration below, and tests/golden/phi3_tiny.npz holds HF's own outputs (oracle/make_golden.py).
Sliding-window attention (Phi-3-mini-4k ships `sliding_window: 2047`; HF:...:265 and the sliding-window
mask of masking_utils) is restated by `attention_window` below and pinned by the "window" fixture, whose
window (24) is shorter than its contexts.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn.functional as F

from .qwen3_oracle import ArchConfig, KVState, Restated, attention_sdpa, rms_norm, rotate_half


@dataclass
class Phi3Arch(ArchConfig):
    family: str = "phi3"
    rotary_dim: int = 0  # 0 = head_dim (partial_rotary_factor 1.0, what Phi-3-mini uses)
    sliding_window: int = 0  # 0 = none; Phi-3-mini-4k ships 2047: a query sees its last 2047 keys, itself included

    @property
    def rot(self) -> int:
        return self.rotary_dim or self.head_dim


# microsoft/Phi-3-mini-4k-instruct shapes (BASELINE.json config 4)
PHI3_MINI = Phi3Arch(name="microsoft/Phi-3-mini-4k-instruct", num_layers=32, hidden=3072, num_q_heads=32, num_kv_heads=32,
                     head_dim=96, intermediate=8192, vocab=32064, rms_eps=1e-5, rope_theta=10000.0, max_position=4096,
                     tie_embeddings=False, sliding_window=2047)
# same structure (MHA, 96-wide heads, untied head), CPU-cheap
PHI3_TINY = Phi3Arch(name="phi3-tiny-test", num_layers=2, hidden=768, num_q_heads=8, num_kv_heads=8, head_dim=96,
                     intermediate=1024, vocab=1024, rms_eps=1e-5, rope_theta=10000.0, max_position=4096, tie_embeddings=False)
# two layers at Phi-3-mini's full widths: exercises the real GEMM / lm_head shapes (K = 3072 and 8192, N = 32064)
PHI3_MINI_2L = Phi3Arch(name="phi3-mini-2-layers", num_layers=2, hidden=3072, num_q_heads=32, num_kv_heads=32, head_dim=96,
                        intermediate=8192, vocab=32064, rms_eps=1e-5, rope_theta=10000.0, max_position=4096,
                        tie_embeddings=False)
# a window shorter than the test contexts, so that it actually cuts
PHI3_TINY_WINDOW = Phi3Arch(name="phi3-tiny-sliding-window", num_layers=2, hidden=768, num_q_heads=8, num_kv_heads=8, head_dim=96,
                            intermediate=1024, vocab=1024, rms_eps=1e-5, rope_theta=10000.0, max_position=4096,
                            tie_embeddings=False, sliding_window=24)
# partial rotary (the "partial-rotary support" of SURVEY.md a14): 64 of 96 dims rotate
PHI3_TINY_PARTIAL = Phi3Arch(name="phi3-tiny-partial-rotary", num_layers=2, hidden=768, num_q_heads=8, num_kv_heads=8, head_dim=96,
                             intermediate=1024, vocab=1024, rms_eps=1e-5, rope_theta=10000.0, max_position=4096,
                             tie_embeddings=False, rotary_dim=64)


def phi3_inv_freq(arch: Phi3Arch) -> torch.Tensor:
    """HF:modeling_rope_utils default rope over dim = head_dim * partial_rotary_factor, fp32."""
    d = arch.rot
    return 1.0 / (arch.rope_theta ** (torch.arange(0, d, 2, dtype=torch.int64).to(dtype=torch.float) / d))


def hf_phi3_config(arch: Phi3Arch):
    from transformers import Phi3Config

    return Phi3Config(
        vocab_size=arch.vocab, hidden_size=arch.hidden, intermediate_size=arch.intermediate, num_hidden_layers=arch.num_layers,
        num_attention_heads=arch.num_q_heads, num_key_value_heads=arch.num_kv_heads, hidden_act="silu",
        max_position_embeddings=arch.max_position, original_max_position_embeddings=arch.max_position,
        rms_norm_eps=arch.rms_eps, rope_parameters={"rope_type": "default", "rope_theta": arch.rope_theta,
                                                    "partial_rotary_factor": arch.rot / arch.head_dim},
        tie_word_embeddings=arch.tie_embeddings, attention_dropout=0.0, resid_pdrop=0.0, embd_pdrop=0.0,
        sliding_window=arch.sliding_window or None, pad_token_id=0, bos_token_id=1, eos_token_id=2,
    )


def hf_phi3_model(arch: Phi3Arch = PHI3_TINY, seed: int = 0, attn: str = "sdpa"):
    """HF Phi3ForCausalLM, default init under torch.manual_seed(seed), bf16, eval, CPU; the rotary
    inv_freq buffer is restored to fp32 for the reason given in qwen3_oracle.hf_model."""
    from transformers import Phi3ForCausalLM

    cfg = hf_phi3_config(arch)
    cfg._attn_implementation = attn
    torch.manual_seed(seed)
    model = Phi3ForCausalLM(cfg).to(torch.bfloat16).eval()
    inv = phi3_inv_freq(arch)
    assert model.model.rotary_emb.inv_freq.shape == inv.shape, (model.model.rotary_emb.inv_freq.shape, inv.shape)
    model.model.rotary_emb.inv_freq = inv.clone()
    model.model.rotary_emb.original_inv_freq = inv.clone()
    return model


def phi3_cos_sin(arch: Phi3Arch, positions: torch.Tensor, dtype=torch.bfloat16):
    """HF:models/phi3/modeling_phi3.py Phi3RotaryEmbedding.forward: fp32 freqs, cat, cos/sin, cast."""
    freqs = positions.to(torch.float32)[..., None] * phi3_inv_freq(arch)
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().to(dtype), emb.sin().to(dtype)


def apply_partial_rope(q, k, cos, sin):
    """HF:models/phi3/modeling_phi3.py:178-205: only the first cos.shape[-1] dims rotate."""
    cos = cos.unsqueeze(1)
    sin = sin.unsqueeze(1)
    r = cos.shape[-1]
    q_rot, q_pass = q[..., :r], q[..., r:]
    k_rot, k_pass = k[..., :r], k[..., r:]
    q2 = torch.cat([(q_rot * cos) + (rotate_half(q_rot) * sin), q_pass], dim=-1)
    k2 = torch.cat([(k_rot * cos) + (rotate_half(k_rot) * sin), k_pass], dim=-1)
    return q2, k2


def attention_window(window: int):
    """softmax(q k^T * scaling + mask) v with the sliding-window causal mask HF builds for Phi-3: key j is visible
    to the query at position i iff j <= i and i - j < window.  Evaluated by torch's sdpa with an explicit boolean
    mask (what HF's sdpa path receives); q: [B, H, T, d], k/v: [B, Hk, S, d], the T queries are the LAST T positions."""
    def attn(q, k, v, scaling: float, causal: bool):
        T, S = q.shape[2], k.shape[2]
        qi = torch.arange(S - T, S)[:, None]
        kj = torch.arange(S)[None, :]
        mask = (kj <= qi) & (qi - kj < window)
        return F.scaled_dot_product_attention(q, k, v, attn_mask=mask[None, None], scale=scaling, enable_gqa=q.shape[1] != k.shape[1])
    return attn


def linear_fp32_order(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """y = x W^T with fp32 accumulation and one rounding, like F.linear on bf16, but summed by the fp32
    GEMM (different blocking, hence a different order of the same additions)."""
    return (x.float() @ w.float().t()).to(x.dtype)


class RestatedPhi3(Restated):
    """Plain-torch Phi-3 forward over an HF-named state dict (forward()/greedy() come from Restated)."""

    def __init__(self, arch: Phi3Arch, sd, attention=None, linear=F.linear):
        """`attention`: default sdpa, or the sliding-window form when the architecture has a window.  `linear`: the nn.Linear arithmetic.  F.linear is HF's (bf16 in, fp32 accumulate, one rounding);
        `linear_fp32_order` below is the same contract with another fp32 summation order -- tests use the
        distance between the two as the bf16 noise floor that no implementation can get under."""
        if attention is None:
            attention = attention_window(arch.sliding_window) if arch.sliding_window else attention_sdpa
        super().__init__(arch, sd, attention)
        self.lin = linear

    def layer(self, i, h, positions, kv: KVState, trace):
        """HF:models/phi3/modeling_phi3.py:295-336 Phi3DecoderLayer.forward, :226-272 Phi3Attention.forward."""
        a = self.arch
        p = f"model.layers.{i}."
        B, T, _ = h.shape
        residual = h
        x = rms_norm(h, self.w(p + "input_layernorm.weight"), a.rms_eps)
        qkv = self.lin(x, self.w(p + "self_attn.qkv_proj.weight"))
        qp = a.num_q_heads * a.head_dim
        kp = a.num_kv_heads * a.head_dim
        q = qkv[..., :qp].view(B, T, a.num_q_heads, a.head_dim).transpose(1, 2)
        k = qkv[..., qp:qp + kp].view(B, T, a.num_kv_heads, a.head_dim).transpose(1, 2)
        v = qkv[..., qp + kp:].view(B, T, a.num_kv_heads, a.head_dim).transpose(1, 2)
        cos, sin = phi3_cos_sin(a, positions)
        q, k = apply_partial_rope(q, k, cos, sin)
        if len(kv.k) <= i:
            kv.k.append(k)
            kv.v.append(v)
        else:
            kv.k[i] = torch.cat((kv.k[i], k), dim=2)
            kv.v[i] = torch.cat((kv.v[i], v), dim=2)
        ao = self.attention(q, kv.k[i], kv.v[i], a.head_dim ** -0.5, T > 1)
        ao = ao.transpose(1, 2).contiguous().reshape(B, T, a.q_size)
        h = residual + self.lin(ao, self.w(p + "self_attn.o_proj.weight"))
        residual = h
        x2 = rms_norm(h, self.w(p + "post_attention_layernorm.weight"), a.rms_eps)
        gu = self.lin(x2, self.w(p + "mlp.gate_up_proj.weight"))
        gate, up = gu.chunk(2, dim=-1)
        out = residual + self.lin(up * F.silu(gate), self.w(p + "mlp.down_proj.weight"))
        if trace is not None:
            trace.attn_in, trace.q, trace.k, trace.v = x, q, k, v
            trace.attn_out, trace.h_mid, trace.h_out = ao, h, out
        return out
