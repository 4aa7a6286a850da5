"""CPU oracle for the OPT variant of the hot path -- TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this package; nothing the
product ships goes through it.

Graded: HF transformers' OPTForCausalLM on CPU (transformers 5.5.0, models/opt/modeling_opt.py,
"HF:" below).  `RestatedOPT` restates that forward in plain torch:

  * h0 = embed_tokens[id] + embed_positions[position + 2]                       HF:...:45-70,363-368
  * per layer (do_layer_norm_before, the 125m arrangement)                      HF:...:203-254
        x  = LayerNorm(h);  q = (Wq x + bq) * head_dim^-0.5;  k, v = W x + b     HF:...:151-157
        h  = h + (Wo attention(q, k, v; scale 1.0) + bo)                        HF:...:167-181,224
        x2 = LayerNorm(h);  h = h + (W2 relu(W1 x2 + b1) + b2)                   HF:...:239-246
  * final LayerNorm, logits = h E^T (head tied to embed_tokens)                  HF:...:386-391

Parity pin: tests/test_oracle.py checks the restatement bit-for-bit against HF's own outputs stored in
tests/golden/opt_tiny.npz (oracle/make_golden.py).  Not restated and not built: the 350m arrangement
(do_layer_norm_before=False) and word_embed_proj_dim != hidden_size (project_in/project_out).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn.functional as F

from .qwen3_oracle import ArchConfig, KVState, Restated, attention_sdpa


@dataclass
class OptArch(ArchConfig):
    family: str = "opt"
    ln_eps: float = 1e-5

    def kv_bytes_per_token(self) -> int:
        return 2 * self.num_layers * self.kv_size * 2


# facebook/opt-125m shapes (BASELINE.json config 3)
OPT_125M = OptArch(name="facebook/opt-125m", num_layers=12, hidden=768, num_q_heads=12, num_kv_heads=12, head_dim=64,
                   intermediate=3072, vocab=50272, rms_eps=1e-5, max_position=2048, tie_embeddings=True)
# same structure, CPU-cheap
OPT_TINY = OptArch(name="opt-tiny-test", num_layers=2, hidden=256, num_q_heads=4, num_kv_heads=4, head_dim=64,
                   intermediate=512, vocab=1024, rms_eps=1e-5, max_position=512, tie_embeddings=True)
# two layers at OPT-125m's widths (hidden 768, 12 heads of 64, ffn 3072, vocab 50272)
OPT_125M_2L = OptArch(name="opt-125m-2-layers", num_layers=2, hidden=768, num_q_heads=12, num_kv_heads=12, head_dim=64,
                      intermediate=3072, vocab=50272, rms_eps=1e-5, max_position=2048, tie_embeddings=True)


def hf_opt_config(arch: OptArch):
    from transformers import OPTConfig

    return OPTConfig(
        vocab_size=arch.vocab, hidden_size=arch.hidden, num_hidden_layers=arch.num_layers, ffn_dim=arch.intermediate,
        max_position_embeddings=arch.max_position, do_layer_norm_before=True, word_embed_proj_dim=arch.hidden,
        num_attention_heads=arch.num_q_heads, activation_function="relu", dropout=0.0, attention_dropout=0.0, layerdrop=0.0,
        enable_bias=True, layer_norm_elementwise_affine=True, pad_token_id=1, bos_token_id=2, eos_token_id=2,
        tie_word_embeddings=True,
    )


def hf_opt_model(arch: OptArch = OPT_TINY, seed: int = 0, attn: str = "sdpa"):
    """HF OPTForCausalLM, default init under torch.manual_seed(seed), bf16, eval, CPU."""
    from transformers import OPTForCausalLM

    cfg = hf_opt_config(arch)
    cfg._attn_implementation = attn
    torch.manual_seed(seed)
    return OPTForCausalLM(cfg).to(torch.bfloat16).eval()


def perturb_opt_weights(sd: dict, seed: int = 7) -> None:
    """HF initialises LayerNorm gains to 1 and every bias to 0, which would hide a kernel that forgot
    one of them; tests and fixtures make them non-trivial."""
    g = torch.Generator().manual_seed(seed)
    for k, v in sd.items():
        if k.endswith("layer_norm.weight"):
            sd[k] = (1.0 + 0.25 * torch.randn(v.shape, generator=g)).to(torch.bfloat16)
        elif k.endswith(".bias"):
            sd[k] = (0.1 * torch.randn(v.shape, generator=g)).to(torch.bfloat16)


class RestatedOPT(Restated):
    """Plain-torch OPT forward over an HF-named state dict; greedy() comes from Restated."""

    def __init__(self, arch: OptArch, sd, attention=attention_sdpa, linear=F.linear):
        super().__init__(arch, sd, attention)
        self.lin = linear

    def w(self, name: str) -> torch.Tensor:
        return self.sd["model.decoder." + name]

    def layer(self, i, h, positions, kv: KVState, trace):
        a = self.arch
        p = f"layers.{i}."
        B, T, H = h.shape
        residual = h
        x = F.layer_norm(h, (H,), self.w(p + "self_attn_layer_norm.weight"), self.w(p + "self_attn_layer_norm.bias"), a.ln_eps)
        q = self.lin(x, self.w(p + "self_attn.q_proj.weight"), self.w(p + "self_attn.q_proj.bias")) * (a.head_dim ** -0.5)
        k = self.lin(x, self.w(p + "self_attn.k_proj.weight"), self.w(p + "self_attn.k_proj.bias"))
        v = self.lin(x, self.w(p + "self_attn.v_proj.weight"), self.w(p + "self_attn.v_proj.bias"))
        q = q.view(B, T, a.num_q_heads, a.head_dim).transpose(1, 2)
        k = k.view(B, T, a.num_kv_heads, a.head_dim).transpose(1, 2)
        v = v.view(B, T, a.num_kv_heads, a.head_dim).transpose(1, 2)
        if len(kv.k) <= i:
            kv.k.append(k)
            kv.v.append(v)
        else:
            kv.k[i] = torch.cat((kv.k[i], k), dim=2)
            kv.v[i] = torch.cat((kv.v[i], v), dim=2)
        ao = self.attention(q, kv.k[i], kv.v[i], 1.0, T > 1)
        ao = ao.transpose(1, 2).contiguous().reshape(B, T, a.q_size)
        h = residual + self.lin(ao, self.w(p + "self_attn.out_proj.weight"), self.w(p + "self_attn.out_proj.bias"))
        residual = h
        x2 = F.layer_norm(h, (H,), self.w(p + "final_layer_norm.weight"), self.w(p + "final_layer_norm.bias"), a.ln_eps)
        f = F.relu(self.lin(x2, self.w(p + "fc1.weight"), self.w(p + "fc1.bias")))
        out = residual + self.lin(f, self.w(p + "fc2.weight"), self.w(p + "fc2.bias"))
        if trace is not None:
            trace.attn_in, trace.q, trace.k, trace.v = x, q, k, v
            trace.attn_out, trace.h_mid, trace.h_out = ao, h, out
        return out

    @torch.no_grad()
    def forward(self, ids, positions, kv: KVState, traces=None, last_only: bool = False):
        a = self.arch
        h = F.embedding(ids, self.w("embed_tokens.weight")) + F.embedding(positions + 2, self.w("embed_positions.weight"))
        for i in range(a.num_layers):
            tr = None
            if traces is not None:
                from .qwen3_oracle import LayerTrace
                tr = LayerTrace()
                traces.append(tr)
            h = self.layer(i, h, positions, kv, tr)
        H = h.shape[-1]
        h = F.layer_norm(h, (H,), self.w("final_layer_norm.weight"), self.w("final_layer_norm.bias"), a.ln_eps)
        hl = h[:, -1:, :] if last_only else h
        return F.linear(hl, self.w("embed_tokens.weight")), h


def linear_fp32_order(x, w, b=None):
    """nn.Linear's contract (fp32 accumulate, bias in fp32, one rounding) with another fp32 summation
    order -- the distance to F.linear is the bf16 noise floor tests compare against."""
    y = x.float() @ w.float().t()
    if b is not None:
        y = y + b.float()
    return y.to(x.dtype)
