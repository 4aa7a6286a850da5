"""ORACLE tooling (test infrastructure only): generate the golden fixtures in tests/golden/.

Runs the real HF transformers Qwen3 (the arithmetic behind the model the reference deploys,
`llm-d-deploy.yaml:118`) on CPU and writes small JSON/npz fixtures:

  qwen3_0p6b_c1.json     config 1 of BASELINE.json: Qwen3-0.6B arch, HF default init under
                         torch.manual_seed(0), bf16, sdpa, batch 1, 128 random prompt ids
                         (generator seed 1) -> 128 greedy tokens; per-step top-1/top-2 logits
                         (tie-aware evidence) and per-layer digests of a teacher-forced pass.
  qwen3_tiny.npz         scaled-down arch (oracle.QWEN3_TINY) with perturbed norm gains: full
                         teacher-forced logits + per-layer hidden states + greedy ids.
  chat_templates.json    renderings of the reference's templates/*.yaml (`template.jinja`
                         key) with jinja2, for the server's own mini-Jinja to reproduce.

Usage:  python oracle/make_golden.py [--skip-full]
The reference templates are read from /root/reference at generation time only; the rendered
strings (not the template sources) are what is committed.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.qwen3_oracle import (  # noqa: E402
    QWEN3_0_6B, QWEN3_TINY, KVState, Restated, hf_model, hf_state_dict, make_prompts, perturb_norm_weights,
)

GOLD = os.path.join(ROOT, "tests", "golden")


def digest(t: torch.Tensor) -> dict:
    f = t.to(torch.float32).flatten()
    return {
        "sum": float(f.sum().item()),
        "abs_sum": float(f.abs().sum().item()),
        "l2": float(f.norm().item()),
        "head": [float(x) for x in f[:8].tolist()],
        "sha256_bf16": hashlib.sha256(t.to(torch.bfloat16).contiguous().view(torch.int16).numpy().tobytes()).hexdigest(),
    }


def full_c1() -> None:
    arch = QWEN3_0_6B
    t0 = time.time()
    model = hf_model(arch, seed=0)
    ids = make_prompts(1, 128, arch.vocab, seed=1)
    with torch.no_grad():
        gen = model.generate(ids, attention_mask=torch.ones_like(ids), do_sample=False, max_new_tokens=128,
                             min_new_tokens=128, pad_token_id=0, output_scores=False)
        new = gen[:, 128:]
        # Teacher-forced pass over prompt + generated tokens: logits at every position and the
        # residual stream after every layer.
        full = torch.cat([ids, new], dim=1)
        out = model(full, output_hidden_states=True)
    logits = out.logits[0].to(torch.float32)  # [256, V]
    top = logits.topk(4, dim=-1)
    layers = []
    for i, h in enumerate(out.hidden_states):
        layers.append({"layer": i, "last_pos": digest(h[0, -1]), "pos127": digest(h[0, 127]), "all": digest(h[0])})
    sd = hf_state_dict(model)
    wd = {k: digest(sd[k]) for k in ("model.embed_tokens.weight", "model.layers.0.self_attn.q_proj.weight",
                                      "model.layers.27.mlp.down_proj.weight")}
    fx = {
        "about": "HF transformers %s Qwen3ForCausalLM, torch %s CPU, sdpa, bf16, seed 0 init, fp32 inv_freq" % (
            __import__("transformers").__version__, torch.__version__),
        "arch": arch.__dict__,
        "prompt_ids": ids[0].tolist(),
        "greedy_ids": new[0].tolist(),
        "tf_top4_ids": top.indices.tolist(),
        "tf_top4_logits": top.values.tolist(),
        "tf_logit_sum": logits.sum(-1).tolist(),
        "tf_logit_absmax": logits.abs().max(-1).values.tolist(),
        "hidden_digests": layers,
        "weight_digests": wd,
        "seconds": time.time() - t0,
    }
    with open(os.path.join(GOLD, "qwen3_0p6b_c1.json"), "w") as f:
        json.dump(fx, f)
    print("wrote qwen3_0p6b_c1.json in %.1fs; unique greedy tokens: %d" % (time.time() - t0, len(set(new[0].tolist()))))


def tiny() -> None:
    arch = QWEN3_TINY
    model = hf_model(arch, seed=0)
    sd = hf_state_dict(model)
    perturb_norm_weights(sd)
    model.load_state_dict(sd)
    ids = make_prompts(3, 48, arch.vocab, seed=1)
    with torch.no_grad():
        gen = model.generate(ids, attention_mask=torch.ones_like(ids), do_sample=False, max_new_tokens=32,
                             min_new_tokens=32, pad_token_id=0)[:, 48:]
        full = torch.cat([ids, gen], dim=1)
        out = model(full, output_hidden_states=True)
    np.savez_compressed(
        os.path.join(GOLD, "qwen3_tiny.npz"),
        prompt_ids=ids.numpy().astype(np.int32),
        greedy_ids=gen.numpy().astype(np.int32),
        logits_bf16=out.logits.view(torch.int16).numpy(),
        hidden_bf16=torch.stack(out.hidden_states).view(torch.int16).numpy(),
    )
    print("wrote qwen3_tiny.npz", out.logits.shape)


def phi3_tiny() -> None:
    """HF Phi3ForCausalLM outputs for the two tiny Phi-3 variants (full and partial rotary)."""
    from oracle.phi3_oracle import PHI3_TINY, PHI3_TINY_PARTIAL, PHI3_TINY_WINDOW, hf_phi3_model

    out_d = {}
    # "window": sliding_window 24 with 40-token prompts + 24 generated tokens, so the window cuts from the start
    for tag, arch in (("full", PHI3_TINY), ("partial", PHI3_TINY_PARTIAL), ("window", PHI3_TINY_WINDOW)):
        model = hf_phi3_model(arch, seed=0)
        sd = hf_state_dict(model)
        perturb_norm_weights(sd)
        model.load_state_dict(sd)
        ids = make_prompts(3, 40, arch.vocab, seed=1)
        with torch.no_grad():
            gen = model.generate(ids, attention_mask=torch.ones_like(ids), do_sample=False, max_new_tokens=24,
                                 min_new_tokens=24, pad_token_id=0)[:, 40:]
            full = torch.cat([ids, gen], dim=1)
            out = model(full, output_hidden_states=True)
        out_d[tag + "_prompt_ids"] = ids.numpy().astype(np.int32)
        out_d[tag + "_greedy_ids"] = gen.numpy().astype(np.int32)
        out_d[tag + "_logits_bf16"] = out.logits.view(torch.int16).numpy()
        out_d[tag + "_hidden_bf16"] = torch.stack(out.hidden_states).view(torch.int16).numpy()
    np.savez_compressed(os.path.join(GOLD, "phi3_tiny.npz"), **out_d)
    print("wrote phi3_tiny.npz")


def opt_tiny() -> None:
    """HF OPTForCausalLM outputs for the tiny OPT configuration (LayerNorm gains and all biases perturbed)."""
    from oracle.opt_oracle import OPT_TINY, hf_opt_model, perturb_opt_weights

    arch = OPT_TINY
    model = hf_opt_model(arch, seed=0)
    sd = hf_state_dict(model)
    perturb_opt_weights(sd)
    model.load_state_dict(sd)
    ids = make_prompts(3, 40, arch.vocab, seed=1)
    with torch.no_grad():
        gen = model.generate(ids, attention_mask=torch.ones_like(ids), do_sample=False, max_new_tokens=24,
                             min_new_tokens=24, pad_token_id=1)[:, 40:]
        full = torch.cat([ids, gen], dim=1)
        out = model(full, output_hidden_states=True)
    np.savez_compressed(
        os.path.join(GOLD, "opt_tiny.npz"),
        prompt_ids=ids.numpy().astype(np.int32),
        greedy_ids=gen.numpy().astype(np.int32),
        logits_bf16=out.logits.view(torch.int16).numpy(),
        hidden_bf16=torch.stack(out.hidden_states).view(torch.int16).numpy(),
    )
    print("wrote opt_tiny.npz", out.logits.shape)


def templates() -> None:
    import jinja2
    import yaml

    cases = [
        ([{"role": "user", "content": "Who are you?"}], True),
        ([{"role": "user", "content": "Who are you?"}], False),
        ([{"role": "system", "content": "Be brief."}, {"role": "user", "content": "Who are you?"}], True),
        ([{"role": "user", "content": "Hi"}, {"role": "assistant", "content": "Hello!"}, {"role": "user", "content": "Bye"}], True),
        ([{"role": "system", "content": "S"}, {"role": "user", "content": "a b\nc"}, {"role": "assistant", "content": ""}], False),
        ([{"role": "tool", "content": "ignored"}, {"role": "user", "content": "x"}], True),
    ]
    res = {}
    for name in ("opt-chat-template", "phi-chat-template"):
        path = "/root/reference/templates/%s.yaml" % name
        src = yaml.safe_load(open(path))["data"]["template.jinja"]
        env = jinja2.Environment()
        tpl = env.from_string(src)
        res[name] = {
            "sha256_16": hashlib.sha256(src.encode()).hexdigest()[:16],
            "chars": len(src),
            "cases": [{"messages": m, "add_generation_prompt": agp, "rendered": tpl.render(messages=m, add_generation_prompt=agp)}
                      for m, agp in cases],
        }
    with open(os.path.join(GOLD, "chat_templates.json"), "w") as f:
        json.dump(res, f, indent=1)
    print("wrote chat_templates.json")


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--skip-full", action="store_true")
    ap.add_argument("--only", default="", help="comma list of: templates,tiny,phi3,opt,full")
    a = ap.parse_args()
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(min(8, os.cpu_count() or 1))
    only = set(x for x in a.only.split(",") if x)
    if not only or "templates" in only:
        templates()
    if not only or "tiny" in only:
        tiny()
    if not only or "phi3" in only:
        phi3_tiny()
    if not only or "opt" in only:
        opt_tiny()
    if (not only and not a.skip_full) or "full" in only:
        full_c1()


if __name__ == "__main__":
    main()
