"""BASELINE.json configs 3 and 4 at their full shapes (GPU), through size-independent properties.
The CPU oracle cannot run these sizes in test time (config 4's prefill alone is ~500 TFLOP), so what is
checked is what must hold for ANY correct engine at any size: identical prompts give identical first
tokens, the same batch gives the same bits run after run, a sequence decodes the same stream inside a
full batch as alone, and pages are returned.  "The same stream" is only asserted for the leading tokens:
with random-init weights the top two logits are often within one bf16 ulp of each other (OPT's tied head
over 50 272 rows of N(0, 0.02) weights gives a logit spread of ~0.5), and decode attention cuts a
sequence's pages differently depending on its place in the batch, so copies legitimately separate at such
near-ties as the stream goes on; the printed counts say how far they stayed together.  Numerical parity
of the same code paths against HF is in test_engine_opt_gpu.py / test_engine_phi3_gpu.py."""
import numpy as np
import pytest
import torch

from _util import bf16_ulp_of, engine_config_from_arch, pkg
from oracle.opt_oracle import OPT_125M
from oracle.phi3_oracle import PHI3_MINI, phi3_inv_freq
from oracle.qwen3_oracle import make_prompts

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return pkg()


def opt_weights(arch, seed):
    g = torch.Generator().manual_seed(seed)

    def n(*shape, std=0.02):
        return (std * torch.randn(*shape, generator=g)).to(torch.bfloat16)

    H, I = arch.hidden, arch.intermediate
    pre = "model.decoder."
    sd = {pre + "embed_tokens.weight": n(arch.vocab, H), pre + "embed_positions.weight": n(arch.max_position + 2, H),
          pre + "final_layer_norm.weight": torch.ones(H, dtype=torch.bfloat16), pre + "final_layer_norm.bias": n(H)}
    for l in range(arch.num_layers):
        p = pre + "layers.%d." % l
        for nm in ("q_proj", "k_proj", "v_proj", "out_proj"):
            sd[p + "self_attn.%s.weight" % nm] = n(H, H)
            sd[p + "self_attn.%s.bias" % nm] = n(H)
        for nm in ("self_attn_layer_norm", "final_layer_norm"):
            sd[p + nm + ".weight"] = torch.ones(H, dtype=torch.bfloat16)
            sd[p + nm + ".bias"] = n(H)
        sd[p + "fc1.weight"], sd[p + "fc1.bias"] = n(I, H), n(I)
        sd[p + "fc2.weight"], sd[p + "fc2.bias"] = n(H, I), n(H)
    return sd


def phi3_weights(arch, seed):
    g = torch.Generator().manual_seed(seed)

    def n(*shape):
        return (0.02 * torch.randn(*shape, generator=g)).to(torch.bfloat16)

    H, I, hd = arch.hidden, arch.intermediate, arch.head_dim
    ones = torch.ones(H, dtype=torch.bfloat16)
    sd = {"model.embed_tokens.weight": n(arch.vocab, H), "lm_head.weight": n(arch.vocab, H), "model.norm.weight": ones}
    for l in range(arch.num_layers):
        p = "model.layers.%d." % l
        sd[p + "input_layernorm.weight"] = ones
        sd[p + "post_attention_layernorm.weight"] = ones
        sd[p + "self_attn.qkv_proj.weight"] = n((arch.num_q_heads + 2 * arch.num_kv_heads) * hd, H)
        sd[p + "self_attn.o_proj.weight"] = n(H, arch.num_q_heads * hd)
        sd[p + "mlp.gate_up_proj.weight"] = n(2 * I, H)
        sd[p + "mlp.down_proj.weight"] = n(H, I)
    return sd


def properties(e, B, prompt, reserve, steps, chunk):
    """Runs the batch twice (determinism) and one copy alone; returns (batch tokens, solo tokens)."""
    runs = []
    for rep in range(2):
        sids = [e.seq_create(reserve) for _ in range(B)]
        firsts = []
        for i in range(0, B, chunk):
            firsts.append(e.prefill(sids[i:i + chunk], [prompt] * len(sids[i:i + chunk])))
        first = np.concatenate(firsts)
        assert (first == first[0]).all(), "identical prompts produced different first tokens"
        if rep == 0:
            # the first decode step through the call that also returns logits: how close to a tie is its arg-max?
            e.set_debug(logits=True)
            t0 = e.decode_step(sids)
            lg = torch.from_numpy(e.get_logits(B)).topk(2, dim=-1).values
            e.set_debug(logits=False)
            properties.step0_gap_ulp = float(((lg[:, 0] - lg[:, 1]) / torch.tensor([bf16_ulp_of(float(v)) for v in lg[:, 0]])).min())
            rest, _ = e.decode_run(sids, steps - 1)
            toks = np.concatenate([t0[None, :], rest], axis=0)
        else:
            toks, _ = e.decode_run(sids, steps)
        runs.append((first.copy(), toks.copy()))
        for s in sids:
            e.seq_free(s)
    assert np.array_equal(runs[0][0], runs[1][0]) and np.array_equal(runs[0][1], runs[1][1]), "not deterministic run to run"
    free_before = e.info()["free_blocks"]
    sid = e.seq_create(reserve)
    f1 = e.prefill([sid], [prompt])
    t1, _ = e.decode_run([sid], steps)
    e.seq_free(sid)
    assert e.info()["free_blocks"] == free_before, "pages leaked"
    assert int(f1[0]) == int(runs[0][0][0])
    return runs[0][1], t1[:, 0]


def test_c3_opt125m_batch256(P):
    """Config 3: OPT-125m (12 layers), batch 256, 128-token prompts."""
    arch = OPT_125M
    B, PL, STEPS = 256, 128, 32
    cfg = engine_config_from_arch(P, arch, max_model_len=PL + 64, max_seqs=B, max_batch_tokens=8192, use_cuda_graphs=True)
    cfg.num_blocks = (B + 1) * ((cfg.max_model_len + 15) // 16) + 8
    e = P.Engine(cfg)
    e.load_state_dict(opt_weights(arch, 3))
    prompt = make_prompts(1, PL, arch.vocab, seed=1)[0].tolist()
    toks, solo = properties(e, B, prompt, PL + 48, STEPS, chunk=64)
    if not (toks[0] == toks[0, 0]).all():
        assert properties.step0_gap_ulp <= 2.0, "first decode step differs across identical sequences away from a tie (gap %.1f ulp)" % properties.step0_gap_ulp
    together = (toks == toks[:, :1]).all(axis=0).mean()
    agree = int((solo == toks[:, 0]).sum())
    print("C3: %d%% of 256 copies stayed together for %d steps; batch vs solo agree on %d/%d tokens" % (
        100 * together, STEPS, agree, STEPS))
    first_split = next((t for t in range(STEPS) if (toks[t] != toks[t, 0]).any()), STEPS)
    print("C3: first step at which any copy differs: %d" % first_split)
    # `together` is reported, not asserted: it counts near-ties.  A tie at the very first step (gap <= 2 ulp) lets the copies and
    # the solo run part at once; otherwise they must stay together for a few steps
    assert properties.step0_gap_ulp <= 2.0 or (first_split >= 2 and agree >= 4)
    e.close()


def test_c4_phi3mini_batch32_prefill2048(P):
    """Config 4: Phi-3-mini-4k (32 layers, 3072 wide, 32 heads of 96, sliding window 2047), batch 32, 2048-token
    prefill, then decode -- the contexts of this configuration are longer than the window from the first step."""
    arch = PHI3_MINI
    B, PL, STEPS = 32, 2048, 16
    cfg = engine_config_from_arch(P, arch, max_model_len=PL + 64, max_seqs=B, max_batch_tokens=8192, use_cuda_graphs=True)
    cfg.num_blocks = (B + 1) * ((cfg.max_model_len + 15) // 16) + 8
    e = P.Engine(cfg, rope_inv_freq=phi3_inv_freq(arch).numpy())
    e.load_state_dict(phi3_weights(arch, 4))
    prompt = make_prompts(1, PL, arch.vocab, seed=1)[0].tolist()
    toks, solo = properties(e, B, prompt, PL + 32, STEPS, chunk=4)
    # copies of one sequence sit in different rows, and decode attention cuts a row's pages by its place in the batch
    # (DESIGN.md section 2): they may part at a step whose top two logits are within bf16 noise of each other, not elsewhere
    if not (toks[0] == toks[0, 0]).all():
        assert properties.step0_gap_ulp <= 2.0, "first decode step differs across identical sequences away from a tie (gap %.1f ulp)" % properties.step0_gap_ulp
    together = (toks == toks[:, :1]).all(axis=0).mean()
    agree = int((solo == toks[:, 0]).sum())
    print("C4: %d%% of 32 copies stayed together for %d steps; batch vs solo agree on %d/%d tokens" % (
        100 * together, STEPS, agree, STEPS))
    first_split = next((t for t in range(STEPS) if (toks[t] != toks[t, 0]).any()), STEPS)
    print("C4: first step at which any copy differs: %d" % first_split)
    # `together` is reported, not asserted: it counts near-ties.  A tie at the very first step (gap <= 2 ulp) lets the copies and
    # the solo run part at once; otherwise they must stay together for a few steps
    assert properties.step0_gap_ulp <= 2.0 or (first_split >= 2 and agree >= 4)
    e.close()
