"""OPT variant of the hot path through the C ABI on the GPU, against HF
OPTForCausalLM's own outputs (tests/golden/opt_tiny.npz) and the pinned restatement
(oracle/opt_oracle.py).  Bars as for the other engine tests: hidden states and logits within 1e-2
relative (or 1.5x / 2x the oracle's own bf16 noise floor where that is larger), greedy ids identical
except after a step whose oracle top-2 gap is within bf16 noise.  OPT's 64-wide heads and one query
head per kv head run zero-padded through the 128-wide kernels."""
import os

import numpy as np
import pytest
import torch

from _util import ROOT, engine_config_from_arch, pkg, rel_err, rel_fro
from oracle.opt_oracle import OPT_125M_2L, OPT_TINY, RestatedOPT, hf_opt_model, linear_fp32_order, perturb_opt_weights
from oracle.qwen3_oracle import attention_explicit, hf_state_dict, top2_gap

pytestmark = pytest.mark.gpu
GOLD = os.path.join(ROOT, "tests", "golden")
REL_TOL = 1e-2


@pytest.fixture(scope="module")
def P():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return pkg()


def bits_to_bf16(bits: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(bits.astype(np.int16)).view(torch.bfloat16)


def make_engine(P, arch, sd, **kw):
    cfg = engine_config_from_arch(P, arch, **kw)
    e = P.Engine(cfg)
    e.load_state_dict(sd)
    return e


@pytest.mark.parametrize("graphs", [False, True])
def test_opt_tiny_matches_hf_golden(P, graphs):
    arch = OPT_TINY
    sd = hf_state_dict(hf_opt_model(arch, seed=0))
    perturb_opt_weights(sd)
    g = np.load(os.path.join(GOLD, "opt_tiny.npz"))
    prompts, gold = g["prompt_ids"], g["greedy_ids"]   # [3][40], [3][24]
    e = make_engine(P, arch, sd, max_model_len=256, num_blocks=64, max_seqs=8, max_batch_tokens=512, use_cuda_graphs=graphs)
    if not graphs:
        e.set_debug(hidden=True, logits=True)
    sids = [e.seq_create(128) for _ in range(3)]
    first = e.prefill(sids, [p.tolist() for p in prompts])
    if not graphs:
        hid = g["hidden_bf16"]  # [L+1][3][64][H]
        T = 3 * 40
        for layer in range(-1, arch.num_layers - 1):  # HF's last entry is after the final LayerNorm: covered by the logits
            got = bits_to_bf16(e.get_hidden(layer, T)).view(3, 40, arch.hidden)
            ref = bits_to_bf16(hid[layer + 1][:, :40])
            assert rel_err(got, ref) < REL_TOL, "hidden after layer %d" % layer
        logits = torch.from_numpy(e.get_logits(3))
        ref_logits = bits_to_bf16(g["logits_bf16"])[:, 39].float()
        assert rel_err(logits, ref_logits) < REL_TOL
    toks = [first] + [e.decode_step(sids) for _ in range(gold.shape[1] - 1)]
    got = np.stack(toks, axis=1)
    r = RestatedOPT(arch, sd)
    _, step_logits = r.greedy(torch.from_numpy(prompts).long(), gold.shape[1])
    gaps = top2_gap(step_logits)
    for b in range(3):
        for t in range(gold.shape[1]):
            if got[b, t] != gold[b, t]:
                assert gaps[b, t] <= 0.0625, "seq %d step %d: token %d != %d with oracle gap %.4f" % (
                    b, t, got[b, t], gold[b, t], gaps[b, t])
                break
    assert (got[:, :8] == gold[:, :8]).mean() > 0.9
    e.close()


def test_opt_125m_widths_two_layers(P):
    """OPT-125m's real widths (hidden 768, 12 heads of 64, ffn 3072, vocab 50272, 2050 learned positions)
    on two layers: ragged prefill + batched decode against the restatement."""
    arch = OPT_125M_2L
    g = torch.Generator().manual_seed(13)

    def n(*shape, std=0.02):
        return (std * torch.randn(*shape, generator=g)).to(torch.bfloat16)

    def gain(k):
        return (1.0 + 0.25 * torch.randn(k, generator=g)).to(torch.bfloat16)

    H, I = arch.hidden, arch.intermediate
    pre = "model.decoder."
    sd = {pre + "embed_tokens.weight": n(arch.vocab, H), pre + "embed_positions.weight": n(arch.max_position + 2, H),
          pre + "final_layer_norm.weight": gain(H), pre + "final_layer_norm.bias": n(H, std=0.1)}
    for l in range(arch.num_layers):
        p = pre + "layers.%d." % l
        for nm in ("q_proj", "k_proj", "v_proj", "out_proj"):
            sd[p + "self_attn.%s.weight" % nm] = n(H, H)
            sd[p + "self_attn.%s.bias" % nm] = n(H, std=0.1)
        sd[p + "self_attn_layer_norm.weight"] = gain(H)
        sd[p + "self_attn_layer_norm.bias"] = n(H, std=0.1)
        sd[p + "final_layer_norm.weight"] = gain(H)
        sd[p + "final_layer_norm.bias"] = n(H, std=0.1)
        sd[p + "fc1.weight"] = n(I, H)
        sd[p + "fc1.bias"] = n(I, std=0.1)
        sd[p + "fc2.weight"] = n(H, I)
        sd[p + "fc2.bias"] = n(H, std=0.1)
    lens = [3, 40, 64, 17]
    prompts = [torch.randint(0, arch.vocab, (k,), generator=g).tolist() for k in lens]
    e = make_engine(P, arch, sd, max_model_len=256, num_blocks=64, max_seqs=8, max_batch_tokens=256, use_cuda_graphs=True)
    e.set_debug(logits=True)
    sids = [e.seq_create(128) for _ in lens]
    first = e.prefill(sids, prompts)
    pre_logits = torch.from_numpy(e.get_logits(len(lens)).copy())
    steps = [first] + [e.decode_step(sids) for _ in range(5)]
    dec_logits = torch.from_numpy(e.get_logits(len(lens)).copy())
    got = np.stack(steps, axis=1)
    r = RestatedOPT(arch, sd, attention=attention_explicit)
    r2 = RestatedOPT(arch, sd, linear=linear_fp32_order)  # same contract, other fp32 order: the noise floor
    for i, pr in enumerate(prompts):
        ids = torch.tensor([pr])
        ref_ids, ref_logits = r.greedy(ids, 6)
        _, alt_logits = r2.greedy(ids, 1)
        ref0 = ref_logits[0, 0].float()
        floor, floor_fro = rel_err(alt_logits[0, 0].float(), ref0), rel_fro(alt_logits[0, 0].float(), ref0)
        err, fro = rel_err(pre_logits[i], ref0), rel_fro(pre_logits[i], ref0)
        print("sequence %d: prefill logits engine vs oracle max-rel %.4g fro %.4g | oracle vs its reordered self %.4g / %.4g" % (
            i, err, fro, floor, floor_fro))
        assert err <= max(REL_TOL, 2.0 * floor), "prefill logits, sequence %d" % i
        assert fro < max(REL_TOL, 1.5 * floor_fro), "prefill logits (Frobenius), sequence %d" % i
        gaps = top2_gap(ref_logits)
        same = True
        for t in range(6):
            if got[i, t] != int(ref_ids[0, t]):
                assert gaps[0, t] <= 0.0625, "seq %d step %d diverges with oracle gap %.4f" % (i, t, gaps[0, t])
                same = False
                break
        if same:
            assert rel_err(dec_logits[i], ref_logits[0, 5].float()) <= max(REL_TOL, 2.0 * floor), "decode logits, sequence %d" % i
    e.close()
