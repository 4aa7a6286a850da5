"""Diagnostic (GPU): the fused four-step cluster chain (B2_CHAIN=2) against the separate kernels (B2_CHAIN=0)
on the bench shapes: same weights, same prompts, token streams side by side and ms/step of each."""
import importlib, os, sys
import numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
pkg = importlib.import_module("aws-k8s-ansible-provisioner_b200")
weights = importlib.import_module("aws-k8s-ansible-provisioner_b200.weights")
B, P, STEPS = int(os.environ.get("B", 64)), 512, int(os.environ.get("STEPS", 48))
cfg = pkg.EngineConfig(max_model_len=P + STEPS + 32, max_seqs=B, max_batch_tokens=8192, use_cuda_graphs=os.environ.get("GRAPHS", "1") == "1")
cfg.num_blocks = B * ((cfg.max_model_len + 15) // 16) + 8
sd = weights.random_state_dict(cfg, seed=0)
prompts = torch.randint(0, cfg.vocab, (B, P), generator=torch.Generator().manual_seed(1), dtype=torch.int32).numpy()
res = {}
for mode in ("0", "2"):
    os.environ["B2_CHAIN"] = mode
    eng = pkg.Engine(cfg)
    eng.load_state_dict(sd)
    sids = [eng.seq_create(cfg.max_model_len) for _ in range(B)]
    first = np.concatenate([eng.prefill(sids[i:i + 16], [p.tolist() for p in prompts[i:i + 16]]) for i in range(0, B, 16)])
    toks, ms = eng.decode_run(sids, STEPS)
    toks2, ms2 = eng.decode_run(sids, 16)
    res[mode] = (first, toks, ms2 / 16)
    print("B2_CHAIN=%s: ms/step %.3f" % (mode, ms2 / 16), flush=True)
    eng.close()
a, b = res["0"][1], res["2"][1]
assert np.array_equal(res["0"][0], res["2"][0])
same = (a == b)
first_diff = [int(np.argmax(~same[:, i])) if (~same[:, i]).any() else STEPS for i in range(B)]
print("sequences identical over %d steps: %d/%d; earliest divergence step: %d; step-0 agreement %d/%d" % (
    STEPS, sum(d == STEPS for d in first_diff), B, min(first_diff), int(same[0].sum()), B))
