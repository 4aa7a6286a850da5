"""Diagnostic (GPU): is decode deterministic run to run?  Same prompts, fresh sequences, N repetitions."""
import importlib, os, sys
import numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
pkg = importlib.import_module("aws-k8s-ansible-provisioner_b200")
weights = importlib.import_module("aws-k8s-ansible-provisioner_b200.weights")
full = len(sys.argv) > 1 and sys.argv[1] == "full"
graphs = int(os.environ.get("GRAPHS", "1"))
if full:
    cfg = pkg.EngineConfig(max_model_len=512, max_seqs=64, max_batch_tokens=8192, use_cuda_graphs=bool(graphs), num_blocks=64 * 20 + 8)
    B, P, STEPS = 64, 120, 24
else:
    cfg = pkg.EngineConfig(num_layers=2, hidden=256, num_q_heads=4, num_kv_heads=2, intermediate=512, vocab=1024, max_model_len=256,
                           max_seqs=8, max_batch_tokens=512, use_cuda_graphs=bool(graphs), num_blocks=128)
    B, P, STEPS = 4, 33, 40
eng = pkg.Engine(cfg)
eng.load_state_dict(weights.random_state_dict(cfg, seed=0))
g = torch.Generator().manual_seed(3)
prompts = torch.randint(0, cfg.vocab, (B, P), generator=g, dtype=torch.int32).numpy()
ref = None
bad = 0
first_bad = None
for rep in range(int(os.environ.get("REPS", "12"))):
    sids = [eng.seq_create(P + STEPS + 8) for _ in range(B)]
    per = max(1, cfg.max_batch_tokens // P)
    firsts = []
    for i in range(0, B, per):
        firsts.append(eng.prefill(sids[i:i + per], [p.tolist() for p in prompts[i:i + per]]))
    toks, _ = eng.decode_run(sids, STEPS)
    out = np.concatenate([np.concatenate(firsts)[None, :], toks], axis=0)
    if ref is None:
        ref = out
    elif not np.array_equal(out, ref):
        bad += 1
        if first_bad is None:
            d = np.argwhere(out != ref)
            first_bad = (rep, int(d[0][0]), int(d[0][1]), int((out != ref).sum()))
    for s in sids:
        eng.seq_free(s)
print("env", {k: v for k, v in os.environ.items() if k.startswith("B2_") or k in ("GRAPHS",)}, "full" if full else "tiny",
      "-> %d/%d repetitions differ" % (bad, int(os.environ.get("REPS", "12")) - 1), "first diff (rep, step, seq, count):", first_bad)
