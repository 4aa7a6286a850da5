"""Diagnostic (GPU): per-step timeline of the decode chain kernel inside a real C2-shaped decode step."""
import importlib, os, sys
import numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from _util import ptr
pkg = importlib.import_module("aws-k8s-ansible-provisioner_b200")
weights = importlib.import_module("aws-k8s-ansible-provisioner_b200.weights")
B, P = 64, int(sys.argv[1]) if len(sys.argv) > 1 else 512
cfg = pkg.EngineConfig(max_model_len=P + 64, max_seqs=B, max_batch_tokens=8192, use_cuda_graphs=True)
cfg.num_blocks = B * ((cfg.max_model_len + 15) // 16) + 8
eng = pkg.Engine(cfg)
eng.load_state_dict(weights.random_state_dict(cfg, seed=0))
sids = [eng.seq_create(cfg.max_model_len) for _ in range(B)]
prompts = torch.randint(0, cfg.vocab, (B, P), dtype=torch.int32).numpy()
for i in range(0, B, 16):
    eng.prefill(sids[i:i + 16], [p.tolist() for p in prompts[i:i + 16]])
eng.decode_run(sids, 8, want_tokens=False)
trace = torch.zeros(148 * 32, dtype=torch.int64, device="cuda")
eng.lib.b2k_debug_chain_trace(ptr(trace))
_, ms = eng.decode_run(sids, 4, want_tokens=False)
torch.cuda.synchronize()
eng.lib.b2k_debug_chain_trace(None)
t = trace.cpu().numpy().reshape(148, 32)
names = {8: "cta start", 9: "dep ok", 0: "O gemm", 1: "add+norm", 2: "gate-up gemm", 3: "swiglu", 4: "down gemm", 5: "add+norm", 6: "qkv gemm", 7: "rope+kv", 10: "cta done"}
t0 = t[:, 8].min()
print("ms/step %.3f (last chain launch of the step = layer 27, steps 0..5)" % (ms / 4))
for k in (8, 9, 0, 1, 2, 3, 4, 5, 10):
    col = (t[:, k] - t0) / 1e3
    line = "  %-14s enter: min %7.2f median %7.2f max %7.2f us" % (names[k], col.min(), np.median(col), col.max())
    if k < 8 and t[:, 16 + k].max() > 0:
        w = (t[:, 16 + k] - t0) / 1e3
        line += "   | work done: min %7.2f median %7.2f max %7.2f" % (w.min(), np.median(w), w.max())
    print(line)

sub = {24: "O: first stage full (MMA warp)", 25: "O: MMAs issued", 26: "O: accumulator ready (epilogue)", 27: "O: partial stores issued",
       28: "norm: loads summed", 29: "norm: row reduced", 30: "norm: x stored", 31: "norm: proxy fence done"}
busy = t[:, 24] > 0
print("  busy CTAs (have an O tile): %d" % busy.sum())
for k in sorted(sub):
    m = t[:, k] > 0
    if not m.any():
        continue
    col = (t[m, k] - t0) / 1e3
    print("  %-34s min %7.2f median %7.2f max %7.2f us" % (sub[k], col.min(), np.median(col), col.max()))
