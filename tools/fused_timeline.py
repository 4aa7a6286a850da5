"""Diagnostic (GPU): per-step timeline of the fused cluster chain inside a real C2-shaped decode step."""
import importlib, os, sys
import numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from _util import ptr
os.environ["B2_CHAIN"] = "2"
pkg = importlib.import_module("aws-k8s-ansible-provisioner_b200")
weights = importlib.import_module("aws-k8s-ansible-provisioner_b200.weights")
B, P = 64, 512
cfg = pkg.EngineConfig(max_model_len=P + 64, max_seqs=B, max_batch_tokens=8192, use_cuda_graphs=True)
cfg.num_blocks = B * ((cfg.max_model_len + 15) // 16) + 8
eng = pkg.Engine(cfg)
eng.load_state_dict(weights.random_state_dict(cfg, seed=0))
sids = [eng.seq_create(cfg.max_model_len) for _ in range(B)]
prompts = torch.randint(0, cfg.vocab, (B, P), dtype=torch.int32).numpy()
for i in range(0, B, 16):
    eng.prefill(sids[i:i + 16], [p.tolist() for p in prompts[i:i + 16]])
eng.decode_run(sids, 8, want_tokens=False)
trace = torch.zeros(128 * 32, dtype=torch.int64, device="cuda")
eng.lib.b2k_debug_fused_trace(ptr(trace))
_, ms = eng.decode_run(sids, 4, want_tokens=False)
torch.cuda.synchronize()
eng.lib.b2k_debug_fused_trace(None)
t = trace.cpu().numpy().reshape(128, 32)
t0 = t[:, 8].min()
print("ms/step %.3f (last fused launch of the step = layer 27, steps 0..2)" % (ms / 4))
names = {8: "cta start", 9: "dep ok", 0: "O", 1: "gate-up", 2: "down", 3: "qkv", 10: "end"}
def line(name, col):
    col = (col - t0) / 1e3
    print("  %-28s min %7.2f median %7.2f max %7.2f us" % (name, col.min(), np.median(col), col.max()))
line("cta start", t[:, 8]); line("dep ok", t[:, 9])
for ph in (0, 1, 2):
    line("%s: enter" % names[ph], t[:, ph])
    line("%s: own work issued/scattered" % names[ph], t[:, 16 + ph])
    line("%s: cluster barrier passed" % names[ph], t[:, 24 + ph])
line("end", t[:, 10])

sub = {13: "gate-up: builder has h + rstd", 14: "gate-up: A tile in smem, arrived", 11: "gate-up: MMA warp has W",
       12: "gate-up: MMAs issued", 15: "gate-up: accumulator ready", 20: "gate-up: scatter done (warp 4 lane 0)",
       21: "gate-up: epilogue tail done"}
for k in (13, 14, 11, 12, 15, 20, 21):
    line(sub[k], t[:, k])
