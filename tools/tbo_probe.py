"""Diagnostic (GPU): do two half-batch decode streams overlap on one GPU?

Two independent engines (32 sequences each, own stream, own graphs) replay decode steps from two
host threads at once; compared with each of them alone and with one 64-sequence engine.  If the pair
finishes a step of 64 tokens faster than the single engine, splitting the batch into two interleaved
half-batches inside one engine would pay."""
import importlib, os, sys, threading, time
import torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
pkg = importlib.import_module("aws-k8s-ansible-provisioner_b200")
weights = importlib.import_module("aws-k8s-ansible-provisioner_b200.weights")
P, STEPS = 512, 96

def make(B):
    cfg = pkg.EngineConfig(max_model_len=P + 256, max_seqs=B, max_batch_tokens=8192, use_cuda_graphs=True)
    cfg.num_blocks = B * ((cfg.max_model_len + 15) // 16) + 8
    eng = pkg.Engine(cfg)
    eng.load_state_dict(weights.random_state_dict(cfg, seed=0))
    sids = [eng.seq_create(cfg.max_model_len) for _ in range(B)]
    prompts = torch.randint(0, cfg.vocab, (B, P), dtype=torch.int32).numpy()
    for i in range(0, B, 16):
        eng.prefill(sids[i:i + 16], [p.tolist() for p in prompts[i:i + 16]])
    eng.decode_run(sids, 8, want_tokens=False)
    return eng, sids

def run(engs):
    out = [None] * len(engs)
    def work(i):
        e, s = engs[i]
        out[i] = e.decode_run(s, STEPS, want_tokens=False)[1] / STEPS
    th = [threading.Thread(target=work, args=(i,)) for i in range(len(engs))]
    t0 = time.perf_counter()
    for t in th: t.start()
    for t in th: t.join()
    wall = (time.perf_counter() - t0) * 1e3 / STEPS
    return out, wall

big = make(64)
print("one engine, 64 seqs      : ms/step", run([big]))
del big
a, b = make(32), make(32)
print("one engine, 32 seqs alone: ms/step", run([a]))
print("two engines, 32+32       : ms/step", run([a, b]))
if len(sys.argv) > 1:
    c, d = make(16), make(16)
    print("four engines, 32+32+16+16: ms/step", run([a, b, c, d]))
