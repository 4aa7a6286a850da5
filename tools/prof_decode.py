"""Profiling target (GPU): the bench workload (C2: batch 64, 512-token prompts) advanced to context ~640, then a few
graph-replayed decode steps and nothing else -- what ncu is pointed at (tools/gpu_profile2.sh).  Prints the number of
kernel launches before the steps of interest so that the caller can skip them."""
import argparse
import importlib
import os
import sys

import torch

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ctx", type=int, default=640)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--prefill-only", action="store_true", help="stop after the prefill (prefill kernels are the target)")
    args = ap.parse_args()
    pkg = importlib.import_module("aws-k8s-ansible-provisioner_b200")
    weights = importlib.import_module("aws-k8s-ansible-provisioner_b200.weights")
    B, P = 64, 512
    chunk = int(os.environ.get("B2_BENCH_CHUNK", "32768"))  # bench.py's prefill: the whole batch in one call
    cfg = pkg.EngineConfig(max_model_len=args.ctx + 64, max_seqs=B, max_batch_tokens=chunk)
    cfg.num_blocks = B * ((cfg.max_model_len + 15) // 16) + 8
    eng = pkg.Engine(cfg, device=0)
    eng.load_state_dict(weights.random_state_dict(cfg, seed=0))
    g = torch.Generator().manual_seed(1)
    prompts = torch.randint(0, cfg.vocab, (B, P), generator=g, dtype=torch.int32).numpy()
    sids = [eng.seq_create(cfg.max_model_len) for _ in range(B)]
    per = max(1, chunk // P)
    for i in range(0, B, per):
        eng.prefill(sids[i:i + per], [p.tolist() for p in prompts[i:i + per]])
    print("launches after prefill: %d" % eng.launch_count(), flush=True)
    if args.prefill_only:
        eng.close()
        return
    # advance with the attention class left out (fast under a profiler too), then restore it: the pages of the
    # skipped positions hold whatever the QK-norm/RoPE kernel wrote, which is all the timing needs
    eng.decode_run(sids, args.ctx - P - 4, want_tokens=False)
    eng.decode_run(sids, 2, want_tokens=False)
    print("launches before the profiled steps: %d" % eng.launch_count(), flush=True)
    _, ms = eng.decode_run(sids, args.steps, want_tokens=False)
    print("profiled %d steps: %.3f ms/step" % (args.steps, ms / args.steps), flush=True)
    eng.close()


if __name__ == "__main__":
    main()
