#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric: Qwen3-0.6B bf16 greedy decode throughput (tok/s) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W]            our engine (CUDA, sm_100a)
    python bench.py --impl reference [--gpus N] [--steps K] ...     HF transformers on the host cores

Workload (config.workload = "C2"): BASELINE.json configs[1] -- batch 64, 512-token prompts,
256 generated tokens, paged KV.  One "step" = one decode forward over the whole batch
(64 tokens).  The workload does NOT shrink with --steps: whatever K and W are, the K timed steps
are centred on the middle of C2's decode window (mean context 640.5; `config.timed_context` gives the
first and last context length).  With the defaults (W=8, K=248) the timed region is tokens 9..256 of
every sequence (context 520 -> 768); a short run (K=20) first advances the sequences, untimed, to
context ~630.  The timed region is repeated REPEATS times from the same context and the median is
reported (a 20-step region lasts ~30 ms).  Both arms print the same `config`.

Numbers on the JSON line:
  value      device-resident decode: K steps replayed as CUDA graphs with token feedback on the GPU,
             timed with CUDA events on the engine's stream, max over ranks, median of the repeats
  e2e        the same K steps through the public C-ABI call b2_decode_step with HOST token buffers
             in and out every step (H2D + D2H inside the timed region, host clock)
  roofline   the dominant kernel (paged decode attention), per layer pass, at the mid context:
               frac / achieved   IN THE STEP: (step time - step time with the attention class left out
                                 of the graph) / layers, i.e. what the pass adds to the critical path
                                 with programmatic dependent launch and the two-lane overlap in place
               frac_burst        every layer's pass launched back to back over the real pages, 8 rounds
                                 (224 launches) between one CUDA-event pair (b2_attn_profile)
             The per-class table (kernel_ms_per_step_eager) and ms_per_launch_isolated come from an
             eager single-lane step with an event pair around every launch (adds launch latency).
  cpu_baseline  HF Qwen3 on the host CPU, same batch/context, a bounded number of steps
N > 1: one replica per GPU (request-level data parallelism, no collective on the data path);
ranks only meet at a gloo barrier around the timed region ("scaling": "weak").
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "aws-k8s-ansible-provisioner_b200"

BATCH = 64
PROMPT = 512
OUT_TOKENS = 256
REPEATS = 3
NUM_LAYERS = 28
KV_BYTES_PER_TOKEN = 2 * NUM_LAYERS * 8 * 128 * 2          # K+V, 28 layers, 8 kv heads x 128 dims, bf16
WEIGHT_BYTES = 1_192_099_840                                 # SURVEY.md section 8d (tied head counted once)


def timed_window(K: int, W: int):
    """Context lengths (first, last) of the K timed steps and the untimed advance that puts them there:
    centred on C2's mean context (512 + 257/2) unless the run already covers the window."""
    lo = PROMPT + W if K + W >= OUT_TOKENS else PROMPT + (OUT_TOKENS + 1) // 2 - K // 2
    lo = max(lo, PROMPT + W)
    advance = lo - PROMPT - W
    return lo, lo + K, advance


def c2_config(world: int, K: int, W: int) -> dict:
    """The workload description both arms print (identical for the same --gpus/--steps/--warmup)."""
    lo, hi, _ = timed_window(K, W)
    s_mean = lo + (K + 1) / 2.0
    step_bytes = WEIGHT_BYTES + BATCH * s_mean * KV_BYTES_PER_TOKEN + BATCH * KV_BYTES_PER_TOKEN
    return {"workload": "C2", "model": "Qwen3-0.6B arch random-init", "batch": BATCH, "prompt_len": PROMPT,
            "new_tokens": OUT_TOKENS, "timed_context": [lo, hi], "kv_block": 16,
            "parallelism": "dp%d replicas, no collective" % world,
            "l2": "inputs_exceed_l2 (%.2f GB read per step)" % (step_bytes / 1e9)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device = device
        self.proc = None
        self.lines: list[str] = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup(n_gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # the replicas never exchange data: gloo carries only the barrier and the max-over-ranks
        dist_mod.init_process_group("gloo", rank=rank, world_size=world)
        dist = dist_mod
    return rank, world, local, dist


def barrier(dist):
    if dist is not None:
        dist.barrier()


def reduce_max(dist, x: float) -> float:
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def reduce_sum(dist, x: float) -> float:
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t[0])


# -------------------------------------------------------------------------------------------------
# CPU arm: HF transformers Qwen3 decode steps on the host cores (the reference's arithmetic)
# -------------------------------------------------------------------------------------------------
def hf_cpu_decode(batch: int, ctx: int, steps: int, warmup: int, budget_s: float):
    """Times HF Qwen3ForCausalLM decode forwards (batch x 1 token against a `ctx`-token KV cache,
    the C2 decode shape) on CPU.  The KV cache is filled with random bf16 (no CPU prefill: that alone
    would take minutes and is not part of the metric).  If `steps` full-batch steps would overrun
    `budget_s`, the per-step sample is cut to a sub-batch and said so."""
    import torch
    from transformers import DynamicCache
    from oracle.qwen3_oracle import QWEN3_0_6B, hf_model  # the checker's model builder (CPU arm only)

    arch = QWEN3_0_6B
    # all the host threads the CPU arm can use: torchrun exports OMP_NUM_THREADS=1 to every rank, which
    # would silently turn this into a single-core measurement
    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    if torch.get_num_threads() < avail:
        torch.set_num_threads(avail)
    threads = torch.get_num_threads()
    model = hf_model(arch, seed=0)

    def make_cache(b):
        cache = DynamicCache()
        g = torch.Generator().manual_seed(3)
        for layer in range(arch.num_layers):
            k = torch.randn(b, arch.num_kv_heads, ctx, arch.head_dim, generator=g).to(torch.bfloat16)
            v = torch.randn(b, arch.num_kv_heads, ctx, arch.head_dim, generator=g).to(torch.bfloat16)
            cache.update(k, v, layer)
        return cache

    def run(b, n):
        cache = make_cache(b)
        tok = torch.randint(0, arch.vocab, (b, 1), generator=torch.Generator().manual_seed(4))
        times = []
        with torch.no_grad():
            for i in range(n):
                pos = torch.full((b, 1), ctx + i, dtype=torch.long)
                t0 = time.perf_counter()
                out = model(input_ids=tok, past_key_values=cache, position_ids=pos, use_cache=True)
                tok = out.logits[:, -1].argmax(-1, keepdim=True)
                times.append(time.perf_counter() - t0)
        return times

    b = batch
    probe = run(min(b, 8), 2)  # cheap probe to size the sample
    est_full = probe[-1] * max(1.0, b / min(b, 8))
    while b > 1 and est_full * (steps + warmup) > budget_s:
        b //= 2
        est_full /= 2
    times = run(b, steps + warmup)[warmup:]
    total = sum(times)
    return {"tok_s": b * len(times) / total, "ms_per_step": 1e3 * total / len(times), "batch": b, "threads": threads,
            "steps": len(times), "ctx": ctx}


def run_reference(args, rank, world, dist):
    if rank != 0:
        return
    K, W = args.steps, args.warmup
    lo, hi, _ = timed_window(K, W)
    steps = min(K, 32)  # bounded sample: ~0.3 s per full-batch step on 16 cores
    warm = max(1, min(W, 2))
    ctx = lo + (K - steps) // 2  # the sample sits in the middle of the GPU arm's timed window
    r = hf_cpu_decode(BATCH, ctx - warm, steps=steps, warmup=warm, budget_s=150.0)
    sample = "HF transformers Qwen3 bf16 sdpa on host CPU: %d decode steps, batch %d of %d, context %d..%d (random KV, no prefill)" % (
        r["steps"], r["batch"], BATCH, ctx, ctx + r["steps"])
    line = {
        "impl": "reference", "metric": "decode_tok_s", "value": r["tok_s"], "unit": "tok/s", "n_gpus": args.gpus,
        "steps": r["steps"], "warmup": warm, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": c2_config(world, K, W),
        "cpu_baseline": {"value": r["tok_s"], "unit": "tok/s", "cores": r["threads"], "kind": "reference", "sample": sample},
        "e2e": {"value": r["tok_s"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "host_cpus": os.cpu_count(), "steps_requested": K, "warmup_requested": W,
    }
    print(json.dumps(line), flush=True)


# -------------------------------------------------------------------------------------------------
# GPU arm
# -------------------------------------------------------------------------------------------------
def run_b200(args, rank, world, local, dist):
    import numpy as np
    import torch

    pkg = importlib.import_module(PKG)
    weights = importlib.import_module(PKG + ".weights")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this engine has no CPU path (use --impl reference for the CPU arm)")
    K, W = args.steps, args.warmup
    assert W >= 3, "at least 3 warm-up steps"
    assert W + K <= OUT_TOKENS + 1024
    lo, hi, advance = timed_window(K, W)
    cfg = pkg.EngineConfig(max_model_len=max(PROMPT + OUT_TOKENS, hi) + 16, max_seqs=BATCH, max_batch_tokens=int(os.environ.get("B2_BENCH_CHUNK", "32768")), use_cuda_graphs=True)
    assert cfg.weight_bytes() == WEIGHT_BYTES and cfg.kv_bytes_per_token() == KV_BYTES_PER_TOKEN and cfg.num_layers == NUM_LAYERS
    cfg.num_blocks = BATCH * ((cfg.max_model_len + 15) // 16) + 8
    t0 = time.time()
    sd = weights.random_state_dict(cfg, seed=0)
    eng = pkg.Engine(cfg, device=local)
    eng.load_state_dict(sd)
    del sd
    g = torch.Generator().manual_seed(1 + rank)
    prompts = torch.randint(0, cfg.vocab, (BATCH, PROMPT), generator=g, dtype=torch.int32).numpy()
    sids = [eng.seq_create(cfg.max_model_len) for _ in range(BATCH)]
    per_chunk = max(1, cfg.max_batch_tokens // PROMPT)
    torch.cuda.synchronize(local)
    tp0 = time.perf_counter()
    for i in range(0, BATCH, per_chunk):
        eng.prefill(sids[i:i + per_chunk], [p.tolist() for p in prompts[i:i + per_chunk]])
    prefill_s = time.perf_counter() - tp0
    setup_s = time.time() - t0

    def rewind(ctx):
        """Back to context `ctx` (the K/V of the dropped positions are simply rewritten): one untimed step re-seeds
        the device batch after the truncate."""
        for s_ in sids:
            eng.seq_truncate(s_, ctx - 1)
        eng.decode_step(sids)

    # ---- device-resident decode: untimed advance to the window, then REPEATS x (W warm-up + K timed) steps ----
    if advance > 0:
        eng.decode_run(sids, advance, want_tokens=False)
    sampler = ClockSampler(local)
    launches = 0
    ms_runs = []
    for rep in range(REPEATS):
        if rep > 0:
            rewind(lo - W)
        eng.decode_run(sids, W, want_tokens=False)
        torch.cuda.synchronize(local)
        barrier(dist)
        if rep == 0:
            sampler.start()
        l0 = eng.launch_count()
        _, ms = eng.decode_run(sids, K, want_tokens=False)
        torch.cuda.synchronize(local)
        barrier(dist)
        launches = eng.launch_count() - l0
        ms_runs.append(reduce_max(dist, ms))
    clocks = sampler.stop()
    ms_max = statistics.median(ms_runs)
    tok_s = world * BATCH * K / (ms_max / 1e3)
    if args.value_only:
        if rank == 0:
            print(json.dumps({"ms_per_step": ms_max / K, "value": tok_s, "ms_runs": [m / K for m in ms_runs],
                              "skip_mask": os.environ.get("B2_SKIP_MASK", "0"),
                              "env": {k: v for k, v in os.environ.items() if k.startswith("B2_")}}), flush=True)
        eng.close()
        return

    # ---- end to end through the public call, host buffers in and out every step ----------------
    e2e_runs, step_times = [], []
    for rep in range(REPEATS):
        rewind(lo - W)
        tok = None
        for _ in range(W):
            tok = eng.decode_step(sids, in_tokens=tok)
        torch.cuda.synchronize(local)
        barrier(dist)
        st = []
        te0 = time.perf_counter()
        for _ in range(K):
            ts = time.perf_counter()
            tok = eng.decode_step(sids, in_tokens=tok)  # H2D ids, graph launch, D2H ids, sync
            st.append(time.perf_counter() - ts)
        e2e_s = time.perf_counter() - te0
        barrier(dist)
        e2e_runs.append(reduce_max(dist, e2e_s))
        step_times += st
    e2e_max = statistics.median(e2e_runs)
    e2e_tok_s = world * BATCH * K / e2e_max

    # ---- the dominant kernel in the step: the same K steps with the attention class left out of the graph ----
    ATTN_CLASS = 4  # aws-k8s-ansible-provisioner_b200.engine.KERNEL_CLASSES.index("attn_decode")
    eng.set_tuning(skip_mask=1 << ATTN_CLASS)
    noattn_runs = []
    for rep in range(REPEATS):
        rewind(lo - W)
        eng.decode_run(sids, W, want_tokens=False)
        _, ms = eng.decode_run(sids, K, want_tokens=False)
        noattn_runs.append(ms)
    eng.set_tuning(skip_mask=0)
    ms_noattn = statistics.median(noattn_runs)
    attn_ms_in_step = max(1e-6, (ms_max - reduce_max(dist, ms_noattn)) / K / cfg.num_layers)

    # ---- the same kernel alone: eager profiled pass and a back-to-back burst at mid context ------
    mid = lo + K // 2
    rewind(mid)
    prof = eng.decode_profile(sids, 3)
    attn = prof["attn_decode"]
    attn_ms_isolated = attn["ms"] / max(1, attn["launches"])  # event pair around every single launch: + launch latency
    # all layers' passes back to back over the real pages between ONE event pair, 8 rounds
    # (28 x 8 launches; every layer's K/V is its own 170 MB region, so no launch finds its input in L2)
    attn_ms_burst, kv_bytes = eng.attn_profile(sids, reps=8)
    ctx_burst = kv_bytes // (BATCH * 2 * cfg.num_kv_heads * cfg.head_dim * 2)
    io_bytes = 2 * BATCH * cfg.num_q_heads * cfg.head_dim * 2  # q in, o out
    burst_bytes = kv_bytes + io_bytes
    # in the step the pass of one layer reads the K/V of the mean context of the timed region
    s_mean = lo + (K + 1) / 2.0
    step_pass_bytes = BATCH * s_mean * (KV_BYTES_PER_TOKEN // cfg.num_layers) + io_bytes
    peak, peak_kind = measured_peaks()
    achieved = step_pass_bytes / (attn_ms_in_step / 1e3) / 1e9
    achieved_burst = burst_bytes / (attn_ms_burst / 1e3) / 1e9
    # DRAM traffic of one pass, from the committed ncu --set full capture (context 641; scaled to this context by its
    # own traffic/algorithmic ratio, which is what the figure is for)
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "r02_attn_traffic.json")) as f:
            tj = json.load(f)
        traffic = step_pass_bytes * (tj["dram_bytes_read"] + tj["dram_bytes_write"]) / tj["algorithmic_bytes"]
    except (OSError, KeyError, ValueError):
        pass
    prof_total = sum(v["ms"] for v in prof.values()) / 3.0
    kernel_ms = {k: round(v["ms"] / 3.0, 4) for k, v in prof.items() if v["launches"]}
    # whole-step algorithmic bytes: W + B*S*kv + B*kv at the mean context of the timed region
    step_bytes = WEIGHT_BYTES + BATCH * s_mean * KV_BYTES_PER_TOKEN + BATCH * KV_BYTES_PER_TOKEN
    step_frac = (step_bytes / (ms_max / K / 1e3) / 1e9) / peak

    line = {
        "metric": "decode_tok_s", "value": tok_s, "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": c2_config(world, K, W),
        "e2e": {"value": e2e_tok_s, "unit": "tok/s", "h2d_bytes_per_step": BATCH * 4, "d2h_bytes_per_step": BATCH * 4},
        "tpot_ms_p50": 1e3 * statistics.median(step_times),
        "gpu_launches": int(launches),
        "clocks": clocks,
        "repeats": REPEATS, "ms_per_step_runs": [round(m / K, 5) for m in ms_runs],
        "roofline": {"bound": "hbm", "kernel": "attn_decode_kernel (one layer's pass)", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "traffic_source": "ncu --set full capture at context 641 (profiles/r02_attn_traffic.json): DRAM read+write = 1.024 x algorithmic, applied to this context", "peak_kind": peak_kind,
                     "how": "in-step: (step - step without the attention class) / layers",
                     "bytes_per_launch": step_pass_bytes, "ms_per_launch": attn_ms_in_step, "context": s_mean,
                     "ms_per_step_without_attention": ms_noattn / K,
                     "frac_burst": achieved_burst / peak, "achieved_burst": achieved_burst, "bytes_per_launch_burst": burst_bytes,
                     "ms_per_launch_burst": attn_ms_burst, "context_burst": ctx_burst, "launches_timed_burst": 8 * cfg.num_layers,
                     "ms_per_launch_isolated": attn_ms_isolated},
        "step_hbm_frac": step_frac,
        "step_bytes": step_bytes,
        "kernel_ms_per_step_eager": kernel_ms,
        "eager_step_ms": prof_total,
        "prefill_s": prefill_s,
        "setup_s": setup_s,
        "two_lane": os.environ.get("B2_MB", "default"),
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            r = hf_cpu_decode(BATCH, lo + K // 2, steps=6, warmup=1, budget_s=40.0)
            line["cpu_baseline"] = {
                "value": r["tok_s"], "unit": "tok/s", "cores": r["threads"], "kind": "reference",
                "sample": "HF transformers Qwen3 bf16 sdpa on host CPU: %d decode steps, batch %d of %d, context %d (random KV, "
                          "no prefill); host has %d logical CPUs" % (r["steps"], r["batch"], BATCH, r["ctx"], os.cpu_count())}
        except Exception as ex:  # the GPU numbers stand on their own
            line["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": 0, "kind": "reference", "sample": "failed: %r" % (ex,)}
    eng.close()
    if rank == 0:
        print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=248)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--value-only", action="store_true", help="device-resident timing only (ablation runs)")
    args = ap.parse_args()
    rank, world, local, dist = dist_setup(args.gpus)
    try:
        if args.impl == "reference":
            run_reference(args, rank, world, dist)
        else:
            run_b200(args, rank, world, local, dist)
    finally:
        if dist is not None:
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
