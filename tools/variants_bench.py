"""Informational (GPU): decode throughput of the two parity configurations at their full shapes --
C3 OPT-125m (batch 256, 128-token prompts) and C4 Phi-3-mini-4k (batch 32, 2048-token prompts) -- against the
algorithmic-bytes floors of BASELINE.md section 4.  These models run zero-padded through the 128-wide kernels
("Model variants"), so the distance to the floor includes the padding's wasted bytes."""
import importlib, json, os, sys, time
import numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from _util import engine_config_from_arch, pkg
from oracle.opt_oracle import OPT_125M
from oracle.phi3_oracle import PHI3_MINI, phi3_inv_freq
from test_variants_fullsize_gpu import opt_weights, phi3_weights
P = pkg()
def _peak():
    try:
        return float(json.load(open(os.path.join(R, "MEASURED_PEAKS.json")))["hbm_gbs"]) * 1e9
    except Exception:
        return 6650.0e9  # the profiling recipe's fallback


PEAK = _peak()
CHUNK_TOKENS = int(os.environ.get("B2_BENCH_CHUNK", "32768"))  # tokens per prefill call (bench.py's default)
ONLY = os.environ.get("B2_VARIANTS", "c3,c4").split(",")

def run(tag, arch, sd, B, PL, new, chunk, inv=None):
    cfg = engine_config_from_arch(P, arch, max_model_len=PL + new + 16, max_seqs=B, max_batch_tokens=CHUNK_TOKENS, use_cuda_graphs=True)
    cfg.num_blocks = B * ((cfg.max_model_len + 15) // 16) + 8
    e = P.Engine(cfg, rope_inv_freq=inv)
    e.load_state_dict(sd)
    g = torch.Generator().manual_seed(1)
    prompts = torch.randint(0, arch.vocab, (B, PL), generator=g, dtype=torch.int32).numpy()
    sids = [e.seq_create(cfg.max_model_len) for _ in range(B)]
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for i in range(0, B, chunk):
        e.prefill(sids[i:i + chunk], [p.tolist() for p in prompts[i:i + chunk]])
    torch.cuda.synchronize(); tp = time.perf_counter() - t0
    e.decode_run(sids, 8, want_tokens=False)
    _, ms = e.decode_run(sids, new - 8, want_tokens=False)
    ms_step = ms / (new - 8)
    w = 2 * sum(v.numel() for k, v in sd.items() if k != "lm_head.weight" or not arch.tie_embeddings)
    kv = arch.kv_bytes_per_token()
    smean = PL + 8 + (new - 8 + 1) / 2
    bytes_step = w + B * smean * kv + B * kv
    print(json.dumps({"config": tag, "batch": B, "prompt": PL, "ms_per_step": ms_step, "decode_tok_s": B / ms_step * 1e3,
                      "algorithmic_GB_per_step": bytes_step / 1e9, "frac_of_measured_hbm_peak": bytes_step / (ms_step / 1e3) / PEAK,
                      "prefill_tok_s": B * PL / tp}), flush=True)
    e.close()

if "c3" in ONLY:
    run("C3 OPT-125m", OPT_125M, opt_weights(OPT_125M, 3), 256, 128, 136, max(1, CHUNK_TOKENS // 128))
if "c4" in ONLY:
    run("C4 Phi-3-mini-4k", PHI3_MINI, phi3_weights(PHI3_MINI, 4), 32, 2048, 72, max(1, CHUNK_TOKENS // 2048), phi3_inv_freq(PHI3_MINI).numpy())
