"""Diagnostic (GPU): 64 concurrent STREAMED chat completions (SSE, one event per token) against b2serve with dummy
Qwen3-0.6B weights -- does per-token event delivery hold the decode rate?"""
import http.client, json, os, subprocess, sys, threading, time
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(R, "tools"))
import serve_bench as sb
proc, port = sb.start()
try:
    rng = np.random.default_rng(1)
    prompts = [rng.integers(0, 151936, 512).tolist() for _ in range(64)]
    sb.wave(port, prompts, 16)
    counts = [0] * 64
    def work(i, n):
        gaps = sb.stream_tpot(port, prompts[i], n)
        counts[i] = len(gaps) + 1
    for n in (1, 256):
        th = [threading.Thread(target=work, args=(i, n)) for i in range(64)]
        t0 = time.perf_counter()
        [t.start() for t in th]; [t.join() for t in th]
        dt = time.perf_counter() - t0
        if n == 1: t1 = dt
    print(json.dumps({"streamed_requests": 64, "new_tokens": 256, "t_1tok_s": t1, "t_256tok_s": dt,
                      "decode_tok_s": 64 * 255 / (dt - t1), "events_per_request_min": min(counts)}))
finally:
    proc.terminate()
    try: proc.wait(timeout=10)
    except Exception: proc.kill()
