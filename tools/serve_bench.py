"""Diagnostic (GPU): BASELINE config 2 through the HTTP boundary -- b2serve with dummy Qwen3-0.6B weights,
64 concurrent /v1/completions requests with 512-token id prompts and 256 new tokens (greedy, ignore_eos).
Prints one JSON line: request-level throughput including prefill, decode-only tok/s (difference to a
1-token run), and TPOT p50 from streamed chunks of one extra request riding along."""
import http.client, json, os, subprocess, sys, threading, time
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(R, "aws-k8s-ansible-provisioner_b200", "b2serve")
B, P, N = 64, 512, 256

def start():
    proc = subprocess.Popen([EXE, "--load-format", "dummy", "--served-model-name", "Qwen/Qwen3-0.6B", "--port", "0", "--host", "127.0.0.1",
                             "--max-model-len", "1024", "--max-num-seqs", "64", "--max-num-batched-tokens", "32768", "--eos-token-id", "-1"],
                            stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    t0 = time.time()
    while time.time() - t0 < 180:
        line = proc.stdout.readline()
        if line.startswith("READY"):
            return proc, int(line.split()[1])
        if proc.poll() is not None:
            break
    proc.kill()
    raise SystemExit("b2serve did not start: " + proc.stderr.read()[-2000:])

def post(port, body, timeout=600):
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=timeout)
    c.request("POST", "/v1/completions", json.dumps(body), {"Content-Type": "application/json"})
    r = c.getresponse()
    data = r.read()
    c.close()
    return r.status, json.loads(data)

def wave(port, prompts, n):
    out = [None] * len(prompts)
    def work(i):
        out[i] = post(port, {"model": "Qwen/Qwen3-0.6B", "prompt": prompts[i], "max_tokens": n, "temperature": 0, "ignore_eos": True})
    th = [threading.Thread(target=work, args=(i,)) for i in range(len(prompts))]
    t0 = time.perf_counter()
    for t in th: t.start()
    for t in th: t.join()
    dt = time.perf_counter() - t0
    for st, j in out:
        assert st == 200 and j["usage"]["completion_tokens"] == n, (st, j)
    return dt

def stream_tpot(port, prompt, n):
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=600)
    c.request("POST", "/v1/completions", json.dumps({"model": "Qwen/Qwen3-0.6B", "prompt": prompt, "max_tokens": n, "temperature": 0,
                                                      "ignore_eos": True, "stream": True}), {"Content-Type": "application/json"})
    r = c.getresponse()
    stamps = []
    buf = b""
    while True:
        chunk = r.read1(65536)
        if not chunk:
            break
        buf += chunk
        while b"\n\n" in buf:
            ev, buf = buf.split(b"\n\n", 1)
            if ev.startswith(b"data: ") and ev != b"data: [DONE]":
                stamps.append(time.perf_counter())
    c.close()
    return np.diff(np.array(stamps))

def main():
    proc, port = start()
    try:
        rng = np.random.default_rng(1)
        prompts = [rng.integers(0, 151936, P).tolist() for _ in range(B)]
        wave(port, prompts[:8], 8)                       # warm-up (graphs for the buckets that follow are captured lazily)
        wave(port, prompts, 16)
        t1 = min(wave(port, prompts, 1) for _ in range(2))
        tN = min(wave(port, prompts, N) for _ in range(2))
        # TPOT: one streamed request while 63 others keep the batch full
        gaps = []
        def bg():
            wave(port, prompts[:B - 1], N)
        th = threading.Thread(target=bg); th.start()
        time.sleep(0.05)
        gaps = stream_tpot(port, prompts[B - 1], N)
        th.join()
        print(json.dumps({"workload": "C2 over HTTP", "requests": B, "prompt_len": P, "new_tokens": N,
                          "t_1tok_s": t1, "t_%dtok_s" % N: tN, "request_tok_s_incl_prefill": B * N / tN,
                          "decode_tok_s": B * (N - 1) / (tN - t1), "ms_per_step": (tN - t1) / (N - 1) * 1e3,
                          "tpot_ms_p50_streamed": float(np.median(gaps) * 1e3) if len(gaps) else None,
                          "tpot_ms_p99_streamed": float(np.percentile(gaps, 99) * 1e3) if len(gaps) else None}), flush=True)
    finally:
        proc.terminate()
        try:
            proc.wait(timeout=10)
        except Exception:
            proc.kill()

if __name__ == "__main__":
    main()
