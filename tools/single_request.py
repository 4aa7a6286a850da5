"""Diagnostic (GPU): the reference's own smoke request (llm-d-test.yaml:61-73: one /v1/completions call, default 16 new
tokens) against b2serve with dummy Qwen3-0.6B weights: wall-clock latency of the whole HTTP round trip."""
import http.client, json, os, subprocess, sys, time
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(R, "aws-k8s-ansible-provisioner_b200", "b2serve")
proc = subprocess.Popen([EXE, "--load-format", "dummy", "--served-model-name", "Qwen/Qwen3-0.6B", "--port", "0", "--host", "127.0.0.1",
                         "--max-model-len", "1024", "--max-num-seqs", "64", "--eos-token-id", "-1"],
                        stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
try:
    port = None
    t0 = time.time()
    while time.time() - t0 < 180:
        line = proc.stdout.readline()
        if line.startswith("READY"):
            port = int(line.split()[1]); break
    def one(body):
        c = http.client.HTTPConnection("127.0.0.1", port, timeout=60)
        t = time.perf_counter()
        c.request("POST", "/v1/completions", json.dumps(body), {"Content-Type": "application/json"})
        r = c.getresponse(); d = json.loads(r.read()); c.close()
        return (time.perf_counter() - t) * 1e3, d
    body = {"model": "Qwen/Qwen3-0.6B", "prompt": "Who are you?"}
    one(body); one(body)
    lat = [one(body)[0] for _ in range(20)]
    lat128 = [one(dict(body, max_tokens=128, ignore_eos=True))[0] for _ in range(5)]
    print(json.dumps({"request": "llm-d-test.yaml completion (16 new tokens)", "latency_ms_p50": float(np.median(lat)), "latency_ms_min": min(lat),
                      "latency_ms_128_tokens_p50": float(np.median(lat128)), "ms_per_token_single_stream": (float(np.median(lat128)) - float(np.median(lat))) / 112}))
finally:
    proc.terminate()
    try: proc.wait(timeout=10)
    except Exception: proc.kill()
