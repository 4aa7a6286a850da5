"""Diagnostic (GPU): soak b2serve (dummy Qwen3-0.6B) with waves of ragged requests for a few minutes; every request must
complete, and at the end nothing may be running, waiting or holding KV pages."""
import http.client, json, os, random, subprocess, sys, threading, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(R, "aws-k8s-ansible-provisioner_b200", "b2serve")
SECONDS = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
SAMPLED = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3  # share of requests with temperature > 0
proc = subprocess.Popen([EXE, "--load-format", "dummy", "--served-model-name", "Qwen/Qwen3-0.6B", "--port", "0", "--host", "127.0.0.1",
                         "--max-model-len", "2048", "--max-num-seqs", "64", "--eos-token-id", "-1"],
                        stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
try:
    port = None
    while True:
        line = proc.stdout.readline()
        if line.startswith("READY"):
            port = int(line.split()[1]); break
        if proc.poll() is not None:
            raise SystemExit(proc.stderr.read()[-2000:])
    rnd = random.Random(3)
    stats = {"ok": 0, "tokens": 0, "bad": 0}
    lock = threading.Lock()
    t_end = time.time() + SECONDS
    def client(i):
        r = random.Random(i)
        while time.time() < t_end:
            plen, ntok = r.randint(1, 1500), r.randint(1, 200)
            body = {"model": "Qwen/Qwen3-0.6B", "prompt": [r.randrange(0, 151936) for _ in range(plen)], "max_tokens": ntok,
                    "temperature": 0.8 if r.random() < SAMPLED else 0, "top_p": 0.9, "ignore_eos": True}
            c = http.client.HTTPConnection("127.0.0.1", port, timeout=300)
            c.request("POST", "/v1/completions", json.dumps(body), {"Content-Type": "application/json"})
            resp = c.getresponse(); j = json.loads(resp.read()); c.close()
            with lock:
                if resp.status == 200 and j["usage"]["completion_tokens"] == ntok and j["usage"]["prompt_tokens"] == plen:
                    stats["ok"] += 1; stats["tokens"] += ntok
                else:
                    stats["bad"] += 1
    th = [threading.Thread(target=client, args=(i,)) for i in range(96)]  # more clients than sequence slots
    t0 = time.time()
    [t.start() for t in th]; [t.join() for t in th]
    dt = time.time() - t0
    time.sleep(0.5)
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=30); c.request("GET", "/metrics"); m = c.getresponse().read().decode(); c.close()
    vals = {ln.split("{")[0]: float(ln.split()[-1]) for ln in m.splitlines()
            if ln.startswith(("vllm:num_requests_running", "vllm:num_requests_waiting", "vllm:kv_cache_usage_perc"))}
    print(json.dumps({"seconds": dt, "requests_ok": stats["ok"], "requests_bad": stats["bad"], "output_tok_s": stats["tokens"] / dt, "after": vals,
                      "server_alive": proc.poll() is None}))
finally:
    proc.terminate()
    try: proc.wait(timeout=10)
    except Exception: proc.kill()
