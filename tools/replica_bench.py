"""Diagnostic (N GPUs): BASELINE config 5 in miniature -- one b2serve replica per GPU behind b2route, closed loop
of 64 requests in flight per replica (512-token id prompts with per-client shared prefixes, 128 new tokens).
Prints aggregate output tok/s, how the router spread the requests and its prefix-hit count."""
import http.client, json, os, subprocess, sys, threading, time
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(R, "aws-k8s-ansible-provisioner_b200")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2
P, NEW, INFLIGHT, ROUNDS = 512, 128, 64, 3

def wait_ready(proc):
    t0 = time.time()
    while time.time() - t0 < 180:
        line = proc.stdout.readline()
        if line.startswith("READY"):
            return int(line.split()[1])
        if proc.poll() is not None:
            break
    raise SystemExit("did not start: " + proc.stderr.read()[-2000:])

procs = []
try:
    ports = []
    for i in range(N):
        p = subprocess.Popen([os.path.join(PKG, "b2serve"), "--load-format", "dummy", "--served-model-name", "Qwen/Qwen3-0.6B", "--port", "0",
                              "--host", "127.0.0.1", "--device", str(i), "--max-model-len", "1024", "--max-num-seqs", "64",
                              "--eos-token-id", "-1"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        procs.append(p)
    for p in procs:
        ports.append(wait_ready(p))
    rt = subprocess.Popen([os.path.join(PKG, "b2route"), "--port", "0", "--host", "127.0.0.1", "--scrape-interval-ms", "100",
                           "--backends", ",".join("127.0.0.1:%d" % x for x in ports)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    procs.append(rt)
    rport = wait_ready(rt)
    time.sleep(0.5)
    rng = np.random.default_rng(1)
    clients = N * INFLIGHT
    # every client keeps its own 448-token prefix across its requests and varies the last 64 tokens
    prefixes = [rng.integers(0, 151936, P - 64).tolist() for _ in range(clients)]
    done = [0] * clients
    def client(i):
        for r in range(ROUNDS):
            body = {"model": "Qwen/Qwen3-0.6B", "prompt": prefixes[i] + rng.integers(0, 151936, 64).tolist(), "max_tokens": NEW,
                    "temperature": 0, "ignore_eos": True}
            c = http.client.HTTPConnection("127.0.0.1", rport, timeout=600)
            c.request("POST", "/v1/completions", json.dumps(body), {"Content-Type": "application/json"})
            resp = c.getresponse()
            j = json.loads(resp.read())
            c.close()
            assert resp.status == 200 and j["usage"]["completion_tokens"] == NEW, j
            done[i] += NEW
    th = [threading.Thread(target=client, args=(i,)) for i in range(clients)]
    t0 = time.perf_counter()
    for t in th: t.start()
    for t in th: t.join()
    dt = time.perf_counter() - t0
    c = http.client.HTTPConnection("127.0.0.1", rport, timeout=30)
    c.request("GET", "/metrics")
    m = c.getresponse().read().decode()
    c.close()
    routed = [int(l.split()[-1]) for l in m.splitlines() if l.startswith("b2route_requests_total")]
    hits = [int(l.split()[-1]) for l in m.splitlines() if l.startswith("b2route_prefix_hits_total")][0]
    print(json.dumps({"replicas": N, "clients": clients, "requests": clients * ROUNDS, "output_tok_s_incl_prefill": sum(done) / dt,
                      "seconds": dt, "routed_per_replica": routed, "prefix_hits": hits}), flush=True)
finally:
    for p in procs:
        p.terminate()
    for p in procs:
        try:
            p.wait(timeout=10)
        except Exception:
            p.kill()
