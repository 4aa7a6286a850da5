"""BASELINE.json config 5 (N GPUs): one b2serve replica per GPU behind b2route, closed loop of 64 requests in flight per
replica, request lengths prompt ~ U{64..4096}, output ~ U{64..4096} (seed 2), token-id prompts, greedy, ignore_eos.
Every client sends ROUNDS requests back to back through the router (/v1/completions, the route llm-d-test.yaml:61-73
exercises).  Prints one JSON line: aggregate output tok/s (wall clock from the first request to the last response),
per-replica spread, TPOT p50/p99 from the replicas' vllm:inter_token_latency_seconds histograms, host CPU use.

    python tools/replica_bench.py N [--policy prefix|least-loaded|round-robin] [--rounds 3] [--fixed]   (--fixed: the old
    miniature: 512-token prompts with per-client prefixes, 128 new tokens)"""
import argparse
import http.client
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(R, "aws-k8s-ansible-provisioner_b200")
MODEL = "Qwen/Qwen3-0.6B"


def wait_ready(proc):
    t0 = time.time()
    while time.time() - t0 < 300:
        line = proc.stdout.readline()
        if line.startswith("READY"):
            return int(line.split()[1])
        if proc.poll() is not None:
            break
    raise SystemExit("did not start: " + proc.stderr.read()[-2000:])


def get(port, path):
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=30)
    c.request("GET", path)
    body = c.getresponse().read().decode()
    c.close()
    return body


def histogram_quantiles(texts, name, qs):
    """Quantiles of a Prometheus histogram summed over replicas (upper bucket bounds: a conservative read)."""
    buckets = {}
    for t in texts:
        for l in t.splitlines():
            if l.startswith(name + "_bucket"):
                le = l.split('le="')[1].split('"')[0]
                buckets[le] = buckets.get(le, 0.0) + float(l.split()[-1])
    if not buckets:
        return {}
    items = sorted(((float("inf") if k == "+Inf" else float(k)), v) for k, v in buckets.items())
    total = items[-1][1]
    out = {}
    for q in qs:
        out[q] = next((le for le, c in items if c >= q * total), float("inf")) if total > 0 else None
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("n", type=int, nargs="?", default=2)
    ap.add_argument("--policy", default="prefix")
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--inflight", type=int, default=64)
    ap.add_argument("--fixed", action="store_true")
    ap.add_argument("--max-len", type=int, default=4096)
    ap.add_argument("--prefix-caching", action="store_true")
    args = ap.parse_args()
    N = args.n
    procs = []
    try:
        for i in range(N):
            cmd = [os.path.join(PKG, "b2serve"), "--load-format", "dummy", "--served-model-name", MODEL, "--port", "0", "--host", "127.0.0.1",
                   "--device", str(i), "--max-model-len", "1024" if args.fixed else str(2 * args.max_len), "--max-num-seqs", str(args.inflight),
                   "--gpu-memory-utilization", "0.5", "--eos-token-id", "-1"]
            if args.prefix_caching:
                cmd.append("--enable-prefix-caching")
            procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
        ports = [wait_ready(p) for p in procs]
        rt = subprocess.Popen([os.path.join(PKG, "b2route"), "--port", "0", "--host", "127.0.0.1", "--scrape-interval-ms", "100", "--policy",
                               args.policy, "--backends", ",".join("127.0.0.1:%d" % x for x in ports)], stdout=subprocess.PIPE,
                              stderr=subprocess.PIPE, text=True)
        procs.append(rt)
        rport = wait_ready(rt)
        time.sleep(0.5)
        rng = np.random.default_rng(2)
        clients = N * args.inflight
        if args.fixed:
            prefixes = [rng.integers(0, 151936, 448).tolist() for _ in range(clients)]
            plan = [[(prefixes[i] + rng.integers(0, 151936, 64).tolist(), 128) for _ in range(args.rounds)] for i in range(clients)]
        else:
            plan = [[(rng.integers(0, 151936, int(rng.integers(64, args.max_len + 1))).tolist(), int(rng.integers(64, args.max_len + 1)))
                     for _ in range(args.rounds)] for i in range(clients)]
        out_tokens = [0] * clients
        in_tokens = [0] * clients
        lat = [[] for _ in range(clients)]
        errors = []

        def client(i):
            for prompt, new in plan[i]:
                body = json.dumps({"model": MODEL, "prompt": prompt, "max_tokens": new, "temperature": 0, "ignore_eos": True})
                t0 = time.perf_counter()
                try:
                    c = http.client.HTTPConnection("127.0.0.1", rport, timeout=1800)
                    c.request("POST", "/v1/completions", body, {"Content-Type": "application/json"})
                    resp = c.getresponse()
                    j = json.loads(resp.read())
                    c.close()
                    if resp.status != 200 or j["usage"]["completion_tokens"] != new:
                        errors.append(str(j)[:200])
                        continue
                except Exception as ex:  # noqa: BLE001
                    errors.append(repr(ex))
                    continue
                lat[i].append(time.perf_counter() - t0)
                out_tokens[i] += new
                in_tokens[i] += len(prompt)

        try:
            import psutil
            psutil.cpu_percent(None)
        except ImportError:
            psutil = None
        th = [threading.Thread(target=client, args=(i,)) for i in range(clients)]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        dt = time.perf_counter() - t0
        cpu = psutil.cpu_percent(None) if psutil else None
        m = get(rport, "/metrics")
        routed = [int(float(l.split()[-1])) for l in m.splitlines() if l.startswith("b2route_requests_total")]
        hits = [int(float(l.split()[-1])) for l in m.splitlines() if l.startswith("b2route_prefix_hits_total")]
        reps = [get(p, "/metrics") for p in ports]
        tpot = histogram_quantiles(reps, "vllm:inter_token_latency_seconds", [0.5, 0.99])
        ttft = histogram_quantiles(reps, "vllm:time_to_first_token_seconds", [0.5, 0.99])
        gen = [sum(float(l.split()[-1]) for l in r.splitlines() if l.startswith("vllm:generation_tokens_total")) for r in reps]
        pc = [sum(float(l.split()[-1]) for l in r.splitlines() if l.startswith("vllm:prefix_cache_hits_total")) for r in reps]
        alll = sorted(x for l in lat for x in l)
        print(json.dumps({
            "workload": "C5 fixed miniature" if args.fixed else "C5: prompt, output ~ U{64..%d}, %d in flight per replica" % (args.max_len, args.inflight),
            "replicas": N, "policy": args.policy, "clients": clients, "requests": sum(len(l) for l in lat), "errors": len(errors),
            "first_errors": errors[:2], "seconds": dt, "output_tok_s": sum(out_tokens) / dt, "prompt_tok_s": sum(in_tokens) / dt,
            "output_tokens": sum(out_tokens), "prompt_tokens": sum(in_tokens), "generated_per_replica": gen, "routed_per_replica": routed,
            "router_prefix_hits": hits[0] if hits else None, "replica_prefix_cache_hit_tokens": pc,
            "tpot_s_p50_p99_bucket_upper_bounds": [tpot.get(0.5), tpot.get(0.99)], "ttft_s_p50_p99_bucket_upper_bounds": [ttft.get(0.5), ttft.get(0.99)],
            "request_latency_s_p50_p99": [alll[len(alll) // 2], alll[int(len(alll) * 0.99)]] if alll else None,
            "host_cpu_percent": cpu, "host_cpus": os.cpu_count()}), flush=True)
    finally:
        for p in procs:
            p.terminate()
        for p in procs:
            try:
                p.wait(timeout=15)
            except Exception:  # noqa: BLE001
                p.kill()


if __name__ == "__main__":
    main()
