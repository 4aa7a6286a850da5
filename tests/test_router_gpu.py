"""b2route in front of REAL replicas (GPU): the path the reference's smoke test takes is gateway -> model-server pod
(llm-d-test.yaml:14-26 resolves the gateway, :32-43 and :61-73 send the two requests through it).  Two b2serve
processes (one per GPU when the box has two, else both on GPU 0) behind the router: the routed greedy completion must
be the text a replica gives directly and the ids the oracle gives, a streamed reply must arrive complete and in order, and
when the replica that holds a prompt's prefix is killed the next request for it is served by the other one."""
import http.client
import importlib
import json
import os
import re
import subprocess
import time

import pytest
import torch

from _util import ROOT, pkg
from oracle.qwen3_oracle import QWEN3_TINY, Restated, attention_explicit, perturb_norm_weights, random_weights

pytestmark = pytest.mark.gpu
MODEL_ID = "Qwen/Qwen3-0.6B"


def wait_ready(proc, what):
    t0 = time.time()
    while time.time() - t0 < 120:
        line = proc.stdout.readline()
        if line.startswith("READY"):
            return int(line.split()[1])
        if proc.poll() is not None:
            break
    proc.kill()
    pytest.fail("%s did not start: %s" % (what, proc.stderr.read()[-2000:]))


@pytest.fixture(scope="module")
def cluster(tmp_path_factory):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    P = pkg()
    weights = importlib.import_module("aws-k8s-ansible-provisioner_b200.weights")
    bindir = os.path.dirname(P.LIB_PATH)
    d = tmp_path_factory.mktemp("model")
    arch = QWEN3_TINY
    sd = random_weights(arch, seed=0)
    perturb_norm_weights(sd)
    weights.write_safetensors(sd, str(d / "model.safetensors"))
    (d / "config.json").write_text(json.dumps({
        "num_hidden_layers": arch.num_layers, "hidden_size": arch.hidden, "num_attention_heads": arch.num_q_heads,
        "num_key_value_heads": arch.num_kv_heads, "head_dim": arch.head_dim, "intermediate_size": arch.intermediate,
        "vocab_size": arch.vocab, "tie_word_embeddings": True, "rms_norm_eps": arch.rms_eps, "rope_theta": arch.rope_theta}))
    ngpu = torch.cuda.device_count()
    reps = []
    for i in range(2):
        reps.append(subprocess.Popen([os.path.join(bindir, "b2serve"), str(d), "--served-model-name", MODEL_ID, "--port", "0", "--host",
                                      "127.0.0.1", "--device", str(i % ngpu), "--max-model-len", "512", "--max-num-seqs", "16",
                                      "--num-gpu-blocks-override", "256", "--max-num-batched-tokens", "256", "--eos-token-id", "-1",
                                      "--enable-prefix-caching"],
                                     stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    ports = [wait_ready(p, "b2serve") for p in reps]
    rt = subprocess.Popen([os.path.join(bindir, "b2route"), "--port", "0", "--host", "127.0.0.1", "--scrape-interval-ms", "100", "--backends",
                           ",".join("127.0.0.1:%d" % x for x in ports)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    rport = wait_ready(rt, "b2route")
    time.sleep(0.4)  # first scrape of the replicas' /metrics
    yield {"router": rport, "ports": ports, "procs": reps, "sd": sd, "arch": arch}
    for p in reps + [rt]:
        if p.poll() is None:
            p.terminate()
    for p in reps + [rt]:
        try:
            p.wait(timeout=10)
        except Exception:
            p.kill()


def call(port, method, path, body=None, raw=False):
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=60)
    c.request(method, path, json.dumps(body) if body is not None else None, {"accept": "application/json", "Content-Type": "application/json"})
    r = c.getresponse()
    data = r.read().decode()
    c.close()
    return r.status, (data if raw else (json.loads(data) if data else None))


def test_reference_probe_through_the_router(cluster):
    """llm-d-test.yaml:32-59 and :61-73, sent to the router instead of the pod."""
    st, raw = call(cluster["router"], "GET", "/v1/models", raw=True)
    assert st == 200 and MODEL_ID in raw
    st, j = call(cluster["router"], "POST", "/v1/completions", {"model": MODEL_ID, "prompt": "Who are you?"})
    assert st == 200 and j["object"] == "text_completion" and j["usage"]["completion_tokens"] == 16


def test_routed_equals_direct_equals_oracle(cluster):
    arch, sd = cluster["arch"], cluster["sd"]
    g = torch.Generator().manual_seed(11)
    prompt = torch.randint(0, 256, (1, 29), generator=g)  # the prompt of test_server_gpu.py: its greedy stream leaves the byte range
    body = {"model": MODEL_ID, "prompt": prompt[0].tolist(), "max_tokens": 12, "temperature": 0, "ignore_eos": True}
    st, routed = call(cluster["router"], "POST", "/v1/completions", body)
    assert st == 200, routed
    texts = []
    for p in cluster["ports"]:
        st, direct = call(p, "POST", "/v1/completions", body)
        assert st == 200
        texts.append(direct["choices"][0]["text"])
    assert texts[0] == texts[1] == routed["choices"][0]["text"]
    ref, _ = Restated(arch, sd, attention=attention_explicit).greedy(prompt, 12)
    got_big = [int(x) for x in re.findall(r"<\|(\d+)\|>", routed["choices"][0]["text"])]
    ref_big = [t for t in ref[0].tolist() if t >= 256]
    assert len(ref_big) >= 4 and got_big[:6] == ref_big[:6], (routed["choices"][0]["text"], ref[0].tolist())


def sse_events(port, body):
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=60)
    c.request("POST", "/v1/completions", json.dumps(body), {"Content-Type": "application/json"})
    r = c.getresponse()
    text = r.read().decode()
    c.close()
    assert r.status == 200
    return [e[6:] for e in text.split("\n\n") if e.startswith("data: ")]


def test_stream_relayed_in_order_and_complete(cluster):
    body = {"model": MODEL_ID, "prompt": list(range(40, 60)), "max_tokens": 10, "temperature": 0, "ignore_eos": True, "stream": True}
    via = sse_events(cluster["router"], body)
    direct = sse_events(cluster["ports"][0], body)
    assert via[-1] == "[DONE]" and direct[-1] == "[DONE]" and len(via) >= 2

    def text_of(events):  # b2serve may put several tokens into one event when its writer lags: compare the stream's text
        js = [json.loads(e) for e in events[:-1]]
        assert all(j["object"] == "text_completion" and j["model"] == MODEL_ID for j in js)
        assert js[-1]["choices"][0]["finish_reason"] == "length" and all(j["choices"][0]["finish_reason"] is None for j in js[:-1])
        return "".join(j["choices"][0]["text"] for j in js)
    assert text_of(via) == text_of(direct) and len(text_of(via)) > 0


def test_prefix_affinity_lands_on_cached_pages(cluster):
    """What the router's prefix policy is for: requests that share a prompt prefix are sent to the replica that served it
    last, and that replica (started with --enable-prefix-caching) prefills only the new tail -- its
    vllm:prefix_cache_hits_total moves, the other replica's does not, and the text is what a cold replica gives."""
    port = cluster["router"]
    shared = list(range(300, 300 + 256))

    def hits():
        out = []
        for p in cluster["ports"]:
            m = call(p, "GET", "/metrics", raw=True)[1]
            out.append(sum(float(l.split()[-1]) for l in m.splitlines() if l.startswith("vllm:prefix_cache_hits_total")))
        return out
    h0 = hits()
    texts = []
    for k in range(4):
        body = {"model": MODEL_ID, "prompt": shared + [900 + k, 901 + k, 902 + k], "max_tokens": 6, "temperature": 0, "ignore_eos": True}
        st, j = call(port, "POST", "/v1/completions", body)
        assert st == 200
        texts.append(j["choices"][0]["text"])
    h1 = hits()
    gained = [b - a for a, b in zip(h0, h1)]
    # 240, not 256, tokens per request: the replica prefills in chunks of --max-num-batched-tokens 256 and a prefill call
    # always computes at least its own last token
    assert sorted(gained) == [0.0, 3 * 240.0], "three of the four requests should have found the prefix cached on ONE replica: %s" % gained
    # the same request sent straight to the replica that never saw the prefix: same greedy text (cached pages hold the same bits)
    cold = cluster["ports"][gained.index(0.0)]
    body = {"model": MODEL_ID, "prompt": shared + [903, 904, 905], "max_tokens": 6, "temperature": 0, "ignore_eos": True}
    assert call(cold, "POST", "/v1/completions", body)[1]["choices"][0]["text"] == texts[3]


def test_failover_to_the_surviving_replica(cluster):
    port = cluster["router"]
    prompt = list(range(7, 7 + 200))  # long enough for the router's prefix key
    body = {"model": MODEL_ID, "prompt": prompt, "max_tokens": 4, "temperature": 0, "ignore_eos": True}
    before = call(port, "POST", "/v1/completions", body)[1]["choices"][0]["text"]
    m = call(port, "GET", "/metrics", raw=True)[1]
    routed = {l.split('"')[1]: int(float(l.split()[-1])) for l in m.splitlines() if l.startswith("b2route_requests_total")}
    busiest = max(routed, key=routed.get)
    victim = cluster["ports"].index(int(busiest.split(":")[1]))
    cluster["procs"][victim].kill()
    cluster["procs"][victim].wait(timeout=10)
    for _ in range(3):  # the prefix's home is gone: retried on the other replica, same greedy text
        st, j = call(port, "POST", "/v1/completions", body)
        assert st == 200 and j["choices"][0]["text"] == before
    time.sleep(0.4)
    m = call(port, "GET", "/metrics", raw=True)[1]
    assert 'b2route_backend_up{backend="%s"} 0' % busiest in m
