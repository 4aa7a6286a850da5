"""HTTP contract of the model-server binary (GPU): replays the two calls of the reference's smoke test
(llm-d-test.yaml:32-43 GET /v1/models, :61-73 POST /v1/completions with {"model", "prompt"}) plus the
chat, streaming, metrics and error surface of SURVEY.md section 8b, against b2serve on localhost with a
scaled-down Qwen3 checkpoint; greedy completions are checked against the oracle."""
import http.client
import importlib
import json
import os
import subprocess
import time

import pytest
import torch

from _util import ROOT, pkg
from oracle.qwen3_oracle import QWEN3_TINY, Restated, attention_explicit, perturb_norm_weights, random_weights

pytestmark = pytest.mark.gpu
MODEL_ID = "Qwen/Qwen3-0.6B"  # served name, as in llm-d-test.yaml:7

TEMPLATE = ("{%- for m in messages -%}[{{ m['role'] }}] {{ m['content'] + '\\n' }}{%- endfor -%}"
            "{%- if add_generation_prompt -%}[assistant] {%- endif -%}")


@pytest.fixture(scope="module")
def server(tmp_path_factory):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    P = pkg()
    weights = importlib.import_module("aws-k8s-ansible-provisioner_b200.weights")
    exe = os.path.join(os.path.dirname(P.LIB_PATH), "b2serve")
    assert os.path.exists(exe), "b2serve not built"
    d = tmp_path_factory.mktemp("model")
    arch = QWEN3_TINY
    sd = random_weights(arch, seed=0)
    perturb_norm_weights(sd)
    weights.write_safetensors(sd, str(d / "model.safetensors"))
    (d / "config.json").write_text(json.dumps({
        "num_hidden_layers": arch.num_layers, "hidden_size": arch.hidden, "num_attention_heads": arch.num_q_heads,
        "num_key_value_heads": arch.num_kv_heads, "head_dim": arch.head_dim, "intermediate_size": arch.intermediate,
        "vocab_size": arch.vocab, "tie_word_embeddings": True, "rms_norm_eps": arch.rms_eps, "rope_theta": arch.rope_theta}))
    (d / "chat.jinja").write_text(TEMPLATE)
    proc = subprocess.Popen([exe, str(d), "--served-model-name", MODEL_ID, "--port", "0", "--host", "127.0.0.1", "--max-model-len", "512",
                             "--max-num-seqs", "16", "--num-gpu-blocks-override", "256", "--max-num-batched-tokens", "256",
                             "--chat-template", str(d / "chat.jinja"), "--eos-token-id", "-1"],
                            stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    port = None
    t0 = time.time()
    while time.time() - t0 < 120:
        line = proc.stdout.readline()
        if line.startswith("READY"):
            port = int(line.split()[1])
            break
        if proc.poll() is not None:
            break
    if port is None:
        proc.kill()
        pytest.fail("b2serve did not start: " + proc.stderr.read())
    yield {"port": port, "sd": sd, "arch": arch}
    proc.terminate()
    try:
        proc.wait(timeout=10)
    except Exception:
        proc.kill()


def call(port, method, path, body=None, raw=False):
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=60)
    hdr = {"accept": "application/json", "Content-Type": "application/json"}
    c.request(method, path, body=json.dumps(body) if body is not None else None, headers=hdr)
    r = c.getresponse()
    data = r.read().decode()
    c.close()
    return r.status, (data if raw else (json.loads(data) if data else None))


def test_models_lists_the_model_id(server):
    """llm-d-test.yaml:54-59: the only assertion in the reference -- the id occurs in the body."""
    st, raw = call(server["port"], "GET", "/v1/models", raw=True)
    assert st == 200 and MODEL_ID in raw
    j = json.loads(raw)
    assert j["object"] == "list" and j["data"][0]["id"] == MODEL_ID and j["data"][0]["object"] == "model"
    assert j["data"][0]["max_model_len"] == 512


def test_completion_as_the_reference_sends_it(server):
    """llm-d-test.yaml:61-73: {"model": ..., "prompt": "Who are you?"} -> defaults max_tokens=16, n=1."""
    st, j = call(server["port"], "POST", "/v1/completions", {"model": MODEL_ID, "prompt": "Who are you?"})
    assert st == 200, j
    assert j["object"] == "text_completion" and j["model"] == MODEL_ID and j["id"].startswith("cmpl-")
    ch = j["choices"]
    assert len(ch) == 1 and ch[0]["index"] == 0 and ch[0]["finish_reason"] == "length" and ch[0]["logprobs"] is None
    assert isinstance(ch[0]["text"], str)
    assert j["usage"] == {"prompt_tokens": len("Who are you?"), "total_tokens": len("Who are you?") + 16, "completion_tokens": 16}


def test_greedy_token_prompt_matches_oracle(server):
    arch, sd = server["arch"], server["sd"]
    g = torch.Generator().manual_seed(11)
    prompt = torch.randint(0, 256, (1, 29), generator=g)  # byte ids: the text of the reply is decodable
    st, j = call(server["port"], "POST", "/v1/completions",
                 {"model": MODEL_ID, "prompt": prompt[0].tolist(), "max_tokens": 12, "temperature": 0, "ignore_eos": True})
    assert st == 200, j
    ref, _ = Restated(arch, sd, attention=attention_explicit).greedy(prompt, 12)
    # the byte tokenizer renders ids >= 256 as <|id|>: recover those from the text and compare in order
    import re
    got_big = [int(x) for x in re.findall(r"<\|(\d+)\|>", j["choices"][0]["text"])]
    ref_big = [t for t in ref[0].tolist() if t >= 256]
    assert len(ref_big) >= 4, "test prompt should produce ids outside the byte range"
    assert got_big[:6] == ref_big[:6], (j["choices"][0]["text"], ref[0].tolist())
    assert j["usage"]["completion_tokens"] == 12


def test_batched_prompts_and_n(server):
    st, j = call(server["port"], "POST", "/v1/completions",
                 {"model": MODEL_ID, "prompt": ["a", "bc"], "max_tokens": 3, "n": 2, "temperature": 0.8, "seed": 5})
    assert st == 200, j
    assert [c["index"] for c in j["choices"]] == [0, 1, 2, 3]
    assert j["usage"]["prompt_tokens"] == 3 and j["usage"]["completion_tokens"] == 12
    st, j2 = call(server["port"], "POST", "/v1/completions", {"model": MODEL_ID, "prompt": [[1, 2, 3], [4, 5]], "max_tokens": 2, "temperature": 0})
    assert st == 200 and len(j2["choices"]) == 2


def test_chat_completion_and_template(server):
    body = {"model": MODEL_ID, "messages": [{"role": "system", "content": "Be brief."}, {"role": "user", "content": "Who are you?"}],
            "max_tokens": 5, "temperature": 0}
    st, j = call(server["port"], "POST", "/v1/chat/completions", body)
    assert st == 200, j
    assert j["object"] == "chat.completion" and j["id"].startswith("chatcmpl-")
    m = j["choices"][0]["message"]
    assert m["role"] == "assistant" and isinstance(m["content"], str)
    rendered = "[system] Be brief.\n[user] Who are you?\n[assistant]"  # the {%- endif strips the space before it
    assert j["usage"]["prompt_tokens"] == len(rendered.encode())
    assert j["usage"]["completion_tokens"] == 5 and j["choices"][0]["finish_reason"] == "length"


def test_streaming_sse(server):
    c = http.client.HTTPConnection("127.0.0.1", server["port"], timeout=60)
    c.request("POST", "/v1/chat/completions", body=json.dumps({"model": MODEL_ID, "messages": [{"role": "user", "content": "Hi"}],
                                                                "max_tokens": 6, "temperature": 0, "stream": True}),
              headers={"Content-Type": "application/json"})
    r = c.getresponse()
    assert r.status == 200 and r.getheader("Content-Type").startswith("text/event-stream")
    events = [ln[6:] for ln in r.read().decode().split("\n") if ln.startswith("data: ")]
    c.close()
    assert events[-1] == "[DONE]"
    chunks = [json.loads(e) for e in events[:-1]]
    assert all(ch["object"] == "chat.completion.chunk" for ch in chunks)
    assert chunks[0]["choices"][0]["delta"].get("role") == "assistant"
    assert chunks[-1]["choices"][0]["finish_reason"] == "length"
    streamed = "".join(ch["choices"][0]["delta"].get("content", "") for ch in chunks)
    st, j = call(server["port"], "POST", "/v1/chat/completions", {"model": MODEL_ID, "messages": [{"role": "user", "content": "Hi"}],
                                                                  "max_tokens": 6, "temperature": 0})
    assert streamed == j["choices"][0]["message"]["content"]


def test_stop_string(server):
    st, full = call(server["port"], "POST", "/v1/completions", {"model": MODEL_ID, "prompt": [7, 8, 9], "max_tokens": 8, "temperature": 0})
    text = full["choices"][0]["text"]
    assert len(text) > 4
    stop = text[3:6]
    st, j = call(server["port"], "POST", "/v1/completions", {"model": MODEL_ID, "prompt": [7, 8, 9], "max_tokens": 8, "temperature": 0, "stop": stop})
    assert j["choices"][0]["finish_reason"] == "stop" and j["choices"][0]["text"] == text[:text.find(stop)]


def test_errors(server):
    port = server["port"]
    st, j = call(port, "POST", "/v1/completions", {"model": "nope", "prompt": "x"})
    assert st == 404 and "does not exist" in j["error"]["message"]
    st, j = call(port, "POST", "/v1/completions", {"model": MODEL_ID})
    assert st == 400 and j["error"]["type"] == "invalid_request_error"
    st, j = call(port, "POST", "/v1/completions", {"model": MODEL_ID, "prompt": [999999]})
    assert st == 400
    st, j = call(port, "POST", "/v1/completions", {"model": MODEL_ID, "prompt": "x" * 600})
    assert st == 400 and "maximum model length" in j["error"]["message"]
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=30)
    c.request("POST", "/v1/chat/completions", body="{not json", headers={"Content-Type": "application/json"})
    assert c.getresponse().status == 400
    c.close()
    st, _ = call(port, "GET", "/nothing")
    assert st == 404


def test_health_and_metrics(server):
    port = server["port"]
    st, body = call(port, "GET", "/health", raw=True)
    assert st == 200 and body == ""
    call(port, "POST", "/v1/completions", {"model": MODEL_ID, "prompt": "abc", "max_tokens": 4, "temperature": 0})
    st, text = call(port, "GET", "/metrics", raw=True)
    assert st == 200
    for name in ("vllm:num_requests_running", "vllm:num_requests_waiting", "vllm:kv_cache_usage_perc", "vllm:gpu_cache_usage_perc",
                 "vllm:prompt_tokens_total", "vllm:generation_tokens_total", "vllm:request_success_total",
                 "vllm:time_to_first_token_seconds_bucket", "vllm:inter_token_latency_seconds_bucket", "vllm:e2e_request_latency_seconds_count"):
        assert name in text, name
    assert 'model_name="%s"' % MODEL_ID in text
    gen = [ln for ln in text.splitlines() if ln.startswith("vllm:generation_tokens_total")][0]
    assert int(gen.split()[-1]) >= 4


def test_concurrent_requests_share_the_batch(server):
    import threading
    port = server["port"]
    results = [None] * 12
    def work(i):
        results[i] = call(port, "POST", "/v1/completions", {"model": MODEL_ID, "prompt": [1 + i, 2, 3, 4][: 1 + i % 4], "max_tokens": 20, "temperature": 0})
    th = [threading.Thread(target=work, args=(i,)) for i in range(12)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert all(r is not None and r[0] == 200 and r[1]["usage"]["completion_tokens"] == 20 for r in results)
    # the same prompt alone must give the same greedy text as inside the concurrent batch (tie-aware: prefix)
    st, solo = call(port, "POST", "/v1/completions", {"model": MODEL_ID, "prompt": [1, 2, 3, 4][:1], "max_tokens": 20, "temperature": 0})
    assert solo["choices"][0]["text"][:8] == results[0][1]["choices"][0]["text"][:8]


def test_overload_cancellation_and_cleanup(server):
    """More requests than sequence slots (16) and than KV pages allow at once, ragged prompts longer than one
    prefill chunk (256 tokens), a mix of plain and streamed requests, and clients that hang up mid-stream: every
    surviving request completes with its full length, and afterwards nothing is running, waiting or holding pages."""
    import random
    import socket
    import threading
    port = server["port"]
    rnd = random.Random(7)
    N = 60
    specs = [(rnd.randint(1, 300), rnd.randint(1, 40), rnd.random() < 0.4) for _ in range(N)]  # (prompt len, new tokens, stream)
    results = [None] * N

    def plain(i, plen, ntok):
        results[i] = call(port, "POST", "/v1/completions", {"model": MODEL_ID, "prompt": [rnd.randrange(1, 1000) for _ in range(plen)],
                                                             "max_tokens": ntok, "temperature": 0, "ignore_eos": True})

    def streamed(i, plen, ntok, hang_up):
        body = json.dumps({"model": MODEL_ID, "prompt": [3 + (i % 50)] * plen, "max_tokens": ntok, "temperature": 0, "ignore_eos": True,
                           "stream": True}).encode()
        s = socket.create_connection(("127.0.0.1", port), timeout=60)
        s.sendall(b"POST /v1/completions HTTP/1.1\r\nHost: x\r\nContent-Type: application/json\r\nContent-Length: %d\r\n\r\n" % len(body) + body)
        buf = b""
        while b"data: [DONE]" not in buf:
            chunk = s.recv(65536)
            if not chunk:
                break
            buf += chunk
            if hang_up and buf.count(b"data: ") >= 2:
                break  # client disappears after the second event
        s.close()
        # (events can be fewer than tokens: a token that ends inside a UTF-8 sequence produces no text yet)
        results[i] = ("hung up" if hang_up else (b"data: [DONE]" in buf, buf.count(b"data: {")), ntok)

    th = []
    for i, (plen, ntok, stream) in enumerate(specs):
        if stream:
            th.append(threading.Thread(target=streamed, args=(i, plen, ntok, i % 3 == 0)))
        else:
            th.append(threading.Thread(target=plain, args=(i, plen, ntok)))
    [t.start() for t in th]
    [t.join(timeout=120) for t in th]
    assert not any(t.is_alive() for t in th), "a request never finished"
    for i, (plen, ntok, stream) in enumerate(specs):
        r = results[i]
        assert r is not None
        if not stream:
            assert r[0] == 200 and r[1]["usage"]["completion_tokens"] == ntok and r[1]["usage"]["prompt_tokens"] == plen, r
        elif r[0] != "hung up":
            done, events = r[0]
            assert done and 1 <= events <= ntok + 1, "streamed request %d: done=%s, %d events for %d tokens" % (i, done, events, ntok)
    # drained: the hang-ups were noticed and their sequences released
    deadline = time.time() + 20
    while time.time() < deadline:
        st, text = call(port, "GET", "/metrics", raw=True)
        vals = {ln.split("{")[0]: float(ln.split()[-1]) for ln in text.splitlines()
                if ln.startswith(("vllm:num_requests_running", "vllm:num_requests_waiting", "vllm:kv_cache_usage_perc"))}
        if vals.get("vllm:num_requests_running") == 0 and vals.get("vllm:num_requests_waiting") == 0 and vals.get("vllm:kv_cache_usage_perc") == 0:
            break
        time.sleep(0.2)
    assert vals.get("vllm:num_requests_running") == 0 and vals.get("vllm:num_requests_waiting") == 0, vals
    assert vals.get("vllm:kv_cache_usage_perc") == 0, vals
    # and the server still answers
    st, j = call(port, "POST", "/v1/completions", {"model": MODEL_ID, "prompt": "Who are you?", "max_tokens": 3})
    assert st == 200 and j["usage"]["completion_tokens"] == 3


# ---- with a tokenizer.json next to the weights (what the reference's model download provides) ---------
@pytest.fixture(scope="module")
def server_bpe(tmp_path_factory):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    tokenizers = pytest.importorskip("tokenizers")
    from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers, trainers
    P = pkg()
    weights = importlib.import_module("aws-k8s-ansible-provisioner_b200.weights")
    exe = os.path.join(os.path.dirname(P.LIB_PATH), "b2serve")
    d = tmp_path_factory.mktemp("model_bpe")
    arch = QWEN3_TINY
    sd = random_weights(arch, seed=0)
    weights.write_safetensors(sd, str(d / "model.safetensors"))
    (d / "config.json").write_text(json.dumps({
        "num_hidden_layers": arch.num_layers, "hidden_size": arch.hidden, "num_attention_heads": arch.num_q_heads,
        "num_key_value_heads": arch.num_kv_heads, "head_dim": arch.head_dim, "intermediate_size": arch.intermediate,
        "vocab_size": arch.vocab, "tie_word_embeddings": True, "rms_norm_eps": arch.rms_eps, "rope_theta": arch.rope_theta}))
    pattern = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+")
    tok = Tokenizer(models.BPE())
    tok.pre_tokenizer = pre_tokenizers.Sequence([pre_tokenizers.Split(Regex(pattern), behavior="isolated"),
                                                 pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
    tok.decoder = decoders.ByteLevel()
    with open(os.path.join(ROOT, "SURVEY.md"), encoding="utf-8") as f:
        text = f.read()
    tok.train_from_iterator([text], trainers.BpeTrainer(vocab_size=arch.vocab - 24, special_tokens=["<|im_start|>", "<|im_end|>"],
                                                        initial_alphabet=pre_tokenizers.ByteLevel.alphabet(), show_progress=False))
    tok.save(str(d / "tokenizer.json"))
    proc = subprocess.Popen([exe, str(d), "--served-model-name", MODEL_ID, "--port", "0", "--host", "127.0.0.1", "--max-model-len", "512",
                             "--max-num-seqs", "8", "--num-gpu-blocks-override", "128", "--max-num-batched-tokens", "256",
                             "--eos-token-id", "-1"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    port = None
    t0 = time.time()
    while time.time() - t0 < 120:
        line = proc.stdout.readline()
        if line.startswith("READY"):
            port = int(line.split()[1])
            break
        if proc.poll() is not None:
            break
    if port is None:
        proc.kill()
        pytest.fail("b2serve did not start: " + proc.stderr.read())
    yield {"port": port, "tok": tok}
    proc.terminate()
    try:
        proc.wait(timeout=10)
    except Exception:
        proc.kill()


def test_text_prompts_go_through_the_bpe_tokenizer(server_bpe):
    tok, port = server_bpe["tok"], server_bpe["port"]
    prompt = "Who are you? The model server's forward pass, 2026."
    st, j = call(port, "POST", "/v1/completions", {"model": MODEL_ID, "prompt": prompt, "max_tokens": 8, "temperature": 0})
    assert st == 200, j
    assert j["usage"]["prompt_tokens"] == len(tok.encode(prompt).ids) < len(prompt)
    # the same prompt as ids gives the same greedy continuation, and the text is those ids decoded
    st, j2 = call(port, "POST", "/v1/completions", {"model": MODEL_ID, "prompt": tok.encode(prompt).ids, "max_tokens": 8, "temperature": 0})
    assert st == 200 and j2["choices"][0]["text"] == j["choices"][0]["text"]
    # default chat template: the <|im_start|> / <|im_end|> markers are single special tokens, not spelled out
    msgs = [{"role": "user", "content": "Who are you?"}]
    st, j3 = call(port, "POST", "/v1/chat/completions", {"model": MODEL_ID, "messages": msgs, "max_tokens": 4, "temperature": 0})
    assert st == 200, j3
    rendered = "<|im_start|>user\nWho are you?<|im_end|>\n<|im_start|>assistant\n"
    assert j3["usage"]["prompt_tokens"] == len(tok.encode(rendered).ids)
    assert j3["usage"]["prompt_tokens"] < 20


def test_starts_with_the_argv_the_image_and_the_values_file_give():
    """The launch boundary (llm-d-deploy.yaml:147-151,176-184): the pod runs deploy/Dockerfile's ENTRYPOINT + CMD with
    the chart's extraArgs (deploy/values-b2serve.yaml) appended.  Exactly that argv -- only the binary's path is the
    in-tree build instead of /opt/b2/bin -- must come up on port 8000, with no checkpoint at /models (random-init
    weights of the served architecture, as BASELINE.json prescribes), and answer the reference's two requests."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import re
    import yaml
    P = pkg()
    txt = open(os.path.join(ROOT, "deploy", "Dockerfile")).read()
    ep = json.loads(re.search(r"^ENTRYPOINT\s+(\[.*\])\s*$", txt, re.M).group(1))
    cmd = json.loads(re.search(r"^CMD\s+(\[.*\])\s*$", txt, re.M).group(1))
    extra = yaml.safe_load(open(os.path.join(ROOT, "deploy", "values-b2serve.yaml")))["sampleApplication"]["decode"]["extraArgs"]
    assert ep == ["/opt/b2/bin/b2serve"]
    exe = os.path.join(os.path.dirname(P.LIB_PATH), "b2serve")
    proc = subprocess.Popen([exe] + cmd + extra, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    try:
        port = None
        t0 = time.time()
        while time.time() - t0 < 300:
            line = proc.stdout.readline()
            if line.startswith("READY"):
                port = int(line.split()[1])
                break
            if proc.poll() is not None:
                break
        assert port == 8000, "b2serve did not come up on the image's port: " + (proc.stderr.read()[-1500:] if proc.poll() is not None else "")
        st, raw = call(port, "GET", "/v1/models", raw=True)
        assert st == 200 and MODEL_ID in raw and json.loads(raw)["data"][0]["max_model_len"] == 4096
        st, j = call(port, "POST", "/v1/completions", {"model": MODEL_ID, "prompt": "Who are you?"})
        assert st == 200 and j["usage"]["completion_tokens"] == 16 and j["model"] == MODEL_ID
        st, m = call(port, "GET", "/metrics", raw=True)
        assert st == 200 and "vllm:kv_cache_usage_perc" in m
    finally:
        proc.terminate()
        try:
            proc.wait(timeout=20)
        except Exception:
            proc.kill()
