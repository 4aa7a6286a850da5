"""b2route (csrc/router/router.cpp), the load- and prefix-aware front for several model-server replicas
(SURVEY.md 8f rank 2), against fake replicas on localhost (CPU, no GPU): relaying of plain and streamed
responses, prefix affinity, load shedding from the scraped vllm:* series, fail-over and its own /metrics."""
import http.client
import http.server
import json
import os
import socket
import subprocess
import threading
import time

import pytest

from _util import ROOT

EXE = os.path.join(ROOT, "aws-k8s-ansible-provisioner_b200", "b2route")


class Replica:
    """A stand-in for b2serve: answers the routes the router relays and reports a settable queue length."""

    def __init__(self, name):
        self.name = name
        self.waiting = 0
        self.running = 0
        self.served = []
        outer = self

        class H(http.server.BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"

            def log_message(self, *a):
                pass

            def _send(self, code, body, ctype="application/json"):
                b = body.encode()
                self.send_response(code)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(b)))
                self.end_headers()
                self.wfile.write(b)

            def do_GET(self):
                if self.path == "/metrics":
                    self._send(200, "# TYPE vllm:num_requests_running gauge\nvllm:num_requests_running{model_name=\"m\"} %d\n"
                                    "vllm:num_requests_waiting{model_name=\"m\"} %d\n" % (outer.running, outer.waiting), "text/plain")
                elif self.path == "/v1/models":
                    self._send(200, json.dumps({"object": "list", "data": [{"id": "Qwen/Qwen3-0.6B", "served_by": outer.name}]}))
                elif self.path == "/health":
                    self._send(200, "", "text/plain")
                else:
                    self._send(404, "{}")

            def do_POST(self):
                n = int(self.headers.get("Content-Length", "0"))
                body = json.loads(self.rfile.read(n) or b"{}")
                outer.served.append(body)
                if body.get("stream"):
                    self.send_response(200)
                    self.send_header("Content-Type", "text/event-stream")
                    self.send_header("Transfer-Encoding", "chunked")
                    self.end_headers()
                    for i in range(5):
                        ev = ("data: " + json.dumps({"i": i, "by": outer.name}) + "\n\n").encode()
                        self.wfile.write(b"%x\r\n" % len(ev) + ev + b"\r\n")
                        self.wfile.flush()
                        time.sleep(0.01)
                    ev = b"data: [DONE]\n\n"
                    self.wfile.write(b"%x\r\n" % len(ev) + ev + b"\r\n0\r\n\r\n")
                else:
                    self._send(200, json.dumps({"object": "text_completion", "by": outer.name, "echo": body.get("prompt")}))

        self.httpd = http.server.ThreadingHTTPServer(("127.0.0.1", 0), H)
        self.port = self.httpd.server_address[1]
        self.thread = threading.Thread(target=self.httpd.serve_forever, daemon=True)
        self.thread.start()

    def stop(self):
        self.httpd.shutdown()
        self.httpd.server_close()


def call(port, method, path, body=None):
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=30)
    c.request(method, path, json.dumps(body) if body is not None else None, {"Content-Type": "application/json"})
    r = c.getresponse()
    data = r.read()
    c.close()
    return r.status, data


@pytest.fixture()
def cluster():
    if not os.path.exists(EXE):
        pytest.skip("b2route not built")
    reps = [Replica("r0"), Replica("r1"), Replica("r2")]
    proc = subprocess.Popen([EXE, "--port", "0", "--host", "127.0.0.1", "--scrape-interval-ms", "100", "--backends",
                             ",".join("127.0.0.1:%d" % r.port for r in reps)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    line = proc.stdout.readline()
    assert line.startswith("READY"), proc.stderr.read()
    port = int(line.split()[1])
    time.sleep(0.4)  # first scrape
    yield {"port": port, "reps": reps}
    proc.terminate()
    proc.wait(timeout=10)
    for r in reps:
        try:
            r.stop()
        except Exception:
            pass


def test_relays_plain_and_streamed_responses(cluster):
    port = cluster["port"]
    st, data = call(port, "GET", "/v1/models")
    assert st == 200 and b"Qwen/Qwen3-0.6B" in data          # the reference's probe, llm-d-test.yaml:32-59
    st, data = call(port, "POST", "/v1/completions", {"model": "m", "prompt": "Who are you?"})
    assert st == 200 and json.loads(data)["echo"] == "Who are you?"
    # SSE through the router: every event, in order, then [DONE]
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=30)
    c.request("POST", "/v1/completions", json.dumps({"model": "m", "prompt": "stream me", "stream": True}), {"Content-Type": "application/json"})
    r = c.getresponse()
    text = r.read().decode()
    c.close()
    events = [e[6:] for e in text.split("\n\n") if e.startswith("data: ")]
    assert events[-1] == "[DONE]" and [json.loads(e)["i"] for e in events[:-1]] == [0, 1, 2, 3, 4]
    assert call(port, "GET", "/health")[0] == 200


def test_prefix_affinity_and_spread(cluster):
    port, reps = cluster["port"], cluster["reps"]
    shared = "You are a helpful assistant. " * 12  # > 256 bytes: the key is the common start
    who = {json.loads(call(port, "POST", "/v1/completions", {"model": "m", "prompt": shared + "question %d" % i})[1])["by"] for i in range(8)}
    assert len(who) == 1, "requests sharing a prompt prefix were spread over %s" % who
    # token-id prompts and chat messages have prefix keys too
    ids = list(range(100))
    who = {json.loads(call(port, "POST", "/v1/completions", {"model": "m", "prompt": ids + [i]})[1])["by"] for i in range(6)}
    assert len(who) == 1
    # the replica holding the shared prefix reports a long queue: its prefix now goes elsewhere
    busy = next(r for r in reps if any(isinstance(b.get("prompt"), str) and b["prompt"].startswith(shared) for b in r.served))
    busy.waiting = 50
    time.sleep(0.4)
    j = json.loads(call(port, "POST", "/v1/completions", {"model": "m", "prompt": shared + "again"})[1])
    assert j["by"] != busy.name
    # unrelated prompts never pick the busy replica while others are idle
    who = [json.loads(call(port, "POST", "/v1/completions", {"model": "m", "prompt": "unrelated %d" % i})[1])["by"] for i in range(12)]
    assert busy.name not in who
    st, m = call(port, "GET", "/metrics")
    text = m.decode()
    assert "b2route_prefix_hits_total" in text and 'b2route_backend_up{backend="127.0.0.1:%d"} 1' % busy.port in text
    hits = int([l for l in text.splitlines() if l.startswith("b2route_prefix_hits_total")][0].split()[-1])
    assert hits >= 10


def test_failover_when_a_replica_dies(cluster):
    port, reps = cluster["port"], cluster["reps"]
    prompt = "sticky prefix " * 30
    first = json.loads(call(port, "POST", "/v1/completions", {"model": "m", "prompt": prompt})[1])["by"]
    dead = next(r for r in reps if r.name == first)
    dead.stop()
    # the very next request for that prefix is retried on another replica, not failed
    st, data = call(port, "POST", "/v1/completions", {"model": "m", "prompt": prompt})
    assert st == 200 and json.loads(data)["by"] != first
    time.sleep(0.4)
    text = call(port, "GET", "/metrics")[1].decode()
    assert 'b2route_backend_up{backend="127.0.0.1:%d"} 0' % dead.port in text
    for r in reps:
        if r is not dead:
            r.stop()
    time.sleep(0.4)
    assert call(port, "POST", "/v1/completions", {"model": "m", "prompt": "x"})[0] == 503
    assert call(port, "GET", "/health")[0] == 503
