"""The launch boundary (SURVEY.md 8b "Launch"; /root/reference llm-d-deploy.yaml:147-151 copies the chart's slim base to
custom-base.yaml, :176-184 hands it to llmd-installer.sh): deploy/values-b2serve.yaml must survive a deep merge into
such a base file, and the argv the image's ENTRYPOINT + CMD plus the values file's extraArgs produce must be one that
b2serve's own flag parser accepts.  helm/docker/yq are absent here: the merge is yq's `. * $i` (maps merge key by key,
everything else is replaced) done with pyyaml, against a stand-in base shaped like the chart's published sample."""
import json
import os
import re
import subprocess

import pytest
import yaml

from _util import ROOT

DEPLOY = os.path.join(ROOT, "deploy")

# A stand-in for quickstart/examples/base/slim/base-slim.yaml (the chart is cloned at `version: main` and is not
# vendored in the reference): the keys the override touches plus neighbours it must leave alone.
BASE = """
sampleApplication:
  enabled: true
  baseConfigMapRefName: basic-gpu-preset
  model:
    modelArtifactURI: hf://meta-llama/Llama-3.2-3B-Instruct
    modelName: "meta-llama/Llama-3.2-3B-Instruct"
  prefill:
    replicas: 1
  decode:
    replicas: 1
    extraArgs: ["--enforce-eager"]
modelservice:
  epp:
    defaultEnvVarsOverride: []
  vllm:
    image:
      registry: ghcr.io
      repository: llm-d/llm-d
      tag: "0.0.8"
    metrics:
      enabled: false
redis:
  enabled: false
"""


def deep_merge(a, b):
    if isinstance(a, dict) and isinstance(b, dict):
        out = dict(a)
        for k, v in b.items():
            out[k] = deep_merge(a[k], v) if k in a else v
        return out
    return b


def dockerfile_argv():
    txt = open(os.path.join(DEPLOY, "Dockerfile")).read()
    ep = json.loads(re.search(r"^ENTRYPOINT\s+(\[.*\])\s*$", txt, re.M).group(1))
    cmd = json.loads(re.search(r"^CMD\s+(\[.*\])\s*$", txt, re.M).group(1))
    return ep, cmd


def test_values_override_survives_the_merge():
    base = yaml.safe_load(BASE)
    over = yaml.safe_load(open(os.path.join(DEPLOY, "values-b2serve.yaml")))
    m = deep_merge(base, over)
    # what the override is for
    assert m["sampleApplication"]["model"]["modelName"] == "Qwen/Qwen3-0.6B"  # llm-d-deploy.yaml:118, llm-d-test.yaml:7
    assert m["sampleApplication"]["model"]["modelArtifactURI"] == "hf://Qwen/Qwen3-0.6B"
    assert m["modelservice"]["vllm"]["image"]["repository"] == "b2serve"
    assert m["modelservice"]["vllm"]["metrics"]["enabled"] is True
    assert m["sampleApplication"]["prefill"]["replicas"] == 0 and m["sampleApplication"]["decode"]["replicas"] == 1
    ann = m["modelservice"]["decode"]["podAnnotations"]
    # otel-observability-setup.yaml:345-368: scrape annotation, path and port the collector keys on
    assert ann == {"prometheus.io/scrape": "true", "prometheus.io/path": "/metrics", "prometheus.io/port": "8000"}
    # a list is replaced, not appended to: the chart's own extraArgs do not leak into the b2serve pod
    assert "--enforce-eager" not in m["sampleApplication"]["decode"]["extraArgs"]
    # what it must leave alone
    assert m["sampleApplication"]["baseConfigMapRefName"] == "basic-gpu-preset" and m["redis"] == {"enabled": False}
    assert m["modelservice"]["epp"] == {"defaultEnvVarsOverride": []}
    # it is still plain YAML a chart can take: a round trip changes nothing
    assert yaml.safe_load(yaml.safe_dump(m)) == m


def test_image_argv_is_one_b2serve_parses():
    """ENTRYPOINT + CMD + extraArgs, as the pod would run them.  Without a GPU the process must get past its flag
    parser and stop at the engine (no CUDA device, no CPU path) -- not at an unknown flag."""
    ep, cmd = dockerfile_argv()
    over = yaml.safe_load(open(os.path.join(DEPLOY, "values-b2serve.yaml")))
    extra = over["sampleApplication"]["decode"]["extraArgs"]
    assert ep == ["/opt/b2/bin/b2serve"] and "--served-model-name" in cmd and cmd[cmd.index("--port") + 1] == "8000"
    exe = os.path.join(ROOT, "aws-k8s-ansible-provisioner_b200", "b2serve")
    if not os.path.exists(exe):
        pytest.skip("b2serve not built")
    import torch
    if torch.cuda.is_available():
        pytest.skip("covered with a real start in tests/test_server_gpu.py")
    r = subprocess.run([exe] + cmd + extra, capture_output=True, text=True, timeout=120)
    assert "unknown flag" not in r.stderr and "needs a value" not in r.stderr, r.stderr
    assert r.returncode != 0 and re.search(r"CUDA|device|GPU", r.stderr), r.stderr[-500:]
