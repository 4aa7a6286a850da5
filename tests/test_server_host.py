"""CPU-side checks of the server's host logic: the C++ mini-Jinja behind `--chat-template` reproduces
the golden renderings of the reference's templates/*.yaml (`template.jinja` key), and the binary
refuses to start without a GPU instead of falling back."""
import importlib
import json
import os
import subprocess

import pytest

from _util import ROOT

GOLD = os.path.join(ROOT, "tests", "golden", "chat_templates.json")
REF_TEMPLATES = "/root/reference/templates"


@pytest.fixture(scope="module")
def b2serve(pkg, lib):
    path = os.path.join(os.path.dirname(pkg.LIB_PATH), "b2serve")
    if not os.path.exists(path):
        importlib.import_module("aws-k8s-ansible-provisioner_b200.build").build()
    assert os.path.exists(path), "b2serve was not built"
    return path


@pytest.mark.parametrize("name", ["opt-chat-template", "phi-chat-template"])
def test_chat_template_goldens(b2serve, name):
    """Renders the reference's own ConfigMap files (read in place, never copied into this repository)
    with the server's Jinja engine and compares with the committed jinja2 renderings."""
    path = os.path.join(REF_TEMPLATES, name + ".yaml")
    if not os.path.exists(path):
        pytest.skip("reference templates are only mounted in the build container")
    fx = json.load(open(GOLD))[name]
    for case in fx["cases"]:
        ctx = json.dumps({"messages": case["messages"], "add_generation_prompt": case["add_generation_prompt"]})
        r = subprocess.run([b2serve, "--render-chat-template", path], input=ctx.encode(), capture_output=True, timeout=20)
        assert r.returncode == 0, r.stderr.decode()
        assert r.stdout.decode() == case["rendered"], case["messages"]


def test_mini_jinja_features(b2serve, tmp_path):
    tpl = ("{% for m in messages %}{{ loop.index }}:{{ m.role }}={{ m['content'][0:3] }}{% if not loop.last %}|{% endif %}{% endfor %}"
           "{% if messages[0]['role'] != 'system' and add_generation_prompt %} >{{ 1 + 2 }}{{ 'a' + 'b' }}{% else %} no{% endif %}")
    p = tmp_path / "t.jinja"
    p.write_text(tpl)
    ctx = {"messages": [{"role": "user", "content": "hello"}, {"role": "assistant", "content": "yo"}], "add_generation_prompt": True}
    r = subprocess.run([b2serve, "--render-chat-template", str(p)], input=json.dumps(ctx).encode(), capture_output=True, timeout=20)
    assert r.returncode == 0, r.stderr.decode()
    assert r.stdout.decode() == "1:user=hel|2:assistant=yo >3ab"
    bad = tmp_path / "bad.jinja"
    bad.write_text("{% if x %}unterminated")
    r = subprocess.run([b2serve, "--render-chat-template", str(bad)], input=b"{}", capture_output=True, timeout=20)
    assert r.returncode != 0 and b"endif" in r.stderr


def test_server_refuses_to_start_without_gpu(b2serve):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = subprocess.run([b2serve, "--load-format", "dummy", "--port", "0"], capture_output=True, timeout=60)
    assert r.returncode != 0
    assert b"no CPU path" in r.stderr or b"CUDA" in r.stderr
