"""Host-side plumbing of bench.py for N > 1 (CPU, gloo, world_size 2): the replicas meet only at a
barrier and at the max/sum reductions the JSON line is built from; the reference arm runs on rank 0
alone.  No GPU and no engine involved."""
import json
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

RANK_SCRIPT = r'''
import json, os, sys, types
sys.path.insert(0, %(root)r)
import bench
rank, world, local, dist = bench.dist_setup(2)
assert world == 2 and dist is not None and local == rank
bench.barrier(dist)
# every rank reports its own elapsed time and token count; the line uses max(time) and sum(tokens)
ms = [10.0, 14.0][rank]
toks = [640, 640][rank]
ms_max = bench.reduce_max(dist, ms)
total = bench.reduce_sum(dist, toks)
bench.barrier(dist)
# the reference arm: rank 0 only
args = types.SimpleNamespace(steps=1, warmup=1, gpus=2)
ran = []
bench.hf_cpu_decode = lambda *a, **k: ran.append(rank) or {"tok_s": 1.0, "ms_per_step": 1.0, "steps": 1, "batch": 1, "ctx": 1, "threads": 1}
bench.run_reference(args, rank, world, dist)
print("RESULT " + json.dumps({"rank": rank, "ms_max": ms_max, "total": total, "ref_ran": ran}), flush=True)
dist.destroy_process_group()
'''


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.timeout(300)
def test_two_rank_reductions_and_reference_arm(tmp_path):
    script = tmp_path / "ranks.py"
    script.write_text(RANK_SCRIPT % {"root": ROOT})
    port = free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        out, err = p.communicate(timeout=240)
        assert p.returncode == 0, err[-2000:]
        outs.append(out)
    results = {}
    ref_lines = 0
    for out in outs:
        for line in out.splitlines():
            if line.startswith("RESULT "):
                r = json.loads(line[7:])
                results[r["rank"]] = r
            elif line.startswith("{"):
                j = json.loads(line)
                assert j["impl"] == "reference" and j["e2e"]["h2d_bytes_per_step"] == 0
                ref_lines += 1
    assert set(results) == {0, 1}
    for r in results.values():
        assert r["ms_max"] == 14.0 and r["total"] == 1280.0
    assert results[0]["ref_ran"] == [0] and results[1]["ref_ran"] == []
    assert ref_lines == 1, "exactly one rank prints the reference line"


def test_single_rank_needs_no_process_group(monkeypatch):
    sys.path.insert(0, ROOT)
    import bench
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    rank, world, local, dist = bench.dist_setup(1)
    assert (rank, world, local, dist) == (0, 1, 0, None)
    assert bench.reduce_max(None, 3.5) == 3.5 and bench.reduce_sum(None, 2.0) == 2.0
    bench.barrier(None)


def test_timed_window_is_pinned_to_c2_whatever_the_step_count():
    """--steps must not shrink the workload.  Every (K, W) is timed around C2's mean context
    (512-token prompts, 256 new tokens -> 640.5) and both arms print the same config."""
    sys.path.insert(0, ROOT)
    import bench
    for K, W in [(248, 8), (20, 5), (8, 3), (120, 4), (64, 3)]:
        lo, hi, advance = bench.timed_window(K, W)
        assert hi - lo == K and advance >= 0 and lo == bench.PROMPT + W + advance
        assert abs((lo + hi) / 2.0 - 640.5) <= 4.5, (K, W, lo, hi)
        cfg = bench.c2_config(1, K, W)
        assert cfg["new_tokens"] == 256 and cfg["timed_context"] == [lo, hi] and cfg["workload"] == "C2"
        assert "5.9" in cfg["l2"]  # 5.90 GB +- the window's centre: SURVEY.md section 8d
    # the reference arm prints exactly the GPU arm's config for the same flags
    import types
    got = []
    bench.hf_cpu_decode = lambda b, ctx, steps, warmup, budget_s: got.append((ctx, steps)) or {
        "tok_s": 1.0, "ms_per_step": 1.0, "steps": steps, "batch": b, "ctx": ctx, "threads": 1}
    import io, contextlib
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        bench.run_reference(types.SimpleNamespace(steps=20, warmup=5, gpus=1), 0, 1, None)
    line = json.loads(buf.getvalue())
    assert line["config"] == bench.c2_config(1, 20, 5) and line["steps"] == 20
    ctx, steps = got[0]
    assert 625 <= ctx <= 640 and steps == 20
