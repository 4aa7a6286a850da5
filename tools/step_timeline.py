"""Diagnostic (GPU): in-situ timeline of one captured decode step at the bench shape (C2: batch 64, context ~640).

Every CTA of every kernel on the decode path stamps %globaltimer at its start, when its dependency wait returns and
at its end (b2k_debug_step_trace, csrc/common.cuh).  The records of one graph-replayed step are grouped into
launches and printed as (a) a per-kind summary -- launch duration, time from the predecessor's end to this launch's
dependency release ("hand-off"), time from the release to the last CTA's end ("work") -- and (b) the launch-by-launch
timeline of one layer in the middle of the model.  With the two-lane schedule it also reports how much of each lane's
chain runs while the other lane's attention pass is streaming.

    python tools/step_timeline.py [--mb 0|32] [--ctx 640] [--layer 14] [--out gpurun_out/step_timeline.json]
"""
import argparse
import importlib
import json
import os
import sys

import numpy as np
import torch

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
sys.path.insert(0, os.path.join(R, "tests"))
from _util import ptr  # noqa: E402

KINDS = {0: "embed", 1: "norm", 3: "rope_kv", 4: "attn", 8: "silu"}
GEMM = {(64, 16): "qkv", (16, 32): "o", (96, 16): "gate_up", (16, 48): "down"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=None, help="two-lane threshold (B2_MB); default = the engine's")
    ap.add_argument("--ctx", type=int, default=640)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--layer", type=int, default=14)
    ap.add_argument("--out", default=os.path.join(R, "gpurun_out", "step_timeline.json"))
    args = ap.parse_args()
    if args.mb is not None:
        os.environ["B2_MB"] = str(args.mb)
    trace_lib = os.path.join(R, "aws-k8s-ansible-provisioner_b200", "libb2engine_trace.so")
    assert os.path.exists(trace_lib), "build the trace variant first: tools/build_variant.sh trace -DB2_STEP_TRACE"
    os.environ["B2_LIB_PATH"] = trace_lib
    pkg = importlib.import_module("aws-k8s-ansible-provisioner_b200")
    weights = importlib.import_module("aws-k8s-ansible-provisioner_b200.weights")
    lib = pkg.load_library()
    B, P = args.batch, 512
    cfg = pkg.EngineConfig(max_model_len=args.ctx + 64, max_seqs=B, max_batch_tokens=8192)
    cfg.num_blocks = B * ((cfg.max_model_len + 15) // 16) + 8
    eng = pkg.Engine(cfg, device=0)
    eng.load_state_dict(weights.random_state_dict(cfg, seed=0))
    g = torch.Generator().manual_seed(1)
    prompts = torch.randint(0, cfg.vocab, (B, P), generator=g, dtype=torch.int32).numpy()
    sids = [eng.seq_create(cfg.max_model_len) for _ in range(B)]
    for i in range(0, B, 16):
        eng.prefill(sids[i:i + 16], [p.tolist() for p in prompts[i:i + 16]])
    eng.decode_run(sids, args.ctx - P - 3, want_tokens=False)
    cap = 400000
    buf = torch.zeros(4 + 4 * cap, dtype=torch.int64, device="cuda")
    buf[1] = cap
    torch.cuda.synchronize()
    assert lib.b2k_debug_step_trace(ptr(buf)) == 0
    _, ms = eng.decode_run(sids, 2, want_tokens=False)
    torch.cuda.synchronize()
    lib.b2k_debug_step_trace(None)
    raw = buf.cpu().numpy().view(np.uint64)
    n = int(min(raw[0], cap))
    rec = raw[4:4 + 4 * n].reshape(n, 4)
    cls = (rec[:, 0] & 0xff).astype(np.int64)
    smid = ((rec[:, 0] >> 32) & 0xff).astype(np.int64)
    aux = ((rec[:, 0] >> 40) & 0xffffff).astype(np.int64)
    t0, tw, t1 = rec[:, 1].astype(np.int64), rec[:, 2].astype(np.int64), rec[:, 3].astype(np.int64)
    # last step = records after the last embed launch started
    emb = np.sort(t0[cls == 0])
    starts = [emb[0]]
    for t in emb[1:]:
        if t - starts[-1] > 200000:
            starts.append(t)
    step0 = starts[-1]
    keep = t0 >= step0
    cls, smid, aux, t0, tw, t1 = cls[keep], smid[keep], aux[keep], t0[keep], tw[keep], t1[keep]
    bid = ((rec[:, 0] >> 8) & 0xffffff).astype(np.int64)[keep]
    # projection phase records (class 101): first operands landed / MMAs issued / accumulator visible, joined to the
    # CTA's main record (class 100) by (shape word, block index, nearest earlier wait)
    ph = cls == 101
    ph_aux, ph_bid, ph_a, ph_b, ph_c = aux[ph], bid[ph], t0[ph], tw[ph], t1[ph]
    main = ~ph
    cls, smid, aux, t0, tw, t1, bid = cls[main], smid[main], aux[main], t0[main], tw[main], t1[main], bid[main]
    phases = {}
    for a in sorted(set(ph_aux)):
        name = GEMM.get((a & 0xfff, (a >> 12) & 0x7ff))
        if name is None:
            continue
        m100 = (cls == 100) & (aux == a)
        rows = []
        for i in np.nonzero(ph_aux == a)[0]:
            cand = np.nonzero(m100 & (bid == ph_bid[i]) & (tw <= ph_a[i]) & (t1 >= ph_c[i]))[0]
            if len(cand):
                j = cand[np.argmax(tw[cand])]
                rows.append((ph_a[i] - tw[j], ph_b[i] - tw[j], ph_c[i] - tw[j], t1[j] - tw[j]))
        if rows:
            phases.setdefault(name, []).extend(rows)
    # ---- group records into launches ----
    names, lanes = [], []
    for c, a in zip(cls, aux):
        if c == 100:
            names.append(GEMM.get((a & 0xfff, (a >> 12) & 0x7ff), "lm_head" if (a & 0xfff) > 2000 else "gemm?"))
            lanes.append(int(a >> 23) & 1)
        else:
            names.append(KINDS.get(int(c), "k%d" % c))
            lanes.append(int(a) & 1)
    names, lanes = np.array(names), np.array(lanes)
    launches = []
    for key in sorted(set(zip(names, lanes))):
        m = (names == key[0]) & (lanes == key[1])
        idx = np.nonzero(m)[0]
        idx = idx[np.argsort(t0[idx])]
        group = [idx[0]]
        for i in idx[1:]:
            if t0[i] - t0[group[-1]] > 2500 and t0[i] > max(t1[j] for j in group) - 500:
                launches.append((key, group))
                group = [i]
            else:
                group.append(i)
        launches.append((key, group))
    L = []
    for (name, lane), grp in launches:
        grp = np.array(grp)
        L.append({"kind": name, "lane": int(lane), "ctas": len(grp), "start": int(t0[grp].min()), "start_max": int(t0[grp].max()),
                  "wait": int(np.median(tw[grp][tw[grp] > 0])) if (tw[grp] > 0).any() else int(t0[grp].min()),
                  "end": int(t1[grp].max()), "end_med": int(np.median(t1[grp]))})
    L.sort(key=lambda r: r["wait"])
    base = L[0]["start"]
    for r in L:
        for k in ("start", "start_max", "wait", "end", "end_med"):
            r[k] = (r[k] - base) / 1e3
    step_us = max(r["end"] for r in L)
    print("step: %.1f us by stamps, %.1f us by events (2 steps / 2); %d records, %d launches; two-lane=%s" % (
        step_us, ms * 1e3 / 2, len(cls), len(L), os.environ.get("B2_MB", "default")))
    # ---- attention: how the CTAs of one launch are spread in time ----
    att_idx = [grp for (name, lane), grp in launches if name == "attn"]
    if att_idx:
        rows = []
        for grp in att_idx:
            grp = np.array(grp)
            s0 = t0[grp].min()
            st, wt, en = (t0[grp] - s0) / 1e3, (tw[grp] - s0) / 1e3, (t1[grp] - s0) / 1e3
            busy = en - np.maximum(st, wt)
            rows.append([np.percentile(st, 50), st.max(), np.median(wt), np.percentile(en, 10), np.percentile(en, 50), np.percentile(en, 90), en.max(),
                         np.median(busy), np.corrcoef(st, en)[0, 1] if st.std() > 0 else 0.0])
        a = np.median(np.array(rows), axis=0)
        print("\nattention launches (median over %d launches; us from the launch's first CTA start): CTA start p50 %.2f max %.2f | wait returned %.2f | "
              "CTA end p10 %.2f p50 %.2f p90 %.2f max %.2f | CTA busy after wait p50 %.2f | corr(start, end) %.2f" % ((len(rows),) + tuple(a)))
        # is the tail made of the SAME CTAs every launch (block index -> SM placement is stable for a one-CTA-per-SM grid)?
        ncta = max(len(g) for g in att_idx)
        M = []
        for grp in att_idx:
            grp = np.array(grp)
            if len(grp) != ncta:
                continue
            v = np.full(ncta, np.nan)
            e = (t1[grp] - np.maximum(t0[grp], tw[grp])) / 1e3
            v[bid[grp] % ncta] = e - np.median(e)
            M.append(v)
        M = np.array(M)
        if len(M) >= 4 and not np.isnan(M).any():
            mu = M.mean(axis=0)
            print("attention CTA run time minus the launch's median, by block index over %d launches: std of the per-block means %.2f us, "
                  "mean within-block std %.2f us (total std %.2f); the 8 slowest blocks %s are late by %s us on average" % (
                      len(M), mu.std(), M.std(axis=0).mean(), M.std(), np.argsort(mu)[-8:][::-1].tolist(),
                      np.round(np.sort(mu)[-8:][::-1], 2).tolist()))
    # ---- per-lane chains: hand-off and work ----
    summary = {}
    for lane in sorted(set(r["lane"] for r in L)):
        chain = [r for r in L if r["lane"] == lane and r["kind"] not in ("embed", "lm_head")]
        for prev, cur in zip(chain[:-1], chain[1:]):
            d = summary.setdefault(cur["kind"], {"dur": [], "handoff": [], "work": [], "early": []})
            d["dur"].append(cur["end"] - cur["start"])
            d["handoff"].append(cur["wait"] - prev["end"])
            d["work"].append(cur["end"] - cur["wait"])
            d["early"].append(prev["end"] - cur["start"])
    print("\n| kind | launches | start->end | resident before predecessor ended | hand-off (pred end -> wait returns) | work (wait -> last CTA end) |")
    print("|---|---|---|---|---|---|")
    for k, d in summary.items():
        print("| %s | %d | %.2f | %.2f | %.2f | %.2f |" % (k, len(d["dur"]), np.median(d["dur"]), np.median(d["early"]),
                                                          np.median(d["handoff"]), np.median(d["work"])))
    per_layer = sum(np.median(d["handoff"]) + np.median(d["work"]) for k, d in summary.items()) + \
        (np.median(summary["norm"]["handoff"]) + np.median(summary["norm"]["work"]) if "norm" in summary else 0)
    print("sum over one layer's 9 launches (norm counted twice) of hand-off + work: %.1f us" % per_layer)
    if phases:
        print("\nprojection CTAs, time after their dependency wait returned (median / p90 over CTAs x layers, us):")
        print("(fused pair kernels, gemm_pair.cu: the three columns are accumulator visible | pair barrier passed | tail done)")
        print("| kind | first operands landed | MMAs issued | accumulator visible | CTA end |")
        print("|---|---|---|---|---|")
        for name, rows in phases.items():
            a = np.array(rows) / 1e3
            print("| %s | %s |" % (name, " | ".join("%.2f / %.2f" % (np.median(a[:, k]), np.percentile(a[:, k], 90)) for k in range(4))))
    # ---- one layer, launch by launch ----
    for lane in sorted(set(r["lane"] for r in L)):
        att = [r for r in L if r["lane"] == lane and r["kind"] == "attn"]
        if len(att) <= args.layer + 1:
            continue
        a0, a1 = att[args.layer]["start"], att[args.layer + 1]["end"]
        print("\nlane %d, layer %d (times in us from the start of that layer's attention pass):" % (lane, args.layer))
        print("| kind | lane | CTAs | first CTA start | wait returned | last CTA end | median CTA end |")
        print("|---|---|---|---|---|---|---|")
        for r in L:
            if r["wait"] >= a0 - 1 and r["end"] <= a1 + 1:
                print("| %s | %d | %d | %.2f | %.2f | %.2f | %.2f |" % (r["kind"], r["lane"], r["ctas"], r["start"] - a0, r["wait"] - a0,
                                                                       r["end"] - a0, r["end_med"] - a0))
    if len(set(r["lane"] for r in L)) == 2:
        att = [r for r in L if r["lane"] == 0 and r["kind"] == "attn"]
        if len(att) > args.layer + 2:
            a0, a1 = att[args.layer]["start"], att[args.layer + 2]["start"]
            print("\nboth lanes, two layers from lane 0's attention pass of layer %d (us):" % args.layer)
            print("| kind | lane | CTAs | first CTA start | last CTA start | wait returned | last CTA end |")
            print("|---|---|---|---|---|---|---|")
            for r in L:
                if r["wait"] >= a0 - 1 and r["wait"] <= a1:
                    print("| %s | %d | %d | %.2f | %.2f | %.2f | %.2f |" % (r["kind"], r["lane"], r["ctas"], r["start"] - a0, r["start_max"] - a0,
                                                                           r["wait"] - a0, r["end"] - a0))
    # ---- overlap between the lanes ----
    lanes_present = sorted(set(r["lane"] for r in L))
    if len(lanes_present) == 2:
        for lane in lanes_present:
            other = [r for r in L if r["lane"] != lane and r["kind"] == "attn"]
            mine = [r for r in L if r["lane"] == lane and r["kind"] not in ("attn", "embed", "lm_head")]
            tot = sum(r["end"] - r["wait"] for r in mine)
            ov = 0.0
            for r in mine:
                for o in other:
                    ov += max(0.0, min(r["end"], o["end"]) - max(r["wait"], o["wait"]))
            att = [r for r in L if r["lane"] == lane and r["kind"] == "attn"]
            print("lane %d: chain work %.0f us of which %.0f us while the other lane's attention runs; its own attention: %d passes, median %.1f us (wait->end)" % (
                lane, tot, ov, len(att), np.median([r["end"] - r["wait"] for r in att])))
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump({"step_us": step_us, "event_ms_per_step": ms / 2, "launches": L}, open(args.out, "w"))
    eng.close()


if __name__ == "__main__":
    main()
