"""Diagnostic (GPU): per-CTA start/end times of the decode-attention pass at the C2 shape."""
import importlib, sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from _util import ptr
pkg = importlib.import_module("aws-k8s-ansible-provisioner_b200")
lib = pkg.load_library()
B, S, nq, nkv, hd, bs = int(sys.argv[3]) if len(sys.argv) > 3 else 64, int(sys.argv[1]) if len(sys.argv) > 1 else 640, 16, 8, 128, 16
NOFLUSH = len(sys.argv) > 4
nblk = (S + bs - 1) // bs
pages = B * nblk + 1
g = torch.Generator().manual_seed(0)
kp = torch.randn(pages, nkv, bs, hd, generator=g).to(torch.bfloat16).cuda()
vp = torch.randn(pages, nkv, bs, hd, generator=g).to(torch.bfloat16).cuda()
perm = (torch.randperm(pages - 1, generator=g) + 1).to(torch.int32)
mode = sys.argv[2] if len(sys.argv) > 2 else "seq"
if mode == "seq":
    perm = torch.arange(1, pages, dtype=torch.int32)
bt = perm.view(B, nblk).contiguous().cuda()
sl = torch.full((B,), S, dtype=torch.int32).cuda()
q = torch.randn(B, nq * hd, generator=g).to(torch.bfloat16).cuda()
out = torch.empty(B, nq * hd, dtype=torch.bfloat16, device="cuda")
scratch = torch.zeros(lib.b2k_attn_decode_scratch_floats(B, nq, nkv, hd), dtype=torch.float32, device="cuda")
# several independent K/V sets, visited round-robin, so nothing read is L2-resident and no dirty
# lines have to be written back during the measurement (a memset flush would add write traffic)
NSETS = 1 if NOFLUSH else 4
sets = [(kp, vp)] + [(torch.randn_like(kp), torch.randn_like(vp)) for _ in range(NSETS - 1)]
times = []
for it in range(8):
    kp, vp = sets[it % NSETS]
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    rc = lib.b2k_attn_decode(ptr(q), ptr(kp), ptr(vp), ptr(bt), nblk, ptr(sl), ptr(scratch), ptr(out), B, nq, nkv, hd, bs, None)
    e1.record()
    torch.cuda.synchronize()
    assert rc == 0
    times.append(e0.elapsed_time(e1) * 1e3)
    if os.environ.get("PER_SM"):
        m = scratch.view(torch.int32)[:1024].cpu().numpy()
        d = (m[601:601 + 2 * 148:2].astype(np.int64) - m[600:600 + 2 * 148:2].astype(np.int64)) / 1e3
        per_sm = d[None, :] if it == 0 else np.concatenate([per_sm, d[None, :]], axis=0)
meta = scratch.view(torch.int32)[:1024].cpu().numpy()
st, en = meta[600:600 + 2 * 148:2].astype(np.int64), meta[601:601 + 2 * 148:2].astype(np.int64)
live = st > 0
st, en = st[live], en[live]
t0 = st.min()
dur = en - st
print("cfg", os.environ.get("B2_ATTN_CFG", "8x2"), "mode", mode, "B", B, "S", S, "event us (last 5): %s" % ["%.1f" % t for t in times[-5:]], "MB %.1f" % (B * S * 4096 / 1e6))
print("start skew us: max %.2f" % ((st.max() - t0) / 1e3), " end: min %.2f median %.2f max %.2f us" % ((en.min() - t0) / 1e3, np.median(en - t0) / 1e3, (en.max() - t0) / 1e3))
print("per-CTA duration us: min %.2f p10 %.2f median %.2f p90 %.2f max %.2f" % tuple(np.percentile(dur, [0, 10, 50, 90, 100]) / 1e3))
print("ideal at 6539.5 GB/s: %.2f us" % (B * S * 4096 / 6539.5e9 * 1e6))
if os.environ.get("DUMP"):
    allst = meta[600:600 + 2 * 148:2].astype(np.int64); allen = meta[601:601 + 2 * 148:2].astype(np.int64)
    d = (allen - allst) / 1e3
    print("per-CTA us by SM id (pairs share a TPC):", " ".join("%.1f" % x for x in d))
    print("start offsets us:", " ".join("%.1f" % ((x - allst[allst > 0].min()) / 1e3) for x in allst))

if os.environ.get("PER_SM"):
    mean, sd = per_sm[2:].mean(axis=0), per_sm[2:].std(axis=0)
    order = np.argsort(-mean)
    print("per-SM mean busy time over %d launches: overall mean %.2f; between-SM std %.2f; mean within-SM std %.2f" % (
        per_sm.shape[0] - 2, mean.mean(), mean.std(), sd.mean()))
    print("slowest SMs (id: mean+-sd):", " ".join("%d:%.1f+-%.1f" % (i, mean[i], sd[i]) for i in order[:16]))
    print("fastest SMs:", " ".join("%d:%.1f+-%.1f" % (i, mean[i], sd[i]) for i in order[-10:]))
    print("by id:", " ".join("%.1f" % x for x in mean))
