"""Diagnostic (GPU): phase timeline of the tcgen05 projection kernel for the four decode GEMM shapes."""
import importlib, os, sys
import numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from _util import ptr
pkg = importlib.import_module("aws-k8s-ansible-provisioner_b200")
lib = pkg.load_library()
T = 64
trace = torch.zeros(148 * 8, dtype=torch.int64, device="cuda")
names = ["start", "setup", "dep-wait", "1st operands", "MMAs issued", "acc visible", "epilogue done"]
for (name, N, K, ns) in [("qkv", 4096, 1024, 4), ("o", 1024, 2048, 16), ("gate-up", 6144, 1024, 3), ("down", 1024, 3072, 16),
                         ("qkv ns=1", 4096, 1024, 1), ("gate-up ns=1", 6144, 1024, 1)]:
    x = torch.randn(T, K).to(torch.bfloat16).cuda()
    w = (torch.randn(N, K) * 0.05).to(torch.bfloat16).cuda()
    out = torch.empty(ns, T, N, dtype=torch.float32, device="cuda")
    flush = torch.empty(200 << 20, dtype=torch.uint8, device="cuda")
    lib.b2k_debug_gemm_trace(ptr(trace))
    for it in range(3):
        flush.copy_(flush)  # evict W from L2 (reads+writes)
        trace.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = lib.b2k_linear(ptr(x), ptr(w), ptr(out), 1, ns, T, N, K, 1, None)
        e1.record()
        torch.cuda.synchronize()
        assert rc == 0, lib.b2_last_error()
    lib.b2k_debug_gemm_trace(None)
    t = trace.cpu().numpy().reshape(148, 8)
    live = t[:, 0] > 0
    t = t[live]
    t0 = t[:, 0].min()
    print("%-14s N=%d K=%d splits=%d CTAs=%d  event %.1f us" % (name, N, K, ns, live.sum(), e0.elapsed_time(e1) * 1e3))
    for k in range(7):
        col = (t[:, k] - t0) / 1e3
        print("    %-14s median %6.2f  max %6.2f us" % (names[k], np.median(col), col.max()))
