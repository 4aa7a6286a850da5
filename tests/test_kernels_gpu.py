"""Kernel-level parity (GPU): every hand-written kernel is called through the C ABI
(include/b2_engine.h, b2k_*) on seeded inputs and compared with the oracle's function for the
same HF lines.  Bars: elementwise kernels reproduce the oracle's rounding points, so outputs must
be within 1 bf16 ulp with only a small fraction differing at all (fp32 reduction order);
GEMM/attention are fp32-accumulate with a different summation order, bar = 1-2 bf16 ulp."""
import ctypes as C

import pytest
import torch

from _util import assert_bf16_close, cuda_stream_ptr, kv_swizzle, pkg, ptr, rel_err, rel_fro
from oracle.qwen3_oracle import (QWEN3_0_6B, apply_rope, attention_explicit, rms_norm, rope_cos_sin)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return pkg().load_library()


_KEEP = []


def dv(t: torch.Tensor) -> torch.Tensor:
    """Device copy that stays alive until the test ends (a temporary passed as a raw pointer would
    be returned to the caching allocator before the kernel runs)."""
    d = t.cuda()
    _KEEP.append(d)
    return d


@pytest.fixture(autouse=True)
def _release_device_copies():
    yield
    torch.cuda.synchronize()
    _KEEP.clear()


def chk(lib, rc):
    assert rc == 0, (rc, lib.b2_last_error())
    torch.cuda.synchronize()


def bf(*shape, scale=1.0, seed=0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16)


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode", ["none", "bf16", "f32x3"])
@pytest.mark.parametrize("H", [1024, 256])
def test_add_rmsnorm(lib, mode, H):
    T = 37
    h = bf(T, H, seed=1)
    w = (1 + 0.25 * torch.randn(H, generator=torch.Generator().manual_seed(2))).to(torch.bfloat16)
    h_d, w_d = h.cuda(), w.cuda()
    x_d = torch.empty_like(h_d)
    if mode == "none":
        chk(lib, lib.b2k_add_rmsnorm(ptr(h_d), None, 0, 0, ptr(w_d), ptr(x_d), T, H, 1e-6, cuda_stream_ptr()))
        h_ref = h
    elif mode == "bf16":
        d = bf(T, H, scale=0.5, seed=3)
        d_d = d.cuda()
        chk(lib, lib.b2k_add_rmsnorm(ptr(h_d), ptr(d_d), 0, 1, ptr(w_d), ptr(x_d), T, H, 1e-6, cuda_stream_ptr()))
        h_ref = h + d
    else:
        parts = torch.randn(3, T, H, generator=torch.Generator().manual_seed(4)) * 0.3
        p_d = parts.cuda()
        chk(lib, lib.b2k_add_rmsnorm(ptr(h_d), ptr(p_d), 1, 3, ptr(w_d), ptr(x_d), T, H, 1e-6, cuda_stream_ptr()))
        o = ((parts[0] + parts[1]) + parts[2]).to(torch.bfloat16)  # split order, one rounding (nn.Linear output)
        h_ref = h + o
    assert torch.equal(h_d.cpu(), h_ref), "residual stream must be bit-exact"
    x_ref = rms_norm(h_ref, w, 1e-6)
    assert_bf16_close(x_d, x_ref, 1, 0.01, "rmsnorm " + mode)


@pytest.mark.parametrize("f32in", [False, True])
def test_qknorm_rope_kv(lib, f32in):
    a = QWEN3_0_6B
    T, nq, nkv, hd, bs = 19, a.num_q_heads, a.num_kv_heads, a.head_dim, 16
    nblocks = 12
    qkv = bf(T, (nq + 2 * nkv) * hd, seed=5)
    g = torch.Generator().manual_seed(6)
    qw = (1 + 0.25 * torch.randn(hd, generator=g)).to(torch.bfloat16)
    kw = (1 + 0.25 * torch.randn(hd, generator=g)).to(torch.bfloat16)
    positions = torch.randint(0, 4000, (T,), generator=g, dtype=torch.int32)
    slots = torch.randperm(nblocks * bs, generator=g)[:T].to(torch.int32)
    slots[3] = -1  # padding token: no cache write
    maxpos = 4096
    cos, sin = rope_cos_sin(a, torch.arange(maxpos)[None, :])
    table = torch.cat([cos[0, :, :64], sin[0, :, :64]], dim=-1).contiguous()  # [pos][cos64 | sin64]
    kp = torch.zeros(nblocks, nkv, bs, hd, dtype=torch.bfloat16, device="cuda")
    vp = torch.zeros_like(kp)
    q_out = torch.empty(T, nq * hd, dtype=torch.bfloat16, device="cuda")
    if f32in:
        parts = torch.randn(2, T, (nq + 2 * nkv) * hd, generator=g)
        src = parts.cuda()
        qkv = (parts[0] + parts[1]).to(torch.bfloat16)
        rc = lib.b2k_qknorm_rope_kv(ptr(src), 1, 2, ptr(dv(qw)), ptr(dv(kw)), ptr(dv(table)), ptr(dv(positions)),
                                    ptr(dv(slots)), ptr(q_out), ptr(kp), ptr(vp), T, nq, nkv, hd, bs, 1e-6, cuda_stream_ptr())
    else:
        src = qkv.cuda()
        rc = lib.b2k_qknorm_rope_kv(ptr(src), 0, 1, ptr(dv(qw)), ptr(dv(kw)), ptr(dv(table)), ptr(dv(positions)),
                                    ptr(dv(slots)), ptr(q_out), ptr(kp), ptr(vp), T, nq, nkv, hd, bs, 1e-6, cuda_stream_ptr())
    chk(lib, rc)
    q = qkv[:, : nq * hd].view(1, T, nq, hd)
    k = qkv[:, nq * hd: (nq + nkv) * hd].view(1, T, nkv, hd)
    v = qkv[:, (nq + nkv) * hd:].view(T, nkv, hd)
    qn = rms_norm(q, qw, 1e-6).transpose(1, 2)
    kn = rms_norm(k, kw, 1e-6).transpose(1, 2)
    c, s = rope_cos_sin(a, positions[None, :].long())
    qr, kr = apply_rope(qn, kn, c, s)
    q_ref = qr.transpose(1, 2).reshape(T, nq * hd)
    k_ref = kr.transpose(1, 2)[0]  # [T][nkv][hd]
    assert_bf16_close(q_out, q_ref, 1, 0.01, "q rope")
    kp_c, vp_c = kv_swizzle(kp.cpu()), kv_swizzle(vp.cpu())  # stored (chunk-swizzled) -> logical
    for t in range(T):
        sl = int(slots[t])
        if sl < 0:
            continue
        assert_bf16_close(kp_c[sl // bs, :, sl % bs, :], k_ref[t], 1, 0.02, "k page t=%d" % t)
        assert torch.equal(vp_c[sl // bs, :, sl % bs, :], v[t]), "v page must be a bit-exact copy"
    written = {int(s_) for s_ in slots if s_ >= 0}
    for sl in range(nblocks * bs):
        if sl not in written:
            assert not kp_c[sl // bs, :, sl % bs, :].any(), "stray K write at slot %d" % sl


def test_qknorm_rope_kv_prefill_sized(lib):
    """Prefill-sized token counts take a different kernel (four heads per warp): it must agree BIT FOR BIT with the
    one-head-per-warp kernel on the same inputs (same arithmetic, same rounding points), pages included."""
    a = QWEN3_0_6B
    nq, nkv, hd, bs = a.num_q_heads, a.num_kv_heads, a.head_dim, 16
    T = 1100  # 1100 x 32 heads >= 32768: the prefill variant
    nblocks = (T + bs - 1) // bs + 2
    g = torch.Generator().manual_seed(16)
    qkv = bf(T, (nq + 2 * nkv) * hd, seed=15)
    qw = (1 + 0.25 * torch.randn(hd, generator=g)).to(torch.bfloat16)
    kw = (1 + 0.25 * torch.randn(hd, generator=g)).to(torch.bfloat16)
    positions = torch.randint(0, 4000, (T,), generator=g, dtype=torch.int32)
    slots = torch.randperm(nblocks * bs, generator=g)[:T].to(torch.int32)
    slots[7] = -1
    cos, sin = rope_cos_sin(a, torch.arange(4096)[None, :])
    table = torch.cat([cos[0, :, :64], sin[0, :, :64]], dim=-1).contiguous()
    src, dqw, dkw, dtab, dpos, dsl = qkv.cuda(), dv(qw), dv(kw), dv(table), dv(positions), dv(slots)
    outs = []
    for whole in (True, False):  # whole batch through the prefill kernel, then in slices small enough for the other one
        kp = torch.zeros(nblocks, nkv, bs, hd, dtype=torch.bfloat16, device="cuda")
        vp = torch.zeros_like(kp)
        q_out = torch.zeros(T, nq * hd, dtype=torch.bfloat16, device="cuda")
        spans = [(0, T)] if whole else [(i, min(T, i + 500)) for i in range(0, T, 500)]
        for (i0, i1) in spans:
            rc = lib.b2k_qknorm_rope_kv(ptr(src[i0:i1]), 0, 1, ptr(dqw), ptr(dkw), ptr(dtab), ptr(dpos[i0:i1]), ptr(dsl[i0:i1]),
                                        ptr(q_out[i0:i1]), ptr(kp), ptr(vp), i1 - i0, nq, nkv, hd, bs, 1e-6, cuda_stream_ptr())
            chk(lib, rc)
        torch.cuda.synchronize()
        outs.append((q_out.cpu(), kp.cpu(), vp.cpu()))
    for x, y, what in zip(outs[0], outs[1], ("q", "k pages", "v pages")):
        assert torch.equal(x, y), "prefill-sized kernel differs from the per-head kernel in " + what


@pytest.mark.parametrize("f32in", [False, True])
def test_silu_mul(lib, f32in):
    T, I = 23, 3072
    g = torch.Generator().manual_seed(7)
    if f32in:
        parts = torch.randn(3, T, 2 * I, generator=g)
        gu = ((parts[0] + parts[1]) + parts[2]).to(torch.bfloat16)
        src = parts.cuda()
    else:
        gu = (torch.randn(T, 2 * I, generator=g) * 2).to(torch.bfloat16)
        src = gu.cuda()
    act = torch.empty(T, I, dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_silu_mul(ptr(src), int(f32in), 3 if f32in else 1, ptr(act), T, I, cuda_stream_ptr()))
    ref = torch.nn.functional.silu(gu[:, :I]) * gu[:, I:]
    assert_bf16_close(act, ref, 1, 0.01, "silu_mul")


# ---------------------------------------------------------------------------------------------
LIN_SHAPES = [(1, 1024, 1024), (3, 4096, 1024), (16, 1024, 2048), (64, 6144, 1024), (64, 1024, 3072), (100, 256, 512),
              (300, 1024, 1024), (513, 1024, 256)]


@pytest.mark.parametrize("impl", [0, 1])
@pytest.mark.parametrize("T,N,K", LIN_SHAPES)
def test_linear_bf16(lib, impl, T, N, K):
    if impl == 0 and T > 64:
        pytest.skip("SIMT GEMV path is for tiny batches")
    x = bf(T, K, seed=8)
    w = bf(N, K, scale=0.05, seed=9)
    out = torch.full((T, N), float("nan"), dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_linear(ptr(dv(x)), ptr(dv(w)), ptr(out), 0, 1, T, N, K, impl, cuda_stream_ptr()))
    ref32 = x.float() @ w.float().t()
    assert_bf16_close(out, ref32.to(torch.bfloat16), 1, 0.02, "linear impl=%d %s" % (impl, (T, N, K)), abs_floor=1e-3)
    # and against the oracle's own op (HF nn.Linear on CPU)
    ref_hf = torch.nn.functional.linear(x, w)
    assert_bf16_close(out, ref_hf, 1, 0.03, "linear vs F.linear", abs_floor=1e-3)


@pytest.mark.parametrize("T,N,K,ns", [(64, 1024, 2048, 16), (64, 4096, 1024, 4), (7, 1024, 3072, 16), (64, 6144, 1024, 3),
                                     (200, 1024, 1024, 2)])
def test_linear_splitk_partials(lib, T, N, K, ns):
    x = bf(T, K, seed=10)
    w = bf(N, K, scale=0.05, seed=11)
    out = torch.full((ns, T, N), float("nan"), dtype=torch.float32, device="cuda")
    chk(lib, lib.b2k_linear(ptr(dv(x)), ptr(dv(w)), ptr(out), 1, ns, T, N, K, 1, cuda_stream_ptr()))
    assert torch.isfinite(out).all()
    total = out.sum(0).cpu()
    ref = x.float() @ w.float().t()
    assert rel_err(total, ref) < 2e-5
    # each split covers its own k-range: partial s == x[:, ks] @ w[:, ks]^T
    kb = K // 64
    kps = -(-kb // ns)
    for s in (0, ns - 1):
        k0, k1 = s * kps * 64, min(K, (s + 1) * kps * 64)
        part = x[:, k0:k1].float() @ w[:, k0:k1].float().t()
        assert rel_err(out[s].cpu(), part) < 2e-5, "split %d" % s


@pytest.mark.parametrize("T", [64, 33, 32, 5, 1])
def test_gateup_swiglu_fused(lib, T):
    """gemm_pair.cu: gate/up projection + SwiGLU in one launch (a pair of CTAs per tile splits K and exchanges halves of
    the accumulator through distributed shared memory).  Against the two-kernel path's arithmetic: g | u = bf16(x . W^T),
    act = bf16(bf16(silu(g)) * u).  The projection's fp32 sum is formed in another order than torch's, so g and u may be
    off by one bf16 ulp now and then, which the product can turn into two."""
    I, K = 3072, 1024
    x = bf(T, K, seed=21)
    w = bf(2 * I, K, scale=0.05, seed=22)
    act = torch.full((T, I), float("nan"), dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_gateup_swiglu(ptr(dv(x)), ptr(dv(w)), ptr(act), T, I, K, cuda_stream_ptr()))
    gu = (x.float() @ w.float().t()).to(torch.bfloat16)
    ref = torch.nn.functional.silu(gu[:, :I]) * gu[:, I:]
    assert_bf16_close(act, ref, 2, 0.05, "fused gate/up + SwiGLU, T=%d" % T, abs_floor=2e-3)
    # bit-identical run to run
    act2 = torch.empty_like(act)
    chk(lib, lib.b2k_gateup_swiglu(ptr(dv(x)), ptr(dv(w)), ptr(act2), T, I, K, cuda_stream_ptr()))
    assert torch.equal(act, act2)
    # and against the library's own two-kernel path (same rounding points; only the order of the fp32 sum differs)
    g2 = torch.empty(T, 2 * I, dtype=torch.bfloat16, device="cuda")
    a2 = torch.empty(T, I, dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_linear(ptr(dv(x)), ptr(dv(w)), ptr(g2), 0, 1, T, 2 * I, K, 1, cuda_stream_ptr()))
    chk(lib, lib.b2k_silu_mul(ptr(g2), 0, 1, ptr(a2), T, I, cuda_stream_ptr()))
    assert_bf16_close(act, a2, 2, 0.02, "fused vs separate kernels", abs_floor=2e-3)


def _ssq_pieces(h, ld=64):
    """[H/16][ld] fp32: per token column the sum of squares of each 16-feature piece of h's rows (what gemm_resid.cu writes)."""
    T, H = h.shape
    out = torch.zeros(H // 16, ld)
    out[:, :T] = (h.float() ** 2).view(T, H // 16, 16).sum(-1).t()
    return out


@pytest.mark.parametrize("T", [64, 40, 32, 5])
@pytest.mark.parametrize("K", [2048, 3072])
def test_linear_resid(lib, T, K):
    """gemm_resid.cu: O / down projection that adds into the residual stream itself (split-K summed across a cluster of
    8 through distributed shared memory).  The residual stream must be BIT-identical to the path it replaces (b2k_linear
    with 8 fp32 partials, then b2k_add_rmsnorm's sum in split order + add); the sum-of-squares pieces must be those of
    the new rows."""
    N = 1024
    x = bf(T, K, seed=41)
    w = bf(N, K, scale=0.03, seed=42)
    h0 = bf(T + 2, N, seed=43)
    gain = torch.ones(N, dtype=torch.bfloat16)
    dx, dw = dv(x), dv(w)
    h_new = h0.cuda()
    ssq = torch.full((N // 16, 64), float("nan"), device="cuda")
    chk(lib, lib.b2k_linear_resid(ptr(dx), ptr(dw), ptr(h_new), ptr(ssq), T, N, K, 64, cuda_stream_ptr()))
    parts = torch.empty(8, T, N, device="cuda")
    chk(lib, lib.b2k_linear(ptr(dx), ptr(dw), ptr(parts), 1, 8, T, N, K, 1, cuda_stream_ptr()))
    h_ref = h0.cuda()
    xn = torch.empty(T, N, dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_add_rmsnorm(ptr(h_ref), ptr(parts), 1, 8, ptr(dv(gain)), ptr(xn), T, N, 1e-6, cuda_stream_ptr()))
    assert torch.equal(h_new[:T].cpu(), h_ref[:T].cpu()), "residual stream differs from the partials + add_rmsnorm path"
    assert torch.equal(h_new[T:].cpu(), h0[T:]), "rows past T were touched"
    want = _ssq_pieces(h_new[:T].cpu())
    got = ssq.cpu()
    assert torch.allclose(got[:, :T], want[:, :T], rtol=1e-5, atol=1e-6)
    assert torch.isnan(got[:, T:]).all(), "wrote statistics for tokens past T"
    # bit-identical run to run
    h2 = h0.cuda()
    ssq2 = torch.empty_like(ssq)
    chk(lib, lib.b2k_linear_resid(ptr(dx), ptr(dw), ptr(h2), ptr(ssq2), T, N, K, 64, cuda_stream_ptr()))
    assert torch.equal(h2.cpu(), h_new.cpu()) and torch.equal(ssq2[:, :T].cpu(), got[:, :T])


@pytest.mark.parametrize("T,N,K", [(512, 256, 256), (8192, 4096, 1024), (8192, 1024, 3072), (1000, 1024, 2048), (32768, 1024, 2048)])
def test_linear_multicast(lib, T, N, K):
    """gemm_mc.cu: prefill projection by CTA pairs that share the token tile through TMA multicast.  Same k order, one
    accumulator per output: BIT-identical to b2k_linear's bf16 output; ragged last token tile; nothing written past row T."""
    x = bf(T, K, seed=81)
    w = bf(N, K, scale=0.03, seed=82)
    dx, dw = dv(x), dv(w)
    y = torch.full((T + 2, N), float("nan"), dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_linear_multicast(ptr(dx), ptr(dw), ptr(y), T, N, K, cuda_stream_ptr()))
    y2 = torch.empty(T, N, dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_linear(ptr(dx), ptr(dw), ptr(y2), 0, 1, T, N, K, 1, cuda_stream_ptr()))
    assert torch.equal(y[:T].cpu(), y2.cpu()), "multicast pair kernel differs from the single-CTA kernel"
    assert torch.isnan(y[T:].float()).all(), "wrote past the last token row"
    rows = slice(0, min(T, 256))
    ref = (x[rows].float() @ w.float().t())
    assert rel_fro(y[rows].float().cpu(), ref) < 5e-3
    y3 = torch.empty(T, N, dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_linear_multicast(ptr(dx), ptr(dw), ptr(y3), T, N, K, cuda_stream_ptr()))
    assert torch.equal(y3.cpu(), y[:T].cpu()), "two launches differ"


@pytest.mark.parametrize("T,N,K", [(300, 1024, 2048), (8192, 1024, 3072), (33, 256, 512), (1000, 3072, 8192)])
def test_linear_add_prefill(lib, T, N, K):
    """gemm_tc.cu with a residual operand: the prefill-sized O / down projection adds into the residual stream in its
    epilogue.  Bit-identical to b2k_linear (bf16 out) followed by b2k_add_rmsnorm's add."""
    x = bf(T, K, seed=71)
    w = bf(N, K, scale=0.03, seed=72)
    h0 = bf(T + 1, N, seed=73)
    dx, dw = dv(x), dv(w)
    h_new = h0.cuda()
    chk(lib, lib.b2k_linear_add(ptr(dx), ptr(dw), ptr(h_new), T, N, K, cuda_stream_ptr()))
    y = torch.empty(T, N, dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_linear(ptr(dx), ptr(dw), ptr(y), 0, 1, T, N, K, 1, cuda_stream_ptr()))
    h_ref = h0.cuda()
    xn = torch.empty(T, N, dtype=torch.bfloat16, device="cuda")
    gain = torch.ones(N, dtype=torch.bfloat16)
    chk(lib, lib.b2k_add_rmsnorm(ptr(h_ref), ptr(y), 0, 1, ptr(dv(gain)), ptr(xn), T, N, 1e-6, cuda_stream_ptr()))
    assert torch.equal(h_new[:T].cpu(), h_ref[:T].cpu()), "residual add in the epilogue differs from projection + add kernel"
    assert torch.equal(h_new[T:].cpu(), h0[T:]), "rows past T were touched"


@pytest.mark.parametrize("T", [64, 33, 32, 5])
def test_gateup_swiglu_normalise_on_load(lib, T):
    """gemm_pair.cu with PairNorm: the projection reads the residual stream and applies the RMSNorm while staging its
    token rows.  Against b2k_add_rmsnorm followed by b2k_gateup_swiglu: the statistic is summed in another order, so rstd
    may differ in its last bit and an element of x by one bf16 ulp now and then."""
    I, K = 3072, 1024
    h = bf(T, K, seed=51)
    w = bf(2 * I, K, scale=0.05, seed=52)
    gain = (1 + 0.25 * torch.randn(K, generator=torch.Generator().manual_seed(53))).to(torch.bfloat16)
    dh, dw, dg = dv(h), dv(w), dv(gain)
    ssq = dv(_ssq_pieces(h))
    act = torch.full((T, I), float("nan"), dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_gateup_swiglu_norm(ptr(dh), ptr(dw), ptr(dg), ptr(ssq), K // 16, 64, 1e-6, ptr(act), T, I, K, cuda_stream_ptr()))
    assert torch.equal(dh.cpu(), h), "the residual stream is an input only"
    hh = h.cuda()
    xn = torch.empty(T, K, dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_add_rmsnorm(ptr(hh), None, 0, 0, ptr(dg), ptr(xn), T, K, 1e-6, cuda_stream_ptr()))
    a2 = torch.empty(T, I, dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_gateup_swiglu(ptr(xn), ptr(dw), ptr(a2), T, I, K, cuda_stream_ptr()))
    assert_bf16_close(act, a2, 2, 0.02, "normalise-on-load vs norm kernel + projection, T=%d" % T, abs_floor=2e-3)
    gu = (rms_norm(h, gain, 1e-6).float() @ w.float().t()).to(torch.bfloat16)
    ref = torch.nn.functional.silu(gu[:, :I]) * gu[:, I:]
    assert_bf16_close(act, ref, 2, 0.05, "normalise-on-load vs torch, T=%d" % T, abs_floor=2e-3)


@pytest.mark.parametrize("T", [64, 40, 17, 1])
def test_qkv_rope_normalise_on_load(lib, T):
    """The QKV + QK-norm + RoPE + page-write launch with the input RMSNorm applied on load, against b2k_add_rmsnorm
    followed by b2k_qkv_rope."""
    a = QWEN3_0_6B
    nq, nkv, hd, bs, K = a.num_q_heads, a.num_kv_heads, a.head_dim, 16, a.hidden
    N = (nq + 2 * nkv) * hd
    nblocks = (T + bs - 1) // bs + 3
    g = torch.Generator().manual_seed(61)
    h = bf(T, K, seed=62)
    w = bf(N, K, scale=0.05, seed=63)
    gain = (1 + 0.25 * torch.randn(K, generator=g)).to(torch.bfloat16)
    qw = (1 + 0.25 * torch.randn(hd, generator=g)).to(torch.bfloat16)
    kw = (1 + 0.25 * torch.randn(hd, generator=g)).to(torch.bfloat16)
    positions = torch.randint(0, 4000, (T,), generator=g, dtype=torch.int32)
    slots = torch.randperm(nblocks * bs, generator=g)[:T].to(torch.int32)
    cos, sin = rope_cos_sin(a, torch.arange(4096)[None, :])
    table = torch.cat([cos[0, :, :64], sin[0, :, :64]], dim=-1).contiguous()
    dh, dw, dg, dqw, dkw, dtab, dpos, dsl = dv(h), dv(w), dv(gain), dv(qw), dv(kw), dv(table), dv(positions), dv(slots)
    ssq = dv(_ssq_pieces(h))
    outs = []
    for fused in (True, False):
        kp = torch.zeros(nblocks, nkv, bs, hd, dtype=torch.bfloat16, device="cuda")
        vp = torch.zeros_like(kp)
        q_out = torch.zeros(T, nq * hd, dtype=torch.bfloat16, device="cuda")
        if fused:
            chk(lib, lib.b2k_qkv_rope_norm(ptr(dh), ptr(dw), ptr(dg), ptr(ssq), K // 16, 64, 1e-6, ptr(dqw), ptr(dkw), ptr(dtab), ptr(dpos),
                                          ptr(dsl), ptr(q_out), ptr(kp), ptr(vp), T, nq, nkv, K, bs, 1e-6, cuda_stream_ptr()))
        else:
            hh = h.cuda()
            xn = torch.empty(T, K, dtype=torch.bfloat16, device="cuda")
            chk(lib, lib.b2k_add_rmsnorm(ptr(hh), None, 0, 0, ptr(dg), ptr(xn), T, K, 1e-6, cuda_stream_ptr()))
            chk(lib, lib.b2k_qkv_rope(ptr(xn), ptr(dw), ptr(dqw), ptr(dkw), ptr(dtab), ptr(dpos), ptr(dsl), ptr(q_out), ptr(kp), ptr(vp), T,
                                     nq, nkv, K, bs, 1e-6, cuda_stream_ptr()))
        outs.append((q_out.cpu(), kv_swizzle(kp.cpu()), kv_swizzle(vp.cpu())))
    (qf, kf, vf), (qs, ks, vs) = outs
    assert_bf16_close(qf, qs, 2, 0.02, "q: normalise-on-load vs norm kernel", abs_floor=0.02)
    assert_bf16_close(kf, ks, 2, 0.02, "k pages: normalise-on-load vs norm kernel", abs_floor=0.02)
    assert_bf16_close(vf, vs, 1, 0.02, "v pages: normalise-on-load vs norm kernel", abs_floor=2e-3)


@pytest.mark.parametrize("T,I,K", [(256, 3072, 1024), (1000, 3072, 1024), (4099, 3072, 1024), (300, 8192, 3072), (277, 512, 256)])
def test_gateup_swiglu_prefill_fused(lib, T, I, K):
    """gemm_gateup.cu: the prefill-sized gate/up projection with SwiGLU in its epilogue (two TMEM accumulators per tile).
    Same k order, one accumulator per output, same rounding points as b2k_linear followed by b2k_silu_mul: BIT-identical to
    that path, ragged last token tile included, and no write past row T."""
    x = bf(T, K, seed=31)
    w = bf(2 * I, K, scale=0.05, seed=32)
    act = torch.full((T + 3, I), float("nan"), dtype=torch.bfloat16, device="cuda")
    dx, dw = dv(x), dv(w)
    chk(lib, lib.b2k_gateup_swiglu(ptr(dx), ptr(dw), ptr(act), T, I, K, cuda_stream_ptr()))
    g2 = torch.empty(T, 2 * I, dtype=torch.bfloat16, device="cuda")
    a2 = torch.empty(T, I, dtype=torch.bfloat16, device="cuda")
    chk(lib, lib.b2k_linear(ptr(dx), ptr(dw), ptr(g2), 0, 1, T, 2 * I, K, 1, cuda_stream_ptr()))
    chk(lib, lib.b2k_silu_mul(ptr(g2), 0, 1, ptr(a2), T, I, cuda_stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(act[:T], a2), "fused prefill gate/up + SwiGLU differs from the two-kernel path"
    assert torch.isnan(act[T:].float()).all(), "wrote past the last token row"
    rows = slice(0, min(T, 512))
    gu = (x[rows].float() @ w.float().t()).to(torch.bfloat16)
    ref = torch.nn.functional.silu(gu[:, :I]) * gu[:, I:]
    # torch sums the K products in another order: g and u may round to the neighbouring bf16, and the product carries it
    assert rel_fro(act[rows], ref) < 5e-3 and rel_err(act[rows], ref) < 1e-2, (rel_fro(act[rows], ref), rel_err(act[rows], ref))


@pytest.mark.parametrize("T", [64, 40, 17, 1, 512, 777])  # > 64 tokens: the prefill-sized kernel (gemm_mc.cu, ROPE epilogue)
def test_qkv_rope_fused(lib, T):
    """gemm_pair.cu: QKV projection + QK-norm + RoPE + KV page write in one launch, against b2k_linear followed by
    b2k_qknorm_rope_kv (the path it replaces) and against the oracle's functions."""
    a = QWEN3_0_6B
    nq, nkv, hd, bs, K = a.num_q_heads, a.num_kv_heads, a.head_dim, 16, a.hidden
    N = (nq + 2 * nkv) * hd
    nblocks = (T + bs - 1) // bs + 3
    g = torch.Generator().manual_seed(23)
    x = bf(T, K, seed=24)
    w = bf(N, K, scale=0.05, seed=25)
    qw = (1 + 0.25 * torch.randn(hd, generator=g)).to(torch.bfloat16)
    kw = (1 + 0.25 * torch.randn(hd, generator=g)).to(torch.bfloat16)
    positions = torch.randint(0, 4000, (T,), generator=g, dtype=torch.int32)
    slots = torch.randperm(nblocks * bs, generator=g)[:T].to(torch.int32)
    if T > 3:
        slots[3] = -1
    cos, sin = rope_cos_sin(a, torch.arange(4096)[None, :])
    table = torch.cat([cos[0, :, :64], sin[0, :, :64]], dim=-1).contiguous()
    dx, dw, dqw, dkw, dtab, dpos, dsl = dv(x), dv(w), dv(qw), dv(kw), dv(table), dv(positions), dv(slots)
    outs = []
    for fused in (True, False):
        kp = torch.zeros(nblocks, nkv, bs, hd, dtype=torch.bfloat16, device="cuda")
        vp = torch.zeros_like(kp)
        q_out = torch.zeros(T, nq * hd, dtype=torch.bfloat16, device="cuda")
        if fused:
            chk(lib, lib.b2k_qkv_rope(ptr(dx), ptr(dw), ptr(dqw), ptr(dkw), ptr(dtab), ptr(dpos), ptr(dsl), ptr(q_out), ptr(kp), ptr(vp), T,
                                     nq, nkv, K, bs, 1e-6, cuda_stream_ptr()))
        else:
            qkv = torch.empty(T, N, dtype=torch.bfloat16, device="cuda")
            chk(lib, lib.b2k_linear(ptr(dx), ptr(dw), ptr(qkv), 0, 1, T, N, K, 1, cuda_stream_ptr()))
            chk(lib, lib.b2k_qknorm_rope_kv(ptr(qkv), 0, 1, ptr(dqw), ptr(dkw), ptr(dtab), ptr(dpos), ptr(dsl), ptr(q_out), ptr(kp), ptr(vp), T,
                                           nq, nkv, hd, bs, 1e-6, cuda_stream_ptr()))
        outs.append((q_out.cpu(), kv_swizzle(kp.cpu()), kv_swizzle(vp.cpu())))
    (qf, kf, vf), (qs, ks, vs) = outs
    # against the path it replaces: same rounding points, another order of the projection's fp32 sum -> an occasional ulp
    # before the norm, up to two after norm and RoPE.  RoPE is a difference of products: where it cancels, one input ulp
    # (2^-6 at the magnitudes here) is many ulps of the small result, hence the absolute floor of one input ulp.
    assert_bf16_close(qf, qs, 2, 0.05, "q: fused vs separate", abs_floor=0.02)
    assert_bf16_close(kf, ks, 2, 0.05, "k pages: fused vs separate", abs_floor=0.02)
    assert_bf16_close(vf, vs, 1, 0.02, "v pages: fused vs separate", abs_floor=2e-3)
    # against the oracle
    qkv_ref = (x.float() @ w.float().t()).to(torch.bfloat16)
    q = qkv_ref[:, : nq * hd].view(1, T, nq, hd)
    k = qkv_ref[:, nq * hd: (nq + nkv) * hd].view(1, T, nkv, hd)
    qn = rms_norm(q, qw, 1e-6).transpose(1, 2)
    kn = rms_norm(k, kw, 1e-6).transpose(1, 2)
    c, s_ = rope_cos_sin(a, positions[None, :].long())
    qr, kr = apply_rope(qn, kn, c, s_)
    # (a million elements at T >= 512: the rare q whose projection sum lands a full input ulp away at magnitude 4..8 -- 2^-5 --
    # and whose RoPE terms cancel needs a floor of two input ulps at the common magnitude)
    assert_bf16_close(qf, qr.transpose(1, 2).reshape(T, nq * hd), 2, 0.05, "q vs oracle", abs_floor=0.02 if T <= 64 else 0.04)
    for t in range(T):
        sl = int(slots[t])
        if sl >= 0:
            assert_bf16_close(kf[sl // bs, :, sl % bs, :], kr.transpose(1, 2)[0][t], 2, 0.08, "k page vs oracle t=%d" % t, abs_floor=0.02 if T <= 64 else 0.04)
    written = {int(v) for v in slots if v >= 0}
    for sl in range(nblocks * bs):
        if sl not in written:
            assert not kf[sl // bs, :, sl % bs, :].any() and not vf[sl // bs, :, sl % bs, :].any(), "stray K/V write at slot %d" % sl


def test_linear_is_deterministic(lib):
    T, N, K = 64, 1024, 2048
    x, w = bf(T, K, seed=12).cuda(), bf(N, K, scale=0.05, seed=13).cuda()
    outs = []
    for _ in range(3):
        o = torch.empty(16, T, N, dtype=torch.float32, device="cuda")
        chk(lib, lib.b2k_linear(ptr(x), ptr(w), ptr(o), 1, 16, T, N, K, 1, cuda_stream_ptr()))
        outs.append(o.cpu())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


@pytest.mark.parametrize("impl", [0, 1])
@pytest.mark.parametrize("T", [1, 5, 64])
def test_lmhead_argmax_full_vocab(lib, impl, T):
    a = QWEN3_0_6B
    V, K = a.vocab, a.hidden
    x = bf(T, K, seed=14)
    w = bf(V, K, scale=0.02, seed=15)
    # forced ties: rows 777 and 90000 equal the arg-max row -> the lowest index must win
    ref_logits = torch.nn.functional.linear(x, w)
    top = int(ref_logits[0].argmax())
    w[90000] = w[top]
    if top > 777:
        w[777] = w[top]
    x_d, w_d = x.cuda(), w.cuda()
    logits = torch.empty(T, V, dtype=torch.bfloat16, device="cuda")
    toks = torch.full((T,), -1, dtype=torch.int32, device="cuda")
    scratch = torch.zeros(T, dtype=torch.int64, device="cuda")
    chk(lib, lib.b2k_lmhead_argmax(ptr(x_d), ptr(w_d), ptr(logits), ptr(toks), ptr(scratch), T, V, K, impl, cuda_stream_ptr()))
    ref_logits = torch.nn.functional.linear(x, w)
    assert_bf16_close(logits, ref_logits, 1, 0.03, "lm_head logits", abs_floor=1e-3)
    # the fused arg-max must equal torch.argmax over the kernel's own bf16 logits (first index on ties)
    assert torch.equal(toks.cpu().long(), logits.cpu().float().argmax(-1))
    assert int(toks[0]) == min(top, 777)
    if impl == 1:  # without the logits buffer the fused path must give the same tokens
        toks2 = torch.full((T,), -1, dtype=torch.int32, device="cuda")
        chk(lib, lib.b2k_lmhead_argmax(ptr(x_d), ptr(w_d), None, ptr(toks2), ptr(scratch), T, V, K, 1, cuda_stream_ptr()))
        assert torch.equal(toks, toks2)


# ---------------------------------------------------------------------------------------------
def attention_exact(q, k, v, scaling, causal):
    """fp64 softmax attention over the bf16 inputs: the value both the oracle and the kernel approximate."""
    B, Hq, T, d = q.shape
    g = Hq // k.shape[1]
    kf, vf = k.double().repeat_interleave(g, 1), v.double().repeat_interleave(g, 1)
    s = q.double() @ kf.transpose(-1, -2) * scaling
    if causal:
        S = k.shape[2]
        s = s.masked_fill(~torch.ones(T, S, dtype=torch.bool).tril(diagonal=S - T), float("-inf"))
    return torch.softmax(s, -1) @ vf


def check_attention(got, q4, k4, v4, causal, what):
    """Bar for attention outputs.  P is rounded to bf16 before P.V (as in torch's CPU flash kernel), which
    by itself perturbs the result by about one output ulp, so two correct implementations that round P
    against different running maxima differ in many elements by one ulp.  Hence: (1) every element within
    2 bf16 ulp of the oracle's written-out rounding points, (2) within 1e-2 relative of HF's sdpa, and
    (3) no further from the exact fp64 result than 1.25x the oracle's own CPU kernel is."""
    hd = q4.shape[-1]
    ref = attention_explicit(q4, k4, v4, hd ** -0.5, causal)
    assert_bf16_close(got, ref.reshape(got.shape) if ref.shape != got.shape else ref, 2, 0.6, what, abs_floor=2e-3)
    sdpa = torch.nn.functional.scaled_dot_product_attention(q4, k4, v4, is_causal=causal and q4.shape[2] > 1, enable_gqa=True) \
        if (q4.shape[2] == 1 or q4.shape[2] == k4.shape[2]) else None
    exact = attention_exact(q4, k4, v4, hd ** -0.5, causal)
    mine = rel_fro(got.reshape(exact.shape).double(), exact)
    floor = rel_fro(ref.double(), exact)
    if sdpa is not None:
        assert rel_err(got.reshape(sdpa.shape), sdpa) < 1e-2, what
        floor = max(floor, rel_fro(sdpa.double(), exact))
    assert mine <= 1.25 * floor + 1e-4, "%s: error vs exact %.3g, oracle's own %.3g" % (what, mine, floor)


def build_pages(seq_lens, nkv, hd, bs, seed):
    """Random K/V per sequence scattered into pages through a shuffled block table."""
    g = torch.Generator().manual_seed(seed)
    B = len(seq_lens)
    nblk = [-(-s // bs) for s in seq_lens]
    total = sum(nblk) + 1
    perm = (torch.randperm(total - 1, generator=g) + 1).tolist()  # page 0 stays the null page
    stride = max(nblk) + 2
    bt = torch.zeros(B, stride, dtype=torch.int32)
    kp = torch.zeros(total, nkv, bs, hd, dtype=torch.bfloat16)
    vp = torch.zeros_like(kp)
    ks, vs = [], []
    it = iter(perm)
    for b, s in enumerate(seq_lens):
        k = torch.randn(nkv, s, hd, generator=g).to(torch.bfloat16)
        v = torch.randn(nkv, s, hd, generator=g).to(torch.bfloat16)
        ks.append(k)
        vs.append(v)
        for j in range(nblk[b]):
            pg = next(it)
            bt[b, j] = pg
            n = min(bs, s - j * bs)
            kp[pg, :, :n] = k[:, j * bs: j * bs + n]
            vp[pg, :, :n] = v[:, j * bs: j * bs + n]
            if n < bs:  # stale finite data beyond the sequence end must be ignored
                kp[pg, :, n:] = 3.0
                vp[pg, :, n:] = -7.0
    return bt, kv_swizzle(kp), kv_swizzle(vp), ks, vs


@pytest.mark.parametrize("seq_lens", [[1, 16, 17, 37, 300, 1023, 640], [5], [2000, 1], [700] * 64,
                                      [1 + (37 * i) % 900 for i in range(50)]])
def test_attn_decode(lib, seq_lens):
    a = QWEN3_0_6B
    nq, nkv, hd, bs = a.num_q_heads, a.num_kv_heads, a.head_dim, 16
    B = len(seq_lens)
    bt, kp, vp, ks, vs = build_pages(seq_lens, nkv, hd, bs, seed=20)
    q = bf(B, nq * hd, seed=21)
    out = torch.full((B, nq * hd), float("nan"), dtype=torch.bfloat16, device="cuda")
    nfl = lib.b2k_attn_decode_scratch_floats(B, nq, nkv, hd)
    scratch = torch.full((max(nfl, 1),), float("nan"), dtype=torch.float32, device="cuda")
    chk(lib, lib.b2k_attn_decode(ptr(dv(q)), ptr(dv(kp)), ptr(dv(vp)), ptr(dv(bt)), bt.shape[1],
                                ptr(dv(torch.tensor(seq_lens, dtype=torch.int32))), ptr(scratch), ptr(out), B, nq, nkv, hd, bs,
                                cuda_stream_ptr()))
    o = out.cpu()
    for b in (range(B) if B <= 8 else range(0, B, 7)):
        check_attention(o[b].view(1, nq, 1, hd), q[b].view(1, nq, 1, hd), ks[b][None], vs[b][None], False,
                        "attn_decode b=%d S=%d" % (b, seq_lens[b]))
    # determinism: the page-balanced partials are merged in run order, never atomically
    out2 = torch.empty_like(out)
    chk(lib, lib.b2k_attn_decode(ptr(dv(q)), ptr(dv(kp)), ptr(dv(vp)), ptr(dv(bt)), bt.shape[1],
                                ptr(dv(torch.tensor(seq_lens, dtype=torch.int32))), ptr(scratch), ptr(out2), B, nq, nkv, hd, bs,
                                cuda_stream_ptr()))
    assert torch.equal(out, out2)


def test_attn_prefill_copies_are_bit_identical(lib):
    """Identical sequences in different rows of the batch (other pages, other CTAs) and repeated launches must give the
    same bits: nothing in the prefill kernel depends on where a sequence sits or on timing."""
    a = QWEN3_0_6B
    nq, nkv, hd, bs = a.num_q_heads, a.num_kv_heads, a.head_dim, 16
    n, copies = 2100, 3
    bt1, kp1, vp1, ks, vs = build_pages([n], nkv, hd, bs, seed=40)
    pages = bt1.shape[1]
    # the same K/V written to three different sets of pages
    kp = torch.zeros(1 + copies * pages, nkv, bs, hd, dtype=torch.bfloat16)
    vp = torch.zeros_like(kp)
    bt = torch.zeros(copies, pages, dtype=torch.int32)
    g = torch.Generator().manual_seed(41)
    perm = (torch.randperm(copies * pages, generator=g) + 1).to(torch.int32)
    for c in range(copies):
        for j in range(pages):
            dst = int(perm[c * pages + j])
            src = int(bt1[0, j])
            kp[dst], vp[dst] = kp1[src], vp1[src]
            bt[c, j] = dst
    q1 = bf(n, nq * hd, seed=42)
    q = q1.repeat(copies, 1)
    T = copies * n
    I32 = lambda v: (C.c_int32 * len(v))(*v)
    outs = []
    for rep in range(2):
        out = torch.full((T, nq * hd), float("nan"), dtype=torch.bfloat16, device="cuda")
        chk(lib, lib.b2k_attn_prefill(ptr(dv(q)), ptr(dv(kp)), ptr(dv(vp)), ptr(dv(bt)), pages, I32([c * n for c in range(copies)]),
                                     I32([0] * copies), I32([n] * copies), I32(list(range(copies))), copies, ptr(out), nq, nkv, hd, bs,
                                     cuda_stream_ptr()))
        outs.append(out.cpu())
    assert torch.equal(outs[0], outs[1]), "two launches on the same inputs differ"
    o = outs[0].view(copies, n, nq * hd)
    for c in range(1, copies):
        assert torch.equal(o[0], o[c]), "copy %d of the sequence differs from copy 0" % c


def test_attn_decode_empty_and_padding_rows(lib):
    """A graph bucket pads the batch with rows of length 0: their output must be zeros, not NaN."""
    a = QWEN3_0_6B
    nq, nkv, hd, bs = a.num_q_heads, a.num_kv_heads, a.head_dim, 16
    seq_lens = [0, 5, 0]
    bt, kp, vp, ks, vs = build_pages([1, 5, 1], nkv, hd, bs, seed=22)
    q = bf(3, nq * hd, seed=23)
    out = torch.full((3, nq * hd), float("nan"), dtype=torch.bfloat16, device="cuda")
    scratch = torch.zeros(lib.b2k_attn_decode_scratch_floats(3, nq, nkv, hd), device="cuda")
    chk(lib, lib.b2k_attn_decode(ptr(dv(q)), ptr(dv(kp)), ptr(dv(vp)), ptr(dv(bt)), bt.shape[1],
                                ptr(dv(torch.tensor(seq_lens, dtype=torch.int32))), ptr(scratch), ptr(out), 3, nq, nkv, hd, bs,
                                cuda_stream_ptr()))
    o = out.cpu()
    assert not o[0].any() and not o[2].any()
    check_attention(o[1].view(1, nq, 1, hd), q[1].view(1, nq, 1, hd), ks[1][None], vs[1][None], False, "attn_decode mid row")


@pytest.mark.parametrize("case", ["fresh", "chunked", "long"])
def test_attn_prefill(lib, case):
    a = QWEN3_0_6B
    nq, nkv, hd, bs = a.num_q_heads, a.num_kv_heads, a.head_dim, 16
    if case == "fresh":
        pos0 = [0, 0, 0, 0]
        ntok = [5, 64, 130, 1]
    elif case == "long":  # many key tiles: the K/V double buffer and every barrier phase of the tcgen05 kernel turn over
        pos0 = [300, 0, 1000]
        ntok = [200, 700, 90]
    else:  # queries appended after an existing prefix (chunked prefill)
        pos0 = [70, 3, 129]
        ntok = [61, 200, 64]
    total_len = [p + n for p, n in zip(pos0, ntok)]
    bt, kp, vp, ks, vs = build_pages(total_len, nkv, hd, bs, seed=30)
    T = sum(ntok)
    q = bf(T, nq * hd, seed=31)
    first = [sum(ntok[:i]) for i in range(len(ntok))]
    out = torch.full((T, nq * hd), float("nan"), dtype=torch.bfloat16, device="cuda")
    I32 = lambda v: (C.c_int32 * len(v))(*v)
    chk(lib, lib.b2k_attn_prefill(ptr(dv(q)), ptr(dv(kp)), ptr(dv(vp)), ptr(dv(bt)), bt.shape[1], I32(first), I32(pos0),
                                 I32(ntok), I32(list(range(len(ntok)))), len(ntok), ptr(out), nq, nkv, hd, bs, cuda_stream_ptr()))
    o = out.cpu()
    for i, n in enumerate(ntok):
        qs = q[first[i]: first[i] + n].view(n, nq, hd).transpose(0, 1)[None]  # [1][nq][n][hd]
        got = o[first[i]: first[i] + n].view(n, nq, hd).transpose(0, 1)[None]  # causal with prefix: tril(diagonal=S-T)
        check_attention(got, qs, ks[i][None], vs[i][None], True, "attn_prefill seq %d (%s)" % (i, case))
    # the tcgen05 kernel (default) against the mma.sync kernel it replaced: same contract, other tile sizes and sum orders
    import os
    out2 = torch.full((T, nq * hd), float("nan"), dtype=torch.bfloat16, device="cuda")
    os.environ["B2_PREFILL_TC"] = "0"
    try:
        chk(lib, lib.b2k_attn_prefill(ptr(dv(q)), ptr(dv(kp)), ptr(dv(vp)), ptr(dv(bt)), bt.shape[1], I32(first), I32(pos0),
                                     I32(ntok), I32(list(range(len(ntok)))), len(ntok), ptr(out2), nq, nkv, hd, bs, cuda_stream_ptr()))
    finally:
        del os.environ["B2_PREFILL_TC"]
    # both round P to bf16 before P.V, but scale it differently (the tcgen05 kernel by a reference maximum that trails the
    # running maximum by up to 2^8, the mma.sync one by the running maximum itself), so P's roundings differ: every element
    # within 2 ulp, up to 40 % of them not bit-equal; each is checked against the oracle above
    assert_bf16_close(out, out2, 2, 0.4, "tcgen05 vs mma.sync prefill attention (%s)" % case, abs_floor=2e-3)
