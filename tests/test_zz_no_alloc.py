"""The allocation-free decode-attention entry point (b2k_attn_decode_noalloc) against the regular one.

Who needs it: code that cannot call cudaMalloc where the attention runs -- a server whose KV pool
already took the free HBM, a request path that cannot afford cudaMalloc's device-wide stall, a plugin
that is only handed input/output pointers.  Its split partials live in an array compiled into the
library, shared by all calls and never cleared, so this test calls it twice in a row (second call
with different inputs) to show stale contents of the first call do not leak into the second.

This file is named to be collected last: with `-x` every other test has run before it.
"""
import pytest
import torch

from _util import cuda_stream_ptr, pkg, ptr
from test_kernels_gpu import build_pages, check_attention

pytestmark = pytest.mark.gpu


def _case(lib, seed, seq_lens):
    nq, nkv, hd, bs = 16, 8, 128, 16
    B = len(seq_lens)
    bt, kp, vp, ks, vs = build_pages(seq_lens, nkv, hd, bs, seed=seed)
    q = (torch.randn(B, nq * hd, generator=torch.Generator().manual_seed(seed + 1))).to(torch.bfloat16)
    dev = dict(q=q.cuda(), kp=kp.cuda(), vp=vp.cuda(), bt=bt.cuda(), sl=torch.tensor(seq_lens, dtype=torch.int32).cuda())
    # existing entry point, caller-provided scratch
    ref = torch.full((B, nq * hd), float("nan"), dtype=torch.bfloat16, device="cuda")
    nfl = lib.b2k_attn_decode_scratch_floats(B, nq, nkv, hd)
    scratch = torch.empty(max(1, nfl), dtype=torch.float32, device="cuda")
    rc = lib.b2k_attn_decode(ptr(dev["q"]), ptr(dev["kp"]), ptr(dev["vp"]), ptr(dev["bt"]), bt.shape[1], ptr(dev["sl"]),
                             ptr(scratch), ptr(ref), B, nq, nkv, hd, bs, cuda_stream_ptr())
    assert rc == 0, lib.b2_last_error()
    # new entry point: same inputs and output, no scratch argument
    out = torch.full((B, nq * hd), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = lib.b2k_attn_decode_noalloc(ptr(dev["q"]), ptr(dev["kp"]), ptr(dev["vp"]), ptr(dev["bt"]), bt.shape[1], ptr(dev["sl"]),
                                     ptr(out), B, nq, nkv, hd, bs, cuda_stream_ptr())
    assert rc == 0, lib.b2_last_error()
    torch.cuda.synchronize()
    # same kernels, same merge order -> the two entry points must agree bit for bit ...
    assert torch.equal(out.cpu(), ref.cpu()), "no-alloc entry point differs from b2k_attn_decode"
    # ... and both must meet the bar the other attention tests use against the oracle
    o = out.cpu()
    for b in range(B):
        check_attention(o[b].view(1, nq, 1, hd), q[b].view(1, nq, 1, hd), ks[b][None], vs[b][None], False,
                        "noalloc attn b=%d S=%d" % (b, seq_lens[b]))


def test_noalloc_matches_regular_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    lib = pkg().load_library()
    _case(lib, seed=40, seq_lens=[300, 17, 640, 1])    # first call on this device: sets up the address
    _case(lib, seed=50, seq_lens=[33, 1023, 5])        # second call reuses the array as the first left it
    _case(lib, seed=60, seq_lens=[64, 200])


def test_noalloc_rejects_oversize():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    lib = pkg().load_library()
    B, nq, nkv, hd = 512, 64, 32, 128   # 512 * 32 units * 260 floats >> 2 Mi
    z = torch.zeros(16, dtype=torch.int32, device="cuda")
    rc = lib.b2k_attn_decode_noalloc(ptr(z), ptr(z), ptr(z), ptr(z), 1, ptr(z), ptr(z), B, nq, nkv, hd, 16, None)
    assert rc != 0
    msg = lib.b2_last_error().decode()
    assert "scratch" in msg and "holds" in msg
