"""CPU-side checks of the drop-in boundary: the C-ABI library builds for sm_100a, loads, and
exports every symbol include/b2_engine.h declares; without a GPU the compute entry points fail
loudly instead of falling back."""
import ctypes as C
import os
import re

import pytest

from _util import ROOT


def header_symbols():
    src = open(os.path.join(ROOT, "include", "b2_engine.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2k?_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_expected_surface():
    syms = header_symbols()
    for must in ("b2_engine_create", "b2_prefill", "b2_decode_step", "b2_get_logits", "b2_engine_free", "b2_last_error",
                 "b2k_attn_decode", "b2k_linear", "b2k_lmhead_argmax"):
        assert must in syms


def test_library_exports_every_declared_symbol(pkg, lib):
    from importlib import import_module
    capi = import_module("aws-k8s-ansible-provisioner_b200.capi")
    syms = header_symbols()
    raw = C.CDLL(pkg.LIB_PATH)
    missing = [s for s in syms if not hasattr(raw, s)]
    assert not missing, "declared in include/b2_engine.h but not exported: %s" % missing
    # and the ctypes binding types every one of them
    untyped = [s for s in syms if s not in capi.SIGNATURES]
    assert not untyped, "no ctypes signature for: %s" % untyped
    assert lib.b2_abi_version() == capi.ABI_VERSION


def test_config_struct_layout(pkg):
    from importlib import import_module
    capi = import_module("aws-k8s-ansible-provisioner_b200.capi")
    # 18 four-byte fields + one pointer + 4 four-byte fields (ABI v2: arch, rotary_dim, max_positions,
    # sliding_window)
    assert C.sizeof(capi.b2_config) == 18 * 4 + 8 + 4 * 4
    assert capi.b2_config.arch.offset == 18 * 4 + 8 and capi.b2_config.rotary_dim.offset == 18 * 4 + 12
    assert C.sizeof(capi.b2_sampling) == 24


def test_library_is_sm100a_native(pkg):
    """The shipped SASS must contain the Blackwell paths (tcgen05 MMA, TMEM loads, TMA)."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    r = subprocess.run([cuobjdump, "-sass", pkg.LIB_PATH], capture_output=True, text=True)
    assert "sm_100a" in r.stdout
    for mnem in ("UTCHMMA", "LDTM", "UTMALDG", "UBLKCP"):
        assert mnem in r.stdout, "missing %s in SASS" % mnem


def test_no_cpu_fallback(pkg, lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.B2Error) as ei:
        pkg.Engine(pkg.EngineConfig(num_layers=1, hidden=256, num_q_heads=2, num_kv_heads=1, intermediate=256, vocab=256,
                                    max_model_len=64, num_blocks=8, max_seqs=2, max_batch_tokens=64))
    assert "no CPU path" in str(ei.value) or "CUDA" in str(ei.value)


def test_missing_library_fails_loudly(pkg, tmp_path):
    from importlib import import_module
    capi = import_module("aws-k8s-ansible-provisioner_b200.capi")
    with pytest.raises(pkg.B2Error):
        capi.load_library(str(tmp_path / "nope.so"))
