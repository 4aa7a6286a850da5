"""In-tree nvcc build of the engine library and the model-server binary for sm_100a.

    python -m importlib ...  (see __graft_entry__.build) or:  python aws-k8s-ansible-provisioner_b200/build.py

Outputs (git-ignored, shipped to the GPU box by gpurun):
    aws-k8s-ansible-provisioner_b200/libb2engine.so     C ABI of include/b2_engine.h
    aws-k8s-ansible-provisioner_b200/b2serve            HTTP model server (when csrc/server/ exists)
Objects are cached under build/ and rebuilt when a source or header is newer.
"""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
BUILD = os.path.join(ROOT, "build")
LIB = os.path.join(PKG, "libb2engine.so")
SERVER = os.path.join(PKG, "b2serve")
ROUTER = os.path.join(PKG, "b2route")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
    "-Xcompiler", "-Wall", "-Xcudafe", "--diag_suppress=177",
]
ENGINE_SOURCES = ["elementwise.cu", "elementwise_opt.cu", "gemm_simt.cu", "gemm_tc.cu", "gemm_pair.cu", "gemm_gateup.cu", "gemm_resid.cu", "gemm_mc.cu", "attn_decode.cu", "attn_prefill.cu", "attn_prefill_tc.cu", "sampling.cu", "chain.cu", "chain_cluster.cu",
                  "engine.cu", "capi.cu"]


def nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the engine has no CPU fallback and cannot be built without the CUDA toolkit")


def _newer(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _headers() -> list[str]:
    hs = [os.path.join(ROOT, "include", "b2_engine.h")]
    for dp, _, fns in os.walk(CSRC):
        hs += [os.path.join(dp, f) for f in fns if f.endswith((".h", ".hpp", ".cuh"))]
    return hs


def _compile(src: str, obj: str, extra: list[str], verbose: bool) -> None:
    cmd = [nvcc(), *NVCC_FLAGS, *extra, "-c", src, "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    if verbose and (r.stdout.strip() or r.stderr.strip()):
        print(r.stdout, r.stderr, flush=True)


def build(verbose: bool = False, force: bool = False, ptxas_info: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    hdrs = _headers()
    extra = ["-Xptxas", "-v"] if ptxas_info else []
    jobs = []
    objs = []
    for s in ENGINE_SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(BUILD, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or ptxas_info or _newer(obj, [src] + hdrs):
            jobs.append((src, obj))
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        futs = [ex.submit(_compile, s, o, extra, verbose or ptxas_info) for s, o in jobs]
        for f in futs:
            f.result()
    if jobs or not os.path.exists(LIB):
        cmd = [nvcc(), "-shared", "-o", LIB, *objs, "-Xcompiler", "-fPIC", "-lcudart_static", "-ldl", "-lpthread", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    server_dir = os.path.join(CSRC, "server")
    if os.path.isdir(server_dir):
        srcs = sorted(os.path.join(server_dir, f) for f in os.listdir(server_dir) if f.endswith(".cpp"))
        if srcs and (force or _newer(SERVER, srcs + hdrs + [LIB])):
            cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-pthread", "-I", os.path.join(ROOT, "include"), "-I", CSRC, *srcs,
                   "-o", SERVER, "-L", PKG, "-lb2engine", "-Wl,-rpath,$ORIGIN"]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("server build failed:\n%s\n%s" % (r.stdout, r.stderr))
    router_dir = os.path.join(CSRC, "router")
    if os.path.isdir(router_dir):  # the replica front needs neither CUDA nor the engine library
        srcs = sorted(os.path.join(router_dir, f) for f in os.listdir(router_dir) if f.endswith(".cpp"))
        if srcs and (force or _newer(ROUTER, srcs + hdrs)):
            cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-pthread", "-I", CSRC, *srcs, "-o", ROUTER]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("router build failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


if __name__ == "__main__":
    build(verbose="-v" in sys.argv, force="-f" in sys.argv, ptxas_info="--ptxas" in sys.argv)
    print(LIB)
