"""Build libbt_b200.so in-tree with nvcc for sm_100a (no torch headers: plain C ABI).

    python bayesian-torch_b200/csrc/build.py [--force] [--verbose]
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
SOURCES = ["bt_api.cu", "bt_kl.cu", "bt_misc.cu", "bt_lstm.cu", "bt_simt_fwd.cu", "bt_simt_dgrad.cu", "bt_simt_wgrad.cu", "bt_tc.cu"]
HEADERS = ["bt_common.cuh", "bt_kl.cuh", "bt_simt.cuh", "bt_tc.cuh", os.path.join("..", "..", "include", "bt_b200.h")]
LIB = os.path.join(HERE, "libbt_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr", "-Xptxas", "-v"] + os.environ.get("BT_NVCC_EXTRA", "").split()


def _mtime(p):
    return os.path.getmtime(p) if os.path.exists(p) else 0.0


def build(force: bool = False, verbose: bool = False) -> str:
    newest_hdr = max(_mtime(os.path.join(HERE, h)) for h in HEADERS)
    newest_src = max(_mtime(os.path.join(HERE, s)) for s in SOURCES)
    if not force and _mtime(LIB) >= max(newest_hdr, newest_src):
        return LIB  # prebuilt library is current (the case on the GPU box)
    objs, jobs = [], []
    for src in SOURCES:
        s = os.path.join(HERE, src)
        o = os.path.join(HERE, "build", src.replace(".cu", ".o"))
        objs.append(o)
        if force or _mtime(o) < max(_mtime(s), newest_hdr):
            jobs.append((s, o))
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)

    def compile_one(job):
        s, o = job
        r = subprocess.run([NVCC, *FLAGS, "-c", s, "-o", o], capture_output=True, text=True)
        return s, r

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for s, r in ex.map(compile_one, jobs):
            if verbose or r.returncode != 0:
                sys.stderr.write(r.stdout + r.stderr)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed on {s}")
    if jobs or not os.path.exists(LIB):
        r = subprocess.run([NVCC, "-shared", "-o", LIB, *objs, "-cudart", "static"], capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
