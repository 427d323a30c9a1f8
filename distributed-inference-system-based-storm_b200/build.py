"""Builds libsib.so in-tree with nvcc for sm_100a (run by __graft_entry__.build()).

    python build.py [--force]

Every translation unit goes through nvcc (-gencode arch=compute_100a,code=sm_100a -lineinfo); objects land in
csrc/build/, the library next to this file so that it travels to the GPU box with the source snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libsib.so")
SOURCES = ["sib_model.cpp", "sib_pbimport.cpp", "sib_json.cpp", "sib_engine.cpp", "sib_shared.cpp", "sib_abi.cpp", "sib_bolt.cpp", "sib_jni.cpp", "sib_loadgen.cpp",
           "sib_kernels_fp32.cu", "sib_kernels_tc.cu", "sib_kernels_tcgen.cu", "sib_kernels_guard.cu", "sib_static_scratch.cu", "sib_kernels_json.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler",
         "-fPIC,-Wall,-Wno-unused-function", "-ccbin", "/usr/bin/g++"] + os.environ.get("SIB_NVCC_EXTRA", "").split()


def _newest_header() -> float:
    t = 0.0
    for d in (CSRC, os.path.join(HERE, "..", "include")):
        for f in os.listdir(d):
            if f.endswith((".h", ".hpp", ".cuh")):
                t = max(t, os.path.getmtime(os.path.join(d, f)))
    return t


def build(force: bool = False, verbose: bool = False) -> str:
    bdir = os.path.join(CSRC, "build")
    os.makedirs(bdir, exist_ok=True)
    hdr_t = _newest_header()
    srcs = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    objs = []
    for s in srcs:
        src = os.path.join(CSRC, s)
        obj = os.path.join(bdir, s + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_t):
            cmd = [NVCC, *FLAGS, "-x", "cu", "-c", src, "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        return cmd, r

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for cmd, r in ex.map(run, jobs):
            if verbose or r.returncode:
                sys.stderr.write(" ".join(cmd[-3:]) + "\n" + r.stdout + r.stderr)
            if r.returncode:
                raise RuntimeError("nvcc failed: " + " ".join(cmd))
    if jobs or not os.path.exists(OUT):
        cmd = [NVCC, "-shared", "-o", OUT, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-ccbin",
               "/usr/bin/g++", "-Xlinker", "--no-undefined", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
