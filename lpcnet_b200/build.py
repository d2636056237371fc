"""Builds lpcnet_b200/liblpcnet_b200.so (CUDA kernels + C-ABI shim) in-tree for sm_100a.

nvcc cross-compiles without a GPU.  `-fmad=false` is part of the numerics contract: the
kernels mirror the reference's AVX2 build operation by operation and fuse multiply-adds only
where the reference does (explicit fmaf / fma.rn.f32x2).  Translation units compile in parallel."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "liblpcnet_b200.so")
SOURCES = ["model.cu", "frame.cu", "sample.cu", "sample_t8.cu", "sample_t16.cu", "sample_t32.cu", "sample_pp.cu",
           "api.cu", "probe.cu", "analysis.cu", "plc.cu", "noalloc.cu", "demo_shims.cu", "multi.cu"]
HEADERS = ["engine.h", "bitexact.cuh", "sample_kernel.cuh", "dsp.cuh", "vq.cuh", "plc_causal.cuh", "dred.cuh", os.path.join("..", "..", "include", "lpcnet_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-fmad=false", "-Xcompiler", "-fPIC,-fvisibility=hidden"]


def _mtime(p):
    return os.path.getmtime(p) if os.path.exists(p) else 0.0


def build(force: bool = False, verbose: bool = False, extra_flags=(), tag: str = "") -> str:
    """tag/extra_flags: development variants (e.g. tag="_clk", extra_flags=["-DLPCB_PHASE_CLOCKS"]) built next to
    the product library as liblpcnet_b200<tag>.so; the product build passes neither."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    OBJ = os.path.join(HERE, "build" + tag)
    LIB = os.path.join(HERE, f"liblpcnet_b200{tag}.so")
    os.makedirs(OBJ, exist_ok=True)
    hdr_t = max([_mtime(os.path.join(CSRC, h)) for h in HEADERS] + [_mtime(os.path.abspath(__file__))])
    jobs = []
    for src in SOURCES:
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        if force or _mtime(obj) < max(_mtime(os.path.join(CSRC, src)), hdr_t):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        cmd = [nvcc] + NVCC_FLAGS + list(extra_flags) + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, os.path.join(CSRC, src)]
        return src, subprocess.run(cmd, capture_output=True, text=True)

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for src, r in ex.map(compile_one, jobs):
            if r.returncode != 0:
                sys.stderr.write(r.stdout + r.stderr)
                raise RuntimeError(f"nvcc failed on {src}")
            if verbose:
                sys.stderr.write(r.stdout + r.stderr)
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in SOURCES]
    if jobs or not os.path.exists(LIB):
        r = subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs,
                           capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed for liblpcnet_b200.so")
    if not tag:
        # The reference installs `liblpcnet` with a pkg-config file `lpcnet.pc` (Makefile.am:9, lpcnet.pc.in): the same
        # names beside the library, so that a program's `-llpcnet` / `pkg-config --libs lpcnet` resolve to this
        # library without a change to its build.
        alias = os.path.join(HERE, "liblpcnet.so")
        if not os.path.islink(alias) or os.readlink(alias) != os.path.basename(LIB):
            if os.path.lexists(alias):
                os.remove(alias)
            os.symlink(os.path.basename(LIB), alias)
        pc = os.path.join(HERE, "lpcnet.pc")
        text = ("# lpcnet_b200 standing in for the reference's liblpcnet (lpcnet.pc.in)\n"
                f"libdir={HERE}\nincludedir={os.path.join(os.path.dirname(HERE), 'include')}\n\n"
                "Name: lpcnet\nDescription: LPCNet synthesis path on B200 GPUs (drop-in for liblpcnet)\nVersion: 0.2\n"
                "Libs: -L${libdir} -llpcnet -Wl,-rpath,${libdir}\nCflags: -I${includedir}\n")
        if not os.path.exists(pc) or open(pc).read() != text:
            open(pc, "w").write(text)
    return LIB


if __name__ == "__main__":
    if "--phase-clocks" in sys.argv:
        print(build(force="--force" in sys.argv, extra_flags=["-DLPCB_PHASE_CLOCKS"], tag="_clk"))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
