#!/usr/bin/env python3
"""Builds libfctc_b200.so (CUDA kernels + C ABI) in-tree for sm_100a with nvcc.

nvcc cross-compiles without a GPU, so this runs in the CPU container; the built library is
git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.environ.get("FCTC_BUILD_OUT") or os.path.join(HERE, "libfctc_b200.so")  # (a second library for A/B runs)
SOURCES = ["api.cu", "viterbi.cu", "beam.cu", "beam_small.cu", "duplex.cu", "duplex_packed.cu", "crf.cu", "fastq.cu", "align.cu", "pack.cu", "multi.cu", "synth.cu"]
HEADERS = ["kernels.h", "device_utils.cuh", "phred.cuh", "duplex_common.cuh", os.path.join("..", "..", "include", "fctc.h")]

EXTRA = os.environ.get("FCTC_NVCC_EXTRA", "").split()  # e.g. -DDP_CTAS_EXACT=20 (development)
NVCC_FLAGS = EXTRA + [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    # bit-exact parity with the reference's f32 arithmetic: no FMA contraction, IEEE div/sqrt,
    # denormals kept (all are nvcc defaults except -fmad)
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math",
]
LINK_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-cudart", "static"]
OBJ_DIR = os.path.join(HERE, "build" + ("_" + os.path.basename(OUT).replace(".so", "") if os.environ.get("FCTC_BUILD_OUT") else ""))


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    """Every .cu is compiled to its own object (in parallel, only when it or a header changed), then linked."""
    if not force and not needs_build():
        return OUT
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdr_time = max(os.path.getmtime(os.path.join(CSRC, h)) for h in HEADERS if os.path.exists(os.path.join(CSRC, h)))
    hdr_time = max(hdr_time, os.path.getmtime(os.path.abspath(__file__)))
    jobs, objs = [], []
    for name in SOURCES:
        src = os.path.join(CSRC, name)
        if not os.path.exists(src):
            continue
        obj = os.path.join(OBJ_DIR, name.replace(".cu", ".o"))
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_time):
            cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
            print(" ".join(cmd), file=sys.stderr)
            jobs.append((name, subprocess.Popen(cmd)))
    failed = [name for name, proc in jobs if proc.wait() != 0]
    if failed:
        raise SystemExit("nvcc failed for: " + ", ".join(failed))
    cmd = ["nvcc"] + LINK_FLAGS + ["-o", OUT] + objs
    print(" ".join(cmd), file=sys.stderr)
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(OUT)
