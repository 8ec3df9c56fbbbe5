#!/usr/bin/env python3
"""Builds the drop-in CPython extension `fast_ctc_decode` (C++/pybind11 over the C ABI) in-tree,
next to libfctc_b200.so (linked with rpath $ORIGIN)."""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "pymodule.cpp")
OUT = os.path.join(HERE, "fast_ctc_decode" + sysconfig.get_config_var("EXT_SUFFIX"))


def build(force=False):
    import pybind11
    deps = [SRC, os.path.join(HERE, "csrc", "pymodule_crf.inc"), os.path.join(HERE, "..", "include", "fctc.h"),
            os.path.join(HERE, "libfctc_b200.so")]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(d) <= os.path.getmtime(OUT) for d in deps if os.path.exists(d)):
        return OUT
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-fvisibility=hidden", "-I", pybind11.get_include(),
           "-I", sysconfig.get_paths()["include"], SRC, "-o", OUT, "-L", HERE, "-lfctc_b200", "-Wl,-rpath,$ORIGIN"]
    print(" ".join(cmd), file=sys.stderr)
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
