"""Build the HOST EMULATION of the tree kernels (test harness only; see csrc/rm_emu.h).

g++ compiles the same .cu sources with -DRM_EMU (one lane per group, threads run one after
another) into tests/emu/_rm_emu.so so that the kernels' sequential-semantics logic can be checked
against the oracle without a GPU.  Nothing in the product package can load this library.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "radixmesh_b200", "csrc")
OUT = os.path.join(HERE, "_rm_emu.so")
SOURCES = ["rm_api.cu", "rm_match.cu", "rm_insert.cu", "rm_maint.cu", "rm_route.cu", "rm_noalloc.cu", "rm_pages.cu"]


def build(force=False):
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    deps.append(os.path.join(ROOT, "include", "radixmesh_b200.h"))
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in deps):
        return OUT
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fPIC", "-shared", "-DRM_EMU", "-Wall", "-Wno-unused-function",
           "-Wno-unknown-pragmas", "-Wno-unused-variable", "-I", CSRC]
    for s in srcs:
        cmd += ["-x", "c++", s]
    cmd += ["-o", OUT]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
