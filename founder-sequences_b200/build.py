"""In-tree build of libfsq_b200.so (nvcc, sm_100a only) and of the test oracle.

    python founder-sequences_b200/build.py [--force]

The shared library lands in founder-sequences_b200/lib/ (git-ignored, but it travels to the GPU box).
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libfsq_b200.so")
SOURCES = ["fsq_capi.cu", "fsq_pbwt.cu", "fsq_dp.cu", "fsq_pack.cu", "fsq_comm.cu", "fsq_match.cu", "fsq_noalloc.cu",
           "fsq_cscan_e9.cu", "fsq_cscan_e13.cu", "fsq_cscan_e17.cu", "fsq_cscan_e21.cu", "fsq_cscan_e25.cu", "fsq_cscan_e27.cu",
           "fsq_cscan_e29.cu", "fsq_cscan_e31.cu", "fsq_cscan_e33.cu", "fsq_cscan_e53.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "--expt-relaxed-constexpr",
]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed: %s\n%s" % (" ".join(cmd), r.stdout))
    return r.stdout


def build_library(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh", ".inc"))]
    headers.append(os.path.join(ROOT, "include", "fsq.h"))
    objs, jobs = [], []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(LIBDIR, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _newer(o, [s] + headers):
            cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
            jobs.append(cmd)
    with ThreadPoolExecutor(max_workers=8) as ex:
        for out in ex.map(_run, jobs):
            if verbose and out:
                print(out)
    if force or jobs or _newer(LIB, objs):
        _run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs)
    return LIB


def build_oracle():
    """The checkers are test infrastructure; building them is not using them."""
    _run(["make", "-C", os.path.join(ROOT, "oracle"), "libfsq_oracle.so"])
    if os.path.isdir("/root/reference/founder-sequences"):
        _run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"])
        _run(["make", "-C", os.path.join(ROOT, "host"), "ref"])


def build_host():
    """host/_build/founder_sequences_b200: the command-line host driver.  It compiles the reference's joiner and writer
    sources where they lie, so it is built where the reference tree exists; the binary travels with the snapshot."""
    if os.path.isdir("/root/reference/founder-sequences"):
        _run(["make", "-C", os.path.join(ROOT, "host"), "all"])


if __name__ == "__main__":
    force = "--force" in sys.argv
    print(build_library(force=force, verbose="-v" in sys.argv))
    build_host()
    build_oracle()
