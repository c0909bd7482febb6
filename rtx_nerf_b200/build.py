#!/usr/bin/env python3
"""Builds rtx_nerf_b200/librtxnerf_b200.so — every CUDA kernel of the hot path plus the
C ABI (include/rtxnerf.h) — for sm_100a only.  In-tree so the .so travels with gpurun."""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "librtxnerf_b200.so")

NVCC_FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v"]


if os.environ.get("FN_TRACE") == "1":     # development: role timeline of network_freq.cu (scripts/freqnet_trace.py)
    NVCC_FLAGS.append("-DFN_TRACE")


for _k in ("FN_EPI_WARPS", "FN_ENC_WARPS"):   # development: role layout of network_freq.cu
    if os.environ.get(_k):
        NVCC_FLAGS.append("-D%s=%d" % (_k, int(os.environ[_k])))


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def build(force=False, verbose=False):
    srcs = sources()
    deps = srcs + glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(HERE, "..", "include", "rtxnerf.h"),
                                                             os.path.abspath(__file__)]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(d) <= os.path.getmtime(OUT) for d in deps):
        return OUT
    objs = []
    log = []
    for s in srcs:
        o = os.path.join(CSRC, os.path.basename(s)[:-3] + ".o")
        r = subprocess.run(["nvcc", *NVCC_FLAGS, "-c", s, "-o", o], stdout=subprocess.PIPE,
                           stderr=subprocess.STDOUT)
        log.append(r.stdout.decode(errors="replace"))
        if r.returncode != 0:
            sys.stderr.write(log[-1])
            raise RuntimeError("nvcc failed on " + s)
        objs.append(o)
    r = subprocess.run(["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", *objs, "-o", OUT], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        sys.stderr.write(r.stdout.decode(errors="replace"))
        raise RuntimeError("link failed")
    with open(os.path.join(HERE, "build.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    return OUT


def build_compat_driver(force=False):
    """tests/cpp/compat_driver: a C++ host program written against the reference's own headers
    and call sequence, compiled with -Icompat and linked to librtxnerf_b200.so."""
    root = os.path.dirname(HERE)
    src = os.path.join(root, "tests", "cpp", "compat_driver.cu")
    out = os.path.join(root, "tests", "cpp", "compat_driver")
    deps = [src, OUT] + glob.glob(os.path.join(root, "compat", "**", "*.h"), recursive=True)
    if not force and os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return out
    r = subprocess.run(["nvcc", "-std=c++17", "-O2", "-gencode", "arch=compute_100a,code=sm_100a",
                        "-I", os.path.join(root, "compat"), "-I", root, src, "-o", out, "-L", HERE, "-lrtxnerf_b200",
                        "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../../rtx_nerf_b200"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        sys.stderr.write(r.stdout.decode(errors="replace"))
        raise RuntimeError("compat_driver build failed")
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_compat_driver(force="--force" in sys.argv))
