#!/usr/bin/env python3
"""Build recipe for the oracle (TEST INFRASTRUCTURE ONLY).

  oracle/liboracle.so          our C restatement (rtxnerf_oracle.c)
  oracle/_ref/libref_host.so   the REFERENCE's own sources compiled for the host:
                                 sampler/sampler.cu, vol_render/vol_render.cu and
                                 rtx/src/optixPrograms.cu, read where they lie under
                                 /root/reference, over oracle/host_shim (CUDA-on-host
                                 shim + OptiX runtime emulation)
  oracle/_ref/libref_cuda.so   the reference's sampler.cu / vol_render.cu compiled
                                 unchanged by nvcc for sm_100a (runs on the GPU box)

No reference source is copied into the repo: each reference file is read, two textual
rewrites that plain C++ needs are applied IN MEMORY, and the text is piped to g++ on stdin.
The rewrites:
  kernel<<<grid, block, ...>>>(   ->  rtx_host_launch(grid, block, kernel,
  extern "C" static __constant__   ->  static      (g++ rejects static in a linkage spec)
The _ref outputs are git-ignored but travel to the GPU box with gpurun.
"""
import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("RTXNERF_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
SHIM = os.path.join(HERE, "host_shim")
CUDA_INC = os.environ.get("CUDA_HOME", "/usr/local/cuda") + "/include"

# -march=x86-64-v3: FMA + F16C without tying the .so to this container's CPU model
HOST_FLAGS = ["-O3", "-march=x86-64-v3", "-fopenmp", "-fPIC"]

_LAUNCH = re.compile(r"(\w+)\s*<<<\s*([^,>]+?)\s*,\s*([^,>]+?)\s*(?:,[^>]*)?>>>\s*\(")


def _run(cmd, **kw):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, **kw)
    if r.returncode != 0:
        sys.stderr.write(r.stdout.decode(errors="replace"))
        raise RuntimeError("command failed: " + " ".join(cmd))
    return r.stdout.decode(errors="replace")


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.exists(s) and os.path.getmtime(s) <= t for s in sources)


def build_oracle(force=False):
    src = os.path.join(HERE, "rtxnerf_oracle.c")
    hdr = os.path.join(HERE, "rtxnerf_oracle.h")
    out = os.path.join(HERE, "liboracle.so")
    if not force and _newer(out, [src, hdr]):
        return out
    # -ffp-contract=off: the oracle spells every fused multiply-add as fmaf()
    _run(["gcc", "-std=gnu11", *HOST_FLAGS, "-ffp-contract=off", "-shared", src, "-o", out, "-lm"])
    return out


def _host_tu(ref_rel, glue):
    """Reference source text (rewritten in memory) + our glue include."""
    with open(os.path.join(REF, ref_rel), "r") as f:
        text = f.read()
    text = _LAUNCH.sub(lambda m: "rtx_host_launch(%s, %s, %s, " % (m.group(2), m.group(3), m.group(1)), text)
    text = text.replace('extern "C" static __constant__', "static")
    return text + '\n#include "%s"\n' % glue


def build_ref_host(force=False):
    if not os.path.isdir(REF):
        return None
    os.makedirs(OUT, exist_ok=True)
    out = os.path.join(OUT, "libref_host.so")
    units = [
        ("sampler/sampler.cu", "ref_sampler_glue.inc", ["-I", os.path.join(REF, "sampler")]),
        ("vol_render/vol_render.cu", "ref_volrender_glue.inc", ["-I", os.path.join(REF, "vol_render")]),
        ("rtx/src/optixPrograms.cu", "ref_optix_glue.inc", ["-I", os.path.join(REF, "rtx/include")]),
    ]
    deps = [os.path.join(REF, u[0]) for u in units] + [os.path.join(SHIM, f) for f in os.listdir(SHIM)]
    deps.append(os.path.abspath(__file__))
    if not force and _newer(out, deps):
        return out
    objs = []
    for rel, glue, inc in units:
        obj = os.path.join(OUT, os.path.basename(rel) + ".host.o")
        # nvcc contracts a*b+c to fma by default (-fmad=true); g++ -ffp-contract=fast is the
        # closest host equivalent.  THRUST_DEVICE_SYSTEM=CPP: thrust::random only, no CUDA.
        cmd = ["g++", "-std=c++17", *HOST_FLAGS, "-ffp-contract=fast", "-w",
               "-DTHRUST_DEVICE_SYSTEM=THRUST_DEVICE_SYSTEM_CPP",
               "-include", os.path.join(SHIM, "cuda_host_shim.h"),
               "-I", SHIM, "-I", CUDA_INC, *inc, "-x", "c++", "-c", "-", "-o", obj]
        _run(cmd, input=_host_tu(rel, glue).encode())
        objs.append(obj)
    _run(["g++", "-shared", "-fopenmp", *objs, "-o", out])
    return out


def build_ref_cuda(force=False):
    if not os.path.isdir(REF) or shutil.which("nvcc") is None:
        return None
    os.makedirs(OUT, exist_ok=True)
    out = os.path.join(OUT, "libref_cuda.so")
    srcs = [os.path.join(REF, "sampler/sampler.cu"), os.path.join(REF, "vol_render/vol_render.cu"),
            os.path.join(SHIM, "ref_cuda_glue.cu")]
    if not force and _newer(out, srcs + [os.path.abspath(__file__)]):
        return out
    _run(["nvcc", "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
          "-Xcompiler", "-fPIC", "-shared", "-w",
          "-I", SHIM, "-I", os.path.join(REF, "sampler"), "-I", os.path.join(REF, "vol_render"),
          *srcs, "-o", out])
    return out


def build_all(force=False):
    res = {"oracle": build_oracle(force)}
    res["ref_host"] = build_ref_host(force)
    res["ref_cuda"] = build_ref_cuda(force)
    return res


if __name__ == "__main__":
    for k, v in build_all(force="--force" in sys.argv).items():
        print("%-9s %s" % (k, v if v else "skipped (reference not present)"))
