#!/usr/bin/env python
"""Per-kernel SASS mnemonic counts of the shipped library (cuobjdump -sass), the static evidence that the
network kernels run on tcgen05 / TMEM (UTCHMMA = tcgen05.mma kind::f16, LDTM / STTM = tcgen05.ld / .st,
UTCBAR = tcgen05.commit, SYNCS = mbarrier traffic) and how long each kernel's instruction stream is.

Usage: python scripts/sass_counts.py [path/to/librtxnerf_b200.so] > profiles/rNN_sass_mnemonics.txt"""
import collections
import os
import re
import subprocess
import sys

WATCH = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UBLKCP", "SYNCS", "HMMA", "HFMA2", "LDG", "STG",
         "RED", "ATOM", "LDS", "STS", "BAR", "MUFU", "FFMA", "IMAD", "LEA", "LOP3", "STL", "LDL"]


def main():
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(here, "rtx_nerf_b200", "librtxnerf_b200.so")
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    names = {}
    try:
        demangled = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)),
                                   capture_output=True, text=True, check=True).stdout.splitlines()
        for raw, nice in zip(re.findall(r"Function : (\S+)", sass), demangled):
            names[raw] = nice
    except (OSError, subprocess.CalledProcessError):
        pass
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur is not None:
            cur["_total"] += 1
            cur[m.group(1)] += 1
    print(f"# cuobjdump -sass {os.path.relpath(lib, here)}: instruction counts per kernel (static, not executed counts)")
    print("# UTCHMMA = tcgen05.mma kind::f16, LDTM/STTM = tcgen05.ld/st, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops")
    for raw, c in kernels.items():
        nice = names.get(raw, raw)
        nice = re.sub(r"\((?!bool\)|int\)).*", "", nice)
        cols = "  ".join(f"{k}={c[k]}" for k in WATCH if c[k])
        print(f"{nice}\n    total={c['_total']}  {cols}")


if __name__ == "__main__":
    main()
