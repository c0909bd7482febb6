#!/usr/bin/env python3
"""Writes profiles/ncu_traffic.json from `ncu --set full` reports of THIS tree's kernels:

  python scripts/update_traffic.py --round r02_x stage=report.ncu-rep:kernel_regex [stage=...]

For every stage (bench.py's stage names: march, sample+network+composite, sampler, network, composite) the
first launch whose kernel name matches the regex gives dram__bytes_read.sum + dram__bytes_write.sum.  The file
records the hash of the CUDA sources it was captured for; bench.py refuses it when the sources have changed."""
import argparse
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def launches(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE,
                         stderr=subprocess.DEVNULL).stdout.decode(errors="replace")
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        yield r[col["Kernel Name"]], float(r[col["dram__bytes_read.sum"]]), float(r[col["dram__bytes_write.sum"]]), \
            rows[1][col["dram__bytes_read.sum"]]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--round", required=True)
    ap.add_argument("specs", nargs="+")
    a = ap.parse_args()
    unit_scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    traffic, captures = {}, []
    for spec in a.specs:
        stage, rest = spec.split("=", 1)
        rep, rx = rest.rsplit(":", 1)
        for name, rd, wr, unit in launches(rep):
            if re.search(rx, name):
                traffic[stage] = int((rd + wr) * unit_scale[unit])
                break
        else:
            raise SystemExit("no launch matching %r in %s" % (rx, rep))
        captures.append(os.path.basename(rep))
    out = dict(round=a.round, sources_sha16=bench.kernel_sources_sha16(), capture=", ".join(sorted(set(captures))),
               _source="dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full --clock-control none",
               traffic=traffic)
    with open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
