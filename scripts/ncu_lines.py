#!/usr/bin/env python
"""Per-source-line instruction and stall-sample totals of one kernel in an ncu report (source page, CSV).
Usage: python scripts/ncu_lines.py report.ncu-rep kernel-regex [min_samples]"""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep, rx = sys.argv[1], sys.argv[2]
    min_s = int(sys.argv[3]) if len(sys.argv) > 3 else 300
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                          "regex:" + rx], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, cur = None, None
    inst, samp = collections.Counter(), collections.Counter()
    stall = collections.defaultdict(collections.Counter)
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r and r[0] == "Line No":
            hdr = r
            continue
        if hdr and cur and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            try:
                ln = int(d["Line No"])
                ie = int(d["Instructions Executed"] or 0)
                s = int(d["# Samples"] or 0)
            except ValueError:
                continue
            key = (cur, ln)
            inst[key] += ie
            samp[key] += s
            for k, v in d.items():
                if k.startswith("stall_") and "Not Issued" not in k and v:
                    try:
                        stall[key][k[6:]] += int(v)
                    except ValueError:
                        pass
    ti, ts = sum(inst.values()), sum(samp.values())
    print("total instructions %.1f M, samples %d" % (ti / 1e6, ts))
    by_file_i, by_file_s = collections.Counter(), collections.Counter()
    for k in inst:
        by_file_i[k[0]] += inst[k]
        by_file_s[k[0]] += samp[k]
    for f in by_file_i:
        print("  %-24s %8.1f M instr (%4.1f %%)  %7d samples (%4.1f %%)" % (f, by_file_i[f] / 1e6, 100.0 * by_file_i[f] / ti,
                                                                         by_file_s[f], 100.0 * by_file_s[f] / max(ts, 1)))
    allst = collections.Counter()
    for k in stall:
        allst.update(stall[k])
    print("stall reasons:", ", ".join("%s %.1f%%" % (k, 100.0 * v / max(ts, 1)) for k, v in allst.most_common(10)))
    for k in sorted(inst):
        if samp[k] >= min_s:
            top = ", ".join("%s %d" % kv for kv in stall[k].most_common(3))
            print("%-22s %5d  %8.1f M  %7d  %s" % (k[0], k[1], inst[k] / 1e6, samp[k], top))


if __name__ == "__main__":
    main()
