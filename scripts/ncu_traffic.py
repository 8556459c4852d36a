#!/usr/bin/env python
"""Writes profiles/r02_traffic.json: per bench configuration, the summed dram__bytes_read.sum + dram__bytes_write.sum of
the step's decode kernels (one launch of each = one step), read from the committed ncu summaries
(profiles/r02_ncu_summary.txt for C2: k_lane + k_lz; profiles/r02_ncu_k_big_k_check_summary.txt for C4's k_big).
bench.py's roofline.traffic reads the file.  usage: python scripts/ncu_traffic.py"""
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def dram_bytes(path):
    """{kernel column index: bytes} summed over read + write, and the kernel names of the summary's header line."""
    names, total = None, None
    for line in open(path):
        if line.startswith("Kernel Name"):
            names = re.split(r"\s{2,}", line.strip())[1:]
        m = re.match(r"(dram__bytes_(?:read|write)\.sum)\s+(\w+)\s+(.*)", line)
        if m:
            vals = m.group(3).split()
            total = total or [0.0] * len(vals)
            for i, v in enumerate(vals):
                try:
                    x = float(v)
                except ValueError:
                    x = float("nan")
                total[i] += x * UNIT[m.group(2)]
    return names, total


def main():
    out = {}
    names, tot = dram_bytes(os.path.join(ROOT, "profiles", "r02_ncu_summary.txt"))
    out["C2"] = int(sum(t for t in tot if t == t))
    out["C2_kernels"] = dict(zip(names, [int(t) for t in tot if t == t]))
    names, tot = dram_bytes(os.path.join(ROOT, "profiles", "r02_ncu_k_big_k_check_summary.txt"))
    out["C4_k_big_only"] = int(tot[0])
    for c in ("C3", "C4", "C5"):
        out[c] = None              # no whole-step capture of these configurations
    out["source"] = "ncu --set full --clock-control none captures summarised in profiles/r02_ncu_summary.txt (C2: one launch of each decode kernel = one step)"
    json.dump(out, open(os.path.join(ROOT, "profiles", "r02_traffic.json"), "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
