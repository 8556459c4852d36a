#!/usr/bin/env python
"""Extra (non-headline) bench lines: BASELINE.json configs C3, C4, C5 on one B200, members resident in HBM,
every member checked against the oracle.  usage: bench_configs.py [C3 C4 C5] [--path 0|1|2] [--small]
One JSON line per config on stdout."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="*", default=["C3", "C4", "C5"])
    ap.add_argument("--path", type=int, default=0, help="0 length-aware split (default), 1 lane path only, 2 group path only")
    ap.add_argument("--lanes", type=int, default=0)
    ap.add_argument("--small", action="store_true")
    ap.add_argument("--split", action="store_true", help="option split_big: large members to the group path")
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    import config_runner
    import workloads
    import zsblib
    full = {"C3": dict(distinct=8, tiles=512), "C4": dict(distinct=512, tiles=32), "C5": dict(distinct=2048, tiles=512)}
    small = {"C3": dict(distinct=8, tiles=32), "C4": dict(distinct=512, tiles=2), "C5": dict(distinct=2048, tiles=32)}
    for c in args.configs:
        w = workloads.BUILDERS[c](**(small if args.small else full)[c])
        eng = zsblib.Zsb(devices=[0])
        eng.set_option("path", args.path)
        if args.split:
            eng.set_option("split_big", 1)
        if args.lanes:
            eng.set_option("lanes_per_member", args.lanes)
        r = config_runner.run(w, eng, steps=args.steps, warmup=2)
        eng.close()
        r["config"] = c
        r["path"] = args.path
        r["parity"] = "ok" if not r["mismatches"] else "MISMATCH"
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
