#!/usr/bin/env python
"""Opcode histogram per kernel of libzsb.so's sm_100a cubin (cuobjdump -sass): memory opcodes, local-memory
traffic (LDL/STL = spills or local arrays), generic LD/ST against LDG/STG, async copies.  usage: sass_hist.py [lib]"""
import collections, re, subprocess, sys
lib = sys.argv[1] if len(sys.argv) > 1 else "zlib-streams_b200/libzsb.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        hist[kern][m.group(1).split(".")[0]] += 1
watch = ["LDG", "STG", "LD", "ST", "LDL", "STL", "LDS", "STS", "LDGSTS", "ATOMS", "ATOMG", "RED", "SHFL", "VOTE", "MATCH", "BREV", "POPC", "BAR", "BRA", "BSSY"]
print("%-34s %7s  %s" % ("kernel", "insts", "  ".join("%s" % w for w in watch)))
for k, h in hist.items():
    print("%-34s %7d  %s" % (k[:34], sum(h.values()), "  ".join("%*d" % (len(w), h.get(w, 0)) for w in watch)))
