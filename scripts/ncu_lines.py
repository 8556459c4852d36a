#!/usr/bin/env python
"""Per-source-line instruction and stall totals for one kernel of an ncu report.
usage: ncu_lines.py report.ncu-rep libzsb.so 'k_inflateILi8E' [top]"""
import csv, re, subprocess, sys, tempfile, os, collections
rep, so, pat = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run("cd %s && cuobjdump -xelf all %s >/dev/null" % (tmp, os.path.abspath(so)), shell=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
# locate the .text section of the kernel
start = None
for i, l in enumerate(dis):
    if l.startswith("//--------------------- .text.") and pat in l:
        start = i
        break
addr2line = {}
cur = ("?", 0)
for l in dis[start + 1:]:
    if l.startswith("//--------------------- "):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        if "inlined at" in m.group(3) and False:
            pass
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);', l)
    if m:
        addr2line[int(m.group(1), 16)] = (cur, m.group(2))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw))
# one section per profiled launch: "Kernel Name" row, header row, instruction rows; take the LAST launch of the kernel
want = os.environ.get("NCU_KERNEL", "")
secs = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name" and want in r[1]]
hi = secs[-1] + 1
end = min([i for i, r in enumerate(rows) if r and r[0] == "Kernel Name" and i > hi] + [len(rows)])
rows = rows[:end]
hdr = rows[hi]
ia, ii, isamp, ithr = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
base = int(rows[hi + 1][ia], 16)
agg = collections.defaultdict(lambda: [0, 0, 0, collections.Counter()])
tot = [0, 0, 0]
for r in rows[hi + 1:]:
    if len(r) <= ii:
        continue
    off = int(r[ia], 16) - base
    key = addr2line.get(off, (("?", 0), ""))[0]
    a = agg[key]
    a[0] += int(r[ii]); a[1] += int(r[isamp]); a[2] += int(r[ithr])
    tot[0] += int(r[ii]); tot[1] += int(r[isamp]); tot[2] += int(r[ithr])
    for i, h in stall_cols:
        if r[i] and r[i] != "0":
            a[3][h] += int(r[i])
print("total inst %d  samples %d  thread-inst %d (%.1f lanes/inst)" % (tot[0], tot[1], tot[2], tot[2] / max(tot[0], 1)))
print("%-22s %7s %7s %6s  top stalls" % ("file:line", "inst%", "samp%", "lanes"))
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    st = ", ".join("%s %d" % (h[6:], c) for h, c in a[3].most_common(3))
    print("%-22s %6.2f%% %6.2f%% %6.1f  %s" % ("%s:%d" % key, 100.0 * a[0] / tot[0], 100.0 * a[1] / max(tot[1], 1), a[2] / max(a[0], 1), st))
