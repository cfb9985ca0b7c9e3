#!/usr/bin/env python
"""Join an ncu --page source --csv dump (SASS rows of one kernel) with nvdisasm -g line info of the same cubin:
instructions executed / stall samples per CUDA source line.  usage: ncu_lines.py <source.csv> <cubin> <mangled kernel> [top]"""
import csv, re, subprocess, sys, collections
src_csv, cubin, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith(".text." + kern + ":"))
line_of = []
cur = None
for l in dis[start + 1:]:
    if l.startswith("\t.section") or l.startswith("//-----"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        line_of.append(cur)
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
body = rows[2:]
iI, iS, iT = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
assert len(body) == len(line_of), (len(body), len(line_of))
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
for r, lo in zip(body, line_of):
    a = agg[lo]
    a[0] += int(r[iI]); a[1] += int(r[iS]); a[2] += int(r[iT]); a[3] += 1
totI = sum(a[0] for a in agg.values()); totS = sum(a[1] for a in agg.values())
print("total warp instructions %d, samples %d, sass instructions %d" % (totI, totS, len(body)))
srcs = {}
def text(f, n):
    if f not in srcs:
        import glob
        p = glob.glob("/root/repo/fast-livo-noted_b200/csrc/" + f)
        srcs[f] = open(p[0]).read().splitlines() if p else []
    return srcs[f][n - 1].strip()[:90] if 0 < n <= len(srcs[f]) else ""
for lo, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% inst %5.1f%% smp  thr/inst %4.1f  sass %3d  %s:%d  %s" % (100.0 * a[0] / totI, 100.0 * a[1] / max(totS, 1), a[2] / max(a[0], 1), a[3], lo[0], lo[1], text(*lo)))
