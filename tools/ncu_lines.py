#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump per CUDA source line:
samples, instructions, average active threads.  Usage: ncu_lines.py dump.csv [top]"""
import csv
import sys
import collections

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
csv.field_size_limit(1 << 30)
cur_file = None
hdr = None
agg = collections.OrderedDict()
kernel = 0
for r in csv.reader(open(path, errors="replace")):
    if not r:
        continue
    if r[0] in ("Kernel Name", "Function Name"):
        continue
    if r[0] in ("File Name", "File Path"):
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
        continue
    d = dict(zip(hdr[:2] + hdr[2:], r))
    try:
        line = int(r[0])
        samples = int(r[hdr.index("# Samples")] or 0)
        inst = int(r[hdr.index("Instructions Executed")] or 0)
        tinst = int(r[hdr.index("Thread Instructions Executed")] or 0)
        lsb = int(r[hdr.index("stall_long_sb")] or 0)
    except ValueError:
        continue
    k = (cur_file, line)
    a = agg.setdefault(k, [0, 0, 0, 0, r[1]])
    a[0] += samples; a[1] += inst; a[2] += tinst; a[3] += lsb
tot = sum(a[0] for a in agg.values()) or 1
toti = sum(a[1] for a in agg.values()) or 1
print("total samples %d, instructions %d, avg threads %.2f" % (tot, toti, sum(a[2] for a in agg.values()) / toti))
by_file = collections.defaultdict(lambda: [0, 0, 0])
for (f, l), a in agg.items():
    by_file[f][0] += a[0]; by_file[f][1] += a[1]; by_file[f][2] += a[2]
for f, v in sorted(by_file.items(), key=lambda x: -x[1][0]):
    print("  %-18s samples %5.1f%%  inst %5.1f%%  avg threads %.1f" % (f, 100 * v[0] / tot, 100 * v[1] / toti, v[2] / max(v[1], 1)))
print("%-22s %7s %7s %7s %6s  %s" % ("file:line", "samp%", "inst%", "thr/inst", "lsb%", "source"))
for (f, l), a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print("%-22s %6.2f%% %6.2f%% %7.1f %5.0f%%  %s" % ("%s:%d" % (f, l), 100 * a[0] / tot, 100 * a[1] / toti, a[2] / max(a[1], 1),
                                                   100 * a[3] / max(a[0], 1), a[4].strip()[:90]))
