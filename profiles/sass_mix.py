"""Aggregate `ncu -i X.ncu-rep --page source --csv` by SASS opcode (per kernel section).

usage: python profiles/sass_mix.py source_page.csv
"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = None
byop = collections.Counter()
samp = collections.Counter()
tot = 0
kernel = None


def flush():
    global byop, samp, tot
    if tot:
        print(f"== {kernel}\n   warp instructions executed: {tot}")
        for op, n in byop.most_common(32):
            print(f"   {op:10s} {n:10d} {100 * n / tot:5.1f}%  stall samples {samp[op]}")
    byop, samp, tot = collections.Counter(), collections.Counter(), 0


for r in rows:
    if r and r[0] == "Kernel Name":
        flush()
        kernel = r[1][:100]
        continue
    if r and r[0] == "Address":
        hdr = r
        ia, isrc, isamp = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
        continue
    if hdr is None or len(r) <= ia:
        continue
    src = r[isrc].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
    op = m.group(2).split(".")[0] if m else src
    n = int(r[ia] or 0)
    byop[op] += n
    samp[op] += int(r[isamp] or 0)
    tot += n
flush()
