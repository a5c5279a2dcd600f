"""Per-source-line totals from `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`
(first kernel only): warp instructions executed and stall samples per CUDA line.

usage: python profiles/line_mix.py cuda_sass_page.csv [top_n]
"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 60
path, seen_funcs, hdr = None, [], None
acc = collections.OrderedDict()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        path = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        if r[1] not in seen_funcs:
            seen_funcs.append(r[1])
        func = seen_funcs.index(r[1])
        continue
    if r[0] == "Line No":
        hdr = r
        ii = hdr.index("Instructions Executed")
        isamp = hdr.index("# Samples")
        continue
    if hdr is None or func != 0 or r[2] != "-":  # only the per-line summary rows of kernel 0
        continue
    key = (path, int(r[0]))
    inst, samp = int(r[ii] or 0), int(r[isamp] or 0)
    if key in acc:
        acc[key][0] += inst
        acc[key][1] += samp
    else:
        acc[key] = [inst, samp, r[1].strip()[:110]]
tot_i = sum(v[0] for v in acc.values())
tot_s = sum(v[1] for v in acc.values())
print(f"kernel: {seen_funcs[0][:90]}\nwarp instructions {tot_i}, stall samples {tot_s}")
for (p, ln), (inst, samp, src) in sorted(acc.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100 * inst / tot_i:5.1f}% inst {100 * samp / max(tot_s, 1):5.1f}% stall  {p}:{ln:<4d} {src}")
