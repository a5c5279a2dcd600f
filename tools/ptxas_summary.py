#!/usr/bin/env python
"""Registers, spill bytes and SASS size per kernel of trace_fast.cu / trace_exact.cu (no GPU needed).

    python tools/ptxas_summary.py [extra nvcc flags...]        # e.g. -DRT_MIN_BLOCKS=6
"""
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
SRC = ROOT / "python_ray_tracer_b200" / "csrc"
FLAGS = "-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr -Xptxas -v".split()


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return [re.sub(r"\(rt::SceneParams.*", "", o).replace("void rt::", "").replace("(int)", "").replace("(bool)", "") for o in out]


def main():
    extra = sys.argv[1:]
    for tu in ("trace_fast.cu",):
        cubin = f"/tmp/ptxas_summary_{tu}.cubin"
        r = subprocess.run(["/usr/local/cuda/bin/nvcc", *FLAGS, *extra, "-cubin", "-o", cubin, str(SRC / tu)], capture_output=True, text=True)
        text = r.stderr
        if r.returncode:
            print(text[-3000:])
            sys.exit(1)
        entries = re.findall(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*?\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers", text)
        sizes = {}
        sass = subprocess.run(["cuobjdump", "-sass", cubin], capture_output=True, text=True).stdout
        for m in re.finditer(r"Function : (\S+)\n(.*?)(?=\n\t\tFunction :|\Z)", sass, re.S):
            sizes[m.group(1)] = len(re.findall(r"^\s+/\*[0-9a-f]{4,}\*/\s+\S", m.group(2), re.M))
        names = demangle([e[0] for e in entries])
        for (mangled, stack, st, ld, regs), name in sorted(zip(entries, names), key=lambda x: x[1]):
            print(f"{name:60s} regs {regs:>3s}  stack {stack:>5s}  spill st/ld {st:>4s}/{ld:>4s}  sass {sizes.get(mangled, 0):5d}")


if __name__ == "__main__":
    main()
