#!/usr/bin/env python3
"""Sums the source page of an ncu report per source function and per source line.

    ncu -i rep.ncu-rep --page source --csv --print-source sass,cuda > src.csv
    python tools/ncu_by_function.py src.csv [--lines 40]

For every CUDA source line ncu gives the executed warp-instructions of the SASS attributed to it
(inlined code is attributed to the line it was written on); the enclosing function is found
by scanning the source file upwards for a definition.
"""
import csv
import os
import re
import sys
from collections import defaultdict

DEF = re.compile(r"^\s*(?:template\s*<[^>]*>\s*)?(?:KGM_HDM?|__device__|__global__|static|inline|__forceinline__|"
                 r"__noinline__|constexpr|\s)+[\w:<>\*&\s]+?\b(\w+)\s*\([^;]*$")


def functions_of(path):
    names = {}
    try:
        lines = open(path, errors="replace").read().split("\n")
    except OSError:
        return names
    cur = "?"
    for i, ln in enumerate(lines, 1):
        m = DEF.match(ln)
        if m and not ln.strip().startswith(("if", "for", "while", "return", "else", "//")):
            cur = m.group(1)
        names[i] = cur
    return names


def main():
    path = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--lines") + 1]) if "--lines" in sys.argv else 30
    per_fn, per_line, thr_fn, smp_fn = defaultdict(int), defaultdict(int), defaultdict(int), defaultdict(int)
    cur_file, fmap, cols = None, {}, None
    for row in csv.reader(open(path)):
        if not row:
            continue
        if row[0] == "File Path":
            cur_file = row[1]
            fmap = functions_of(cur_file)
            continue
        if row[0] == "Line No":
            cols = {n: i for i, n in enumerate(row)}
            continue
        if row[0] in ("Function Name",) or cols is None or not row[0].isdigit():
            continue
        ln = int(row[0])
        num = lambda name: int(row[cols[name]]) if row[cols[name]].isdigit() else 0
        inst, thr, smp = num("Instructions Executed"), num("Thread Instructions Executed"), num("# Samples")
        if not inst and not smp:
            continue
        fn = "%s:%s" % (os.path.basename(cur_file).replace("kgm_", "").split(".")[0], fmap.get(ln, "?"))
        per_fn[fn] += inst
        thr_fn[fn] += thr
        smp_fn[fn] += smp
        per_line[(os.path.basename(cur_file), ln, row[1].strip()[:90])] += inst
    total = sum(per_fn.values()) or 1
    tsmp = sum(smp_fn.values()) or 1
    print("total warp-instructions attributed: %d" % total)
    for fn, v in sorted(per_fn.items(), key=lambda kv: -kv[1])[:45]:
        print("%-36s inst=%9d %5.1f%%  samples %5.1f%%  thr/inst %.1f" %
              (fn, v, 100.0 * v / total, 100.0 * smp_fn[fn] / tsmp, thr_fn[fn] / v if v else 0))
    print()
    for (f, ln, src), v in sorted(per_line.items(), key=lambda kv: -kv[1])[:top]:
        print("%8d %5.1f%%  %s:%d  %s" % (v, 100.0 * v / total, f, ln, src))


if __name__ == "__main__":
    main()
