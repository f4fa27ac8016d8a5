#!/usr/bin/env python
"""Join an ncu `--page source --csv` SASS dump with nvdisasm line info and aggregate per source line.

    ncu -i prof.ncu-rep --page source --csv > src.csv
    cuobjdump -xelf all librtcuda.so ; nvdisasm -g -c rt_kernels.sm_100a.cubin > dis.txt
    python tools/ncu_by_line.py src.csv dis.txt 'renderKernelILb0' [top]

Prints, per file:line, warp-level instructions executed, average active threads, stall samples.
The cubin must be the one that was profiled.
"""
import csv
import re
import sys
from collections import defaultdict


def line_map(dis_path, kernel_pat):
    """address (int) -> 'file:line' for the function whose mangled name matches kernel_pat."""
    m, cur, inside = {}, "?", False
    addr_re = re.compile(r"/\*([0-9a-f]{4,})\*/")
    for ln in open(dis_path, errors="replace"):
        if ln.startswith("//---") and ".text." in ln:
            inside = re.search(kernel_pat, ln) is not None
            cur = "?"
            continue
        if not inside:
            continue
        if "//## File" in ln:
            mm = re.search(r'File "([^"]+)", line (\d+)', ln)
            if mm:
                cur = mm.group(1).split("/")[-1] + ":" + mm.group(2)
            continue
        a = addr_re.search(ln)
        if a and ";" in ln:
            m[int(a.group(1), 16)] = cur
    return m


def main():
    src, dis, pat = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    lm = line_map(dis, pat)
    rows = list(csv.reader(open(src)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    col = {n: i for i, n in enumerate(hdr)}
    agg = defaultdict(lambda: [0, 0, 0, 0])  # inst, thread inst, samples, n sass
    base = None
    tot = [0, 0, 0]
    for r in rows[hi + 1:]:
        if len(r) < len(hdr):
            continue
        try:
            addr = int(r[col["Address"]], 16) if r[col["Address"]].startswith("0x") else int(r[col["Address"]])
        except ValueError:
            continue
        if base is None:
            base = addr
        off = addr - base
        key = lm.get(off, "?")
        inst = int(float(r[col["Instructions Executed"]] or 0))
        tinst = int(float(r[col["Thread Instructions Executed"]] or 0))
        smp = int(float(r[col["# Samples"]] or 0))
        a = agg[key]
        a[0] += inst; a[1] += tinst; a[2] += smp; a[3] += 1
        tot[0] += inst; tot[1] += tinst; tot[2] += smp
    print(f"total warp-inst {tot[0]:,}  avg threads {tot[1] / max(1, tot[0]):.2f}  samples {tot[2]:,}")
    print(f"{'file:line':28s} {'inst%':>7s} {'samp%':>7s} {'avgthr':>7s} {'sass':>5s}")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
        print(f"{k:28s} {100 * a[0] / max(1, tot[0]):7.2f} {100 * a[2] / max(1, tot[2]):7.2f} {a[1] / max(1, a[0]):7.2f} {a[3]:5d}")


if __name__ == "__main__":
    main()
