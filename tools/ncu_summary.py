#!/usr/bin/env python
"""Summarise an .ncu-rep of the render kernel: headline raw metrics, stall mix, and (with the object file
that was profiled) instructions / samples / active threads per source file plus the executed-code footprint.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [myraytracer_b200/csrc/build/rt_kv_fast.o renderKernelFastILb0] [top]
"""
import csv
import os
import subprocess
import sys
import tempfile
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ncu_by_line import line_map  # noqa: E402

SORT = 0 if os.environ.get("NCU_SORT") == "inst" else 2
KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed.avg.per_cycle_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"]


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, val = rows[0], rows[1], rows[2]
    d = {h: (v, u) for h, u, v in zip(hdr, units, val)}
    print("== raw metrics")
    for k in KEYS:
        if k in d:
            print(f"  {k:78s} {d[k][0]:>16s} {d[k][1]}")
    print("== stall reasons (warps per issue-active cycle)")
    st = [(float(v[0]), h) for h, v in d.items() if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and v[0]]
    for v, h in sorted(st, reverse=True)[:9]:
        print(f"  {h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]:28s} {v:8.3f}")


def source(rep, obj, pat, top):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
    cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    dis = os.path.join(tmp, "dis.txt")
    with open(dis, "w") as f:
        subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], stdout=f, stderr=subprocess.DEVNULL)
    lm = line_map(dis, pat)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    col = {n: i for i, n in enumerate(hdr)}
    base = None
    byfile, byline = defaultdict(lambda: [0, 0, 0, 0]), defaultdict(lambda: [0, 0, 0, 0])
    tot = [0, 0, 0, 0, 0]
    for r in rows[hi + 1:]:
        if len(r) < len(hdr):
            continue
        a = r[col["Address"]]
        try:
            a = int(a, 16) if a.startswith("0x") else int(a)
        except ValueError:
            continue
        if base is None:
            base = a
        key = lm.get(a - base, "?")
        inst = int(float(r[col["Instructions Executed"]] or 0))
        ti = int(float(r[col["Thread Instructions Executed"]] or 0))
        sm = int(float(r[col["# Samples"]] or 0))
        for dct, k in ((byfile, key.split(":")[0]), (byline, key)):
            e = dct[k]
            e[0] += inst; e[1] += ti; e[2] += sm; e[3] += 1 if inst > 0 else 0
        tot[0] += inst; tot[1] += ti; tot[2] += sm; tot[3] += 1; tot[4] += 1 if inst > 0 else 0
    print(f"== source: {tot[0]:,} warp-inst, avg active threads {tot[1] / max(1, tot[0]):.2f}, SASS {tot[3]} "
          f"(executed {tot[4]} = {tot[4] * 16 / 1024:.1f} KB)")
    for title, dct, n in (("by file", byfile, 20), ("by line", byline, top)):
        print(f"-- {title}: inst%  samples%  avg-threads  executed-SASS")
        for k, e in sorted(dct.items(), key=lambda kv: -kv[1][SORT])[:n]:
            print(f"  {k:30s} {100 * e[0] / max(1, tot[0]):6.2f} {100 * e[2] / max(1, tot[2]):6.2f} {e[1] / max(1, e[0]):6.1f} {e[3]:6d}")


if __name__ == "__main__":
    raw(sys.argv[1])
    if len(sys.argv) > 3:
        source(sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]) if len(sys.argv) > 4 else 30)
