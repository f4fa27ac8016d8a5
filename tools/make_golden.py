#!/usr/bin/env python
"""Generate tests/golden/*: (1) PCG32 known answers from the reference's recipe restated in pure Python
(tests/refmath.py, Extensions/Object+Extension.swift:556-589); (2) small oracle renders (RGBA8 + primary hit
ids) of reduced C1/C2/C3/C5 scenes.  The reference itself has no fixtures and cannot be built off macOS
(SURVEY.md F6/F7), so (2) freezes the ORACLE's output: it detects drift, it does not pin the oracle to the
reference ("parity unpinned").  Run from the repo root:  python tools/make_golden.py"""
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib  # noqa: E402
import refmath  # noqa: E402
from golden_cases import GOLDEN_CASES  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)

seeds = [0, 1, 42, 0x9E3779B97F4A7C15, refmath.pixel_seed(0, 0), refmath.pixel_seed(799, 799), refmath.pixel_seed(1919, 1079),
         0x123456789ABCDEF]
kat = {}
for s in seeds:
    r = refmath.PCG32(s)
    kat[str(s)] = [r.next() for _ in range(8)]
json.dump(kat, open(os.path.join(OUT, "pcg32.json"), "w"), indent=1)

with tempfile.TemporaryDirectory() as d:
    for name, (fn, kw) in GOLDEN_CASES.items():
        path = fn(d, **kw)
        r = oracle_lib.OracleScene(path).render(want_ids=True)
        ids = r["ids"]
        np.savez_compressed(os.path.join(OUT, f"{name}.npz"), rgba=r["rgba"], inst=ids["inst"].astype(np.int16),
                            kind=ids["kind"].astype(np.int8), prim=ids["prim"].astype(np.int32), t=ids["t"],
                            rays=np.asarray([r["counters"][k] for k in ("raysPrimary", "raysShadow", "raysSecondary")], np.int64))
        print(name, r["rgba"].shape, r["counters"])
