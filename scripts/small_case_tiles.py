#!/usr/bin/env python
"""Sweep time of the small shipped cases per tile choice and face protocol (the wavefront is never saturated there:
the tile that minimises hyperplanes x task latency wins, not the one with most tasks in flight).

    python scripts/small_case_tiles.py [tr2d|asm|k1 ...] [--quick]

DOMA_B200_LIB picks an experimental build, DOMA_B200_MAX_BLOCKS caps the CTAs of a launch."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import doma_b200 as d
from doma_b200 import abi
cases = os.path.join(ROOT, "tests", "golden", "cases")
quick = "--quick" in sys.argv
want = [a for a in sys.argv[1:] if not a.startswith("--")]
# (tx, ty, tz) tiles of the full grid; quick lists are (protocol, tx, ty, tz)
ALL = {"tr2d": ("full_tr2d_1step.xml", [(8, 1, 1), (8, 2, 1), (8, 4, 1), (4, 1, 1), (4, 2, 1), (4, 4, 1), (4, 8, 1)],
                [("ll", 8, 4, 1), ("ll", 4, 4, 1), ("ll", 4, 8, 1), ("ll", 4, 2, 1), ("flags", 4, 4, 1)]),
       "asm": ("full_asm_fs_3outer.xml", [(8, 2, 2), (8, 4, 2), (8, 4, 4), (8, 2, 4), (8, 2, 1), (8, 4, 1), (8, 1, 2), (4, 4, 1), (4, 8, 1), (4, 4, 2), (4, 2, 1), (4, 8, 2)],
               [("flags", 8, 4, 1), ("flags", 8, 2, 1), ("flags", 4, 4, 1), ("flags", 4, 8, 1), ("flags", 4, 4, 2), ("flags", 4, 2, 1), ("ll", 4, 4, 1), ("ll", 4, 8, 1)]),
       "k1": ("full_k1.xml", [(8, 2, 2), (8, 4, 2), (8, 4, 4), (8, 4, 8), (4, 4, 2), (4, 8, 2), (4, 8, 4)],
              [("flags", 8, 4, 2), ("flags", 4, 8, 2), ("flags", 4, 4, 2), ("flags", 4, 8, 4)])}
V = {"auto": abi.SWEEP_AUTO, "flags": abi.SWEEP_FLAGS, "ll": abi.SWEEP_LL}
for key, (f, tiles, quick_list) in ALL.items():
    if want and key not in want:
        continue
    params, prob = d.load_case(os.path.join(cases, f))
    runs = [("auto", 8, 0, 0)] + (quick_list if quick else [(v, tx, ty, tz) for v in ("flags", "ll") for tx, ty, tz in tiles])
    for vname, tx, ty, tz in runs:
        ctx = abi.Context(prob)
        ctx.set_sweep_variant(V[vname])
        if ty: ctx.set_tile(tx, ty, tz)
        ctx.reset_state(0.0)
        ctx.set_array(abi.ARR_Q, 0, np.ones(prob.n_cells))
        ts = []
        for _ in range(8):
            ctx.sweep(0); ctx.synchronize(); ts.append(ctx.last_sweep_ms())
        print(key, vname, "->", "ll" if ctx.info(abi.INFO_SWEEP_VARIANT) == abi.SWEEP_LL else "flags", "tile %dx%dx%d%s" % (ctx.info(abi.INFO_TILE_X), ctx.info(abi.INFO_TILE_Y), ctx.info(abi.INFO_TILE_Z), "" if ty else " (auto)"), "rows", ctx.info(abi.INFO_ROWS_PER_STEP), "units", ctx.info(abi.INFO_N_UNITS),
              "sweep us %.0f" % (min(ts[1:]) * 1e3), "median %.0f" % (float(np.median(ts[1:])) * 1e3), flush=True)
        ctx.close()
