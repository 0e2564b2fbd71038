#!/usr/bin/env python
"""Sweep time of the small shipped cases per tile choice (the wavefront is never saturated there: the
tile that minimises hyperplanes x task latency wins, not the one with most tasks in flight)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import doma_b200 as d
from doma_b200 import abi
cases = os.path.join(ROOT, "tests", "golden", "cases")
for f, tiles in [("full_tr2d_1step.xml", [(0, 0), (1, 1), (2, 1), (4, 1)]), ("full_asm_fs_3outer.xml", [(0, 0), (2, 2), (4, 2), (4, 4), (2, 4), (2, 1), (4, 1), (1, 2)]),
                 ("full_k1.xml", [(0, 0), (2, 2), (4, 2), (4, 4), (4, 8)])]:
  params, prob = d.load_case(os.path.join(cases, f))
  for vname, variant in (("auto", abi.SWEEP_AUTO), ("flags", abi.SWEEP_FLAGS), ("ll", abi.SWEEP_LL)):
    for ty, tz in (tiles if vname != "auto" else [(0, 0)]):
        ctx = abi.Context(prob)
        ctx.set_sweep_variant(variant)
        if ty: ctx.set_tile(8, ty, tz)
        ctx.reset_state(0.0)
        ctx.set_array(abi.ARR_Q, 0, np.ones(prob.n_cells))
        ts = []
        for _ in range(6):
            ctx.sweep(0); ctx.synchronize(); ts.append(ctx.last_sweep_ms())
        print(f, vname, "->", "ll" if ctx.info(abi.INFO_SWEEP_VARIANT) == abi.SWEEP_LL else "flags", "tile 8x%dx%d%s" % (ctx.info(abi.INFO_TILE_Y), ctx.info(abi.INFO_TILE_Z), "" if ty else " (auto)"), "units", ctx.info(abi.INFO_N_UNITS),
              "sweep us %.0f" % (min(ts[1:]) * 1e3), flush=True)
