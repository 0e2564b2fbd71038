#!/usr/bin/env python
"""Sweep-kernel time of the library named by DOMA_B200_LIB on the 256^3 x 512-ordinate case (no torch: one lean
process per library variant, used to A/B builds of sweep.cu).  Prints: <lib> <min ms of 4 sweeps> <pencil us per task>."""
import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from doma_b200 import abi
from doma_b200.synthetic import synthetic_case
n = int(os.environ.get("N", 256))
params, prob = synthetic_case(os.path.join(tempfile.gettempdir(), f"eval_{n}.xml"), n, G=8, mode="B")
ctx = abi.Context(prob)
ctx.reset_state(0.0)
ts = []
for g in range(5):
    ctx.update_mg_source(g % 8, 0.0); ctx.sweep(g % 8); ctx.synchronize(); ts.append(ctx.last_sweep_ms()); ctx.update_scatter_source(g % 8)
print(os.path.basename(os.environ.get("DOMA_B200_LIB", "default")), "%.3f" % min(ts[1:]), "tile %dx%dx%d" % (ctx.info(abi.INFO_TILE_X), ctx.info(abi.INFO_TILE_Y), ctx.info(abi.INFO_TILE_Z)), flush=True)
