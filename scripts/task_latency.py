#!/usr/bin/env python
"""Single-warp task latency of the sweep kernel: a pencil mesh 8 x TY x (TZ * n) swept by ONE unit is a pure
chain of n tiles (tile k+1 waits for tile k), so sweep time / n = latency of one task executed by one warp
with an otherwise idle GPU.  Two tile depths separate the per-cell latency c from the per-task overhead o:
L = cells * c + o.  Used for the wavefront model in DESIGN.md (fill / drain of the tile hyperplanes)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from doma_b200 import abi
from doma_b200.synthetic import random_problem

VARIANT = {"flags": abi.SWEEP_FLAGS, "ll": abi.SWEEP_LL}[sys.argv[1] if len(sys.argv) > 1 else "flags"]

def chain_ms(ty, tz, n_tiles, full_lanes=True):
    prob = random_problem((8, ty, tz * n_tiles), G=1, npol=4, nazi=4, seed=3, void=False)
    keep = np.nonzero(prob.ord_octant == 0)[0][:32 if full_lanes else 8]
    prob = prob.with_ordinates(keep)
    ctx = abi.Context(prob)
    ctx.set_sweep_variant(VARIANT)
    ctx.set_tile(8, ty, tz)
    ctx.reset_state(0.0)
    ctx.set_array(abi.ARR_Q, 0, np.ones(prob.n_cells))
    ts = []
    for _ in range(5):
        ctx.sweep(0); ctx.synchronize(); ts.append(ctx.last_sweep_ms())
    return min(ts[1:]), int(ctx.info(abi.INFO_N_UNITS))

if __name__ == "__main__":
    n = 120
    res = {}
    print("protocol", sys.argv[1] if len(sys.argv) > 1 else "flags")
    for ty, tz in [(4, 2), (4, 8), (4, 32), (2, 2), (4, 1)]:
        ms, u = chain_ms(ty, tz, n)
        res[(ty, tz)] = ms / n * 1e3
        print(f"tile 8x{ty}x{tz}: units {u}, {n} chained tiles, {ms:.3f} ms -> {ms / n * 1e3:.2f} us per task ({8*ty*tz} cells)")
    c = (res[(4, 32)] - res[(4, 8)]) / (8 * 4 * 24)
    o = res[(4, 8)] - 256 * c
    print(f"per-cell latency c = {c*1e3:.0f} ns ({c*1e3*1.965:.0f} cycles at 1965 MHz), per-task overhead o = {o:.2f} us")
