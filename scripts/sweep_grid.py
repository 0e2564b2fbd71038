#!/usr/bin/env python
"""Sweep time over (world size emulated on one GPU, protocol, tile) with ONE problem build per world size.

    python scripts/sweep_grid.py [--n 256] [--worlds 1,2,4,8] [--variants flags,ll] [--tiles 4x8,4x4,4x2,4x1,2x2,2x1] [--out file]

For world W the context holds rank 0's ordinate shard of a W-rank run (doma_b200.sharding), exactly what one rank
of a real W-GPU job sweeps; no all-reduce.  Each (protocol, tile): 2 warm-up sweeps + 5 timed (CUDA events inside
the library, doma_b200_last_sweep_ms); prints min and median, and the max relative difference of phi_cur against
the first configuration of the same world (the protocols and tiles must agree to RED-order noise).
"""
import argparse
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from doma_b200 import abi  # noqa: E402
from doma_b200.sharding import shard_ordinates  # noqa: E402
from doma_b200.synthetic import synthetic_case  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=256)
ap.add_argument("--worlds", default="1,2,4,8")
ap.add_argument("--variants", default="flags,ll")
ap.add_argument("--tiles", default="4x8,4x4,4x2,4x1,2x2,2x1")
ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep_grid.txt"))
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
V = {"flags": abi.SWEEP_FLAGS, "ll": abi.SWEEP_LL, "auto": abi.SWEEP_AUTO}
os.makedirs(os.path.dirname(a.out), exist_ok=True)
out = open(a.out, "a")


def say(s):
    print(s, flush=True)
    out.write(s + "\n")
    out.flush()


params, full = synthetic_case(os.path.join(tempfile.gettempdir(), f"grid_{a.n}.xml"), a.n, G=8, mode="B")
q = None
for W in [int(w) for w in a.worlds.split(",")]:
    prob = full if W == 1 else full.with_ordinates(shard_ordinates(full, 0, W))
    ctx = abi.Context(prob)
    ctx.reset_state(0.0)
    ctx.update_mg_source(0, 0.0)           # the case's own source for group 0
    ref = None
    for vname in a.variants.split(","):
        for tile in a.tiles.split(","):
            ctx.set_sweep_variant(V[vname])
            if tile == "auto":
                ctx.set_tile(8, 0, 0)
            else:
                t = [int(v) for v in tile.split("x")]          # TYxTZ (8-cell rows) or TXxTYxTZ
                ctx.set_tile(*(t if len(t) == 3 else [8] + t))
            ts = []
            for i in range(2 + a.reps):
                if i == 1 + a.reps:         # last repetition: from a clean phi_cur, for the comparison
                    ctx.set_array(abi.ARR_PHI_CUR, 0, np.zeros(prob.n_cells))
                ctx.sweep(0)
                ctx.synchronize()
                ts.append(ctx.last_sweep_ms())
            phi = ctx.get_array(abi.ARR_PHI_CUR, 0)
            if ref is None:
                ref = phi
            nz = np.abs(ref) > 1e-30 * np.abs(ref).max()
            diff = float(np.max(np.abs(phi[nz] - ref[nz]) / np.abs(ref[nz])))
            t = sorted(ts[2:])
            upd = float(prob.n_cells) * prob.n_ord
            say(f"n {a.n} W {W} units {ctx.info(abi.INFO_N_UNITS)} ctas {ctx.info(abi.INFO_SWEEP_BLOCKS)} {vname:5s} -> {'ll' if ctx.info(abi.INFO_SWEEP_VARIANT) == abi.SWEEP_LL else 'flags'} "
                f"tile {ctx.info(abi.INFO_TILE_X)}x{ctx.info(abi.INFO_TILE_Y)}x{ctx.info(abi.INFO_TILE_Z)} sweep_ms min {t[0]:.3f} med {t[len(t) // 2]:.3f} "
                f"updates/s {upd / (t[0] * 1e-3):.4g} maxrel_vs_first {diff:.2e}")
    ctx.close()
