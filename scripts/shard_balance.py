#!/usr/bin/env python
"""Sweep time of EVERY rank's ordinate shard of a W-rank run, one after the other on one GPU: is the dealing balanced?

    python scripts/shard_balance.py [--n 256] [--worlds 2,8]
"""
import argparse, os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from doma_b200 import abi
from doma_b200.sharding import shard_ordinates
from doma_b200.synthetic import synthetic_case
ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=256)
ap.add_argument("--worlds", default="2,8")
a = ap.parse_args()
params, full = synthetic_case(os.path.join(tempfile.gettempdir(), f"bal_{a.n}.xml"), a.n, G=8, mode="B")
for W in [int(w) for w in a.worlds.split(",")]:
    ts = []
    for r in range(W):
        ords = shard_ordinates(full, r, W)
        prob = full.with_ordinates(ords)
        ctx = abi.Context(prob)
        ctx.reset_state(0.0)
        ctx.update_mg_source(0, 0.0)
        t = []
        for i in range(5):
            ctx.sweep(0); ctx.synchronize(); t.append(ctx.last_sweep_ms())
        octs = sorted(set(int(o) for o in np.asarray(prob.ord_octant)))
        ts.append(min(t[1:]))
        print(f"W {W} rank {r} units {ctx.info(abi.INFO_N_UNITS)} octants {octs} tile {ctx.info(abi.INFO_TILE_X)}x{ctx.info(abi.INFO_TILE_Y)}x{ctx.info(abi.INFO_TILE_Z)} sweep_ms {min(t[1:]):.3f}", flush=True)
        ctx.close()
    print(f"W {W}: min {min(ts):.3f} max {max(ts):.3f} mean {np.mean(ts):.3f} imbalance {100 * (max(ts) / np.mean(ts) - 1):.1f} %", flush=True)
