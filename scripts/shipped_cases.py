#!/usr/bin/env python
"""Throughput of the SHIPPED cases on one GPU (SURVEY.md 8d "shipped-case inputs").

    python scripts/shipped_cases.py [--out gpurun_out/shipped_cases.json] [--full]

Runs every tests/golden/cases/full_*.xml (the shipped meshes / quadratures; iteration budgets as in
tests/golden/make_golden_full.py, so the reference's own time for the same work is in the digest)
through the Python mirror of the reference's loops with the library-owned inner iteration, and
prints sweeps, cell-angle-group updates, seconds and updates/s.  --full also runs the unmodified
shipped files from oracle/_ref/cases to convergence where present (no reference timing exists for
those: the reference needs hours on 8 vCPU), the 2D transient bounded to its first 10 time steps.
"""
import argparse
import json
import os
import re
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import doma_b200 as d  # noqa: E402
from doma_b200 import abi  # noqa: E402


def run(path, label, fused=True, ref_seconds=None, ref_threads=None, ref_si=None):
    params, prob = d.load_case(path)
    ctx = abi.Context(prob)
    s = d.TransportSolver(ctx, params, fused_si=fused)
    ctx.synchronize()
    t0 = time.perf_counter()
    ok = s.solve()
    ctx.synchronize()
    sec = time.perf_counter() - t0
    upd = float(prob.n_cells) * prob.n_ord * s.trace.sweeps
    row = {"case": label, "cells": prob.n_cells, "ordinates": prob.n_ord, "groups": params.num_groups, "sweeps": s.trace.sweeps,
           "ok": bool(ok), "k": s.k if params.mode == abi.MODE_EIGEN else None, "seconds": sec, "updates_per_s": upd / sec,
           "us_per_sweep": sec / max(s.trace.sweeps, 1) * 1e6, "units": ctx.info(abi.INFO_N_UNITS),
           "tile": "%dx%dx%d" % (ctx.info(abi.INFO_TILE_X), ctx.info(abi.INFO_TILE_Y), ctx.info(abi.INFO_TILE_Z)), "inner_loop": "fused" if s.fused_si else "host",
           "protocol": "ll" if ctx.info(abi.INFO_SWEEP_VARIANT) == abi.SWEEP_LL else "flags", "rows_per_step": ctx.info(abi.INFO_ROWS_PER_STEP),
           "si_counts_match_reference": None}
    if ref_seconds and ref_seconds == ref_seconds:
        row["reference_seconds"] = ref_seconds
        row["reference_threads"] = ref_threads
        row["reference_where"] = "build container, 8 vCPU (tests/golden/full digest)"
    if ref_si is not None:
        row["si_counts_match_reference"] = list(s.trace.si_counts) == list(ref_si)   # identical iteration trace
    print(json.dumps(row), flush=True)
    return row


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "shipped_cases.json"))
    ap.add_argument("--full", action="store_true")
    a = ap.parse_args()
    rows = []
    cases = os.path.join(ROOT, "tests", "golden", "cases")
    # untimed warm-up: module load, first-launch attribute calls and clocks must not land on the first (9-sweep) case
    params, prob = d.load_case(os.path.join(cases, "full_k1.xml"))
    d.TransportSolver(abi.Context(prob), params, fused_si=True).solve()
    for f in sorted(os.listdir(cases)):
        if not f.startswith("full_"):
            continue
        name = f[5:-4]
        g = np.load(os.path.join(ROOT, "tests", "golden", "full", name + ".npz"))
        for fused in (True, False):
            rows.append(run(os.path.join(cases, f), "full_" + name, fused, float(g["ref_seconds"]), int(g["ref_threads"]), g["si_counts"].tolist()))
    if a.full:
        ref_cases = os.path.join(ROOT, "oracle", "_ref", "cases")
        for f in ("3D_assembly_fs.xml", "3D_assembly_eig.xml", "2D_assembly_fs.xml", "2D_assembly_eig.xml"):
            p = os.path.join(ref_cases, f)
            if os.path.exists(p):
                rows.append(run(p, f + " (to convergence, unpinned)"))
        p = os.path.join(ref_cases, "2D_assembly_transient", "2D_assembly_tr.xml")
        if os.path.exists(p):
            txt = re.sub(r't1="[^"]*"', 't1="0.1"', re.sub(r't_steps="[^"]*"', 't_steps="10"', open(p).read()))
            q = os.path.join("/tmp", "2D_assembly_tr_10steps.xml")
            open(q, "w").write(txt)
            rows.append(run(q, "2D_assembly_tr.xml (first 10 of 1000 time steps, unpinned)"))
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(rows, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
