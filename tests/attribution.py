#!/usr/bin/env python
"""Where the CUDA sweep's differences from the reference come from (not a test - run by hand on the GPU box).

One sweep of a 48^3 TWDD problem through the oracle (bit-identical to the unmodified reference at one thread,
tests/test_oracle_vs_reference.py) and through builds of the product library that differ in ONE arithmetic property:

    default      MUFU.RCP64H seed + one cubic Newton step (~2 ulp), FMA contraction on
    ieee         correctly rounded reciprocals (scripts/build_exp.sh ieee -DDOMA_IEEE_DIV)
    nofma        no FMA contraction (scripts/build_exp.sh nofma -fmad=false)
    ieee_nofma   both

    for l in default ieee nofma ieee_nofma; do DOMA_B200_LIB=... python tests/attribution.py $l; done

Printed per build: max / 99.9th percentile / median of |phi_gpu - phi_ref| / |phi_ref| over the cells, for a benign source
(q = 1 everywhere) and for a source that spans 14 decades with holes (the 1e-8 switch of the theta weights is taken both
ways)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import ORACLE_SO, _build_oracle  # noqa: E402

from doma_b200 import abi  # noqa: E402
from doma_b200.synthetic import random_problem  # noqa: E402

label = sys.argv[1] if len(sys.argv) > 1 else "default"
_build_oracle()
oracle, cuda = abi.Library(ORACLE_SO, "doma_oracle_"), abi.load_library()
prob = random_problem((48, 48, 48), G=2, dims=3, scheme=abi.SCHEME_TWDD, npol=6, nazi=6, seed=7)
rng = np.random.default_rng(3)
sources = {"q=1": np.ones(prob.n_cells), "14 decades": 10.0 ** rng.uniform(-12, 2, prob.n_cells) * (rng.random(prob.n_cells) > 0.2)}
for name, q in sources.items():
    out = []
    for lib in (oracle, cuda):
        ctx = abi.Context(prob, lib)
        ctx.reset_state(0.0)
        ctx.set_array(abi.ARR_Q, 0, q)
        ctx.sweep(1)
        out.append(ctx.get_array(abi.ARR_PHI_CUR, 0))
        ctx.close()
    ref, got = out
    rel = np.abs(got - ref) / np.maximum(np.abs(ref), 1e-300)
    print(f"{label:11s} {name:11s} rel err max {rel.max():.2e} p99.9 {np.percentile(rel, 99.9):.2e} median {np.median(rel):.2e} "
          f"bit-identical cells {np.mean(got == ref) * 100:.1f} %", flush=True)
