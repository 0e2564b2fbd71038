#!/usr/bin/env python
"""Error statistics of the GPU path against a full-size reference digest (tests/golden/full/<name>.npz):
where the differences are, relative to the local flux and relative to the group's maximum."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import doma_b200 as d
from doma_b200 import abi
name = sys.argv[1] if len(sys.argv) > 1 else "syn256_b"
g = np.load(os.path.join(ROOT, "tests", "golden", "full", name + ".npz"))
params, prob = d.load_case(os.path.join(ROOT, "tests", "golden", "cases", "full_" + name + ".xml"))
ctx = abi.Context(prob)
s = d.TransportSolver(ctx, params, fused_si=True)
s.solve()
phi = s.flux(); idx = g["idx"]; ref = g["phi_s"]; got = phi[:, idx]
rel = np.abs(got - ref) / np.maximum(np.abs(ref), 1e-300)
sc = np.abs(got - ref) / g["phi_max"][:, None]
for gi in range(ref.shape[0]):
    r = rel[gi]; j = int(np.argmax(r))
    print(f"g{gi}: rel err median {np.median(r):.2e} p99 {np.percentile(r, 99):.2e} max {r.max():.2e} at |phi|={abs(ref[gi, j]):.3e} (phi_max {g['phi_max'][gi]:.3e}); "
          f"abs err / phi_max: max {sc[gi].max():.2e}; cells with rel > 1e-10: {(r > 1e-10).sum()} of {r.size}, their max |phi|/phi_max {np.abs(ref[gi][r > 1e-10]).max() / g['phi_max'][gi] if (r > 1e-10).any() else 0:.2e}")
print("sum rel", np.abs(phi.sum(axis=1) / g["phi_sum"] - 1).max(), "l2 rel", np.abs(np.sqrt((phi * phi).sum(axis=1)) / g["phi_l2"] - 1).max())
nx, ny, nz = g["shape"].tolist()
slabs = phi.reshape(phi.shape[0], nz, ny * nx).sum(axis=2)
print("slab sums rel max", np.abs(slabs / g["slab_sum"] - 1).max(), "min", phi.min(axis=1), g["phi_min"])
