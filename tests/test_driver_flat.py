"""The flat mesh builder of the C++ drop-in (doma_b200/driver/flatten.h setGeometryFromParams) against the
mesh the reference's own BrickMesh2D/3D constructor builds: per-cell widths and block ids must be identical bit for
bit.  `doma_b200_driver --check-flat` does the comparison inside the binary (it has both) and needs no GPU."""
import os
import subprocess

import pytest

from conftest import GOLDEN, ROOT

DRIVER = os.path.join(ROOT, "doma_b200", "bin", "doma_b200_driver")
SHIPPED = os.path.join(ROOT, "oracle", "_ref", "cases")
CASES = sorted(os.path.join(GOLDEN, "cases", f) for f in os.listdir(os.path.join(GOLDEN, "cases")) if f.endswith(".xml") and "syn256" not in f)
if os.path.isdir(SHIPPED):      # the unchanged shipped inputs (copied by `make -C oracle ref`; they travel to the GPU box)
    for root, _, files in os.walk(SHIPPED):
        CASES += sorted(os.path.join(root, f) for f in files if f.endswith(".xml"))


@pytest.mark.parametrize("case", CASES, ids=[os.path.relpath(c, ROOT) for c in CASES])
def test_flat_builder_equals_reference_mesh(case):
    if not os.path.exists(DRIVER):
        pytest.skip("doma_b200/bin/doma_b200_driver is built only where the reference tree exists (build())")
    r = subprocess.run([DRIVER, "-i", case, "--check-flat", "--no-output"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "FLATCHECK identical" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]
