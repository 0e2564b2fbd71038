"""The output path: text dumps in the reference's on-disk format.

Fixtures: tests/golden/text/<case>.tar.gz, written by the UNMODIFIED reference's BrickMesh::dumpToTextFile /
dumpDNPsToTextFile (tests/golden/make_golden_text.py).  CPU tests: the library's formatter (pure host code) and the
geometry writer reproduce the reference's bytes from the reference's full-precision state.  GPU tests: a solve with
output_base set writes byte-identical files at every dump point of the reference's loops.
"""
import os
import tarfile

import numpy as np
import pytest

from conftest import GOLDEN, golden_case_path, load_golden

import doma_b200 as d
from doma_b200 import abi
from doma_b200.output import cell_centres, dump_geometry


def ref_files(case):
    out = {}
    with tarfile.open(os.path.join(GOLDEN, "text", case + ".tar.gz")) as tar:
        for m in tar.getmembers():
            out[m.name[len("ref"):]] = tar.extractfile(m).read()
    return out


def read(path):
    with open(path, "rb") as f:
        return f.read()


def test_formatter_reproduces_reference_flux_bytes(tmp_path):
    """%.6g of the reference's own full-precision flux == the reference's text file (no GPU involved)."""
    ref = ref_files("k1_r0_s3")
    g = load_golden("k1_r0_s3")
    phi = np.asarray(g["phi"], dtype=np.float64).reshape(-1)
    abi.write_values_text(phi, str(tmp_path / "f.txt"), threads=3)
    assert read(tmp_path / "f.txt") == ref["_g0_flux.txt"]
    abi.write_values_text(phi, str(tmp_path / "f1.txt"), threads=1)
    assert read(tmp_path / "f1.txt") == ref["_g0_flux.txt"]
    edge = np.array([0.0, -0.0, 1e-300, -3.5e22, np.inf, -np.inf, np.nan, 999999.5, 123456.5, 1e-5, 0.1])
    abi.write_values_text(edge, str(tmp_path / "e.txt"))
    assert read(tmp_path / "e.txt") == "".join("%.6g\n" % v for v in edge).encode()
    abi.write_values_text(np.zeros(0), str(tmp_path / "z.txt"))
    assert read(tmp_path / "z.txt") == b""


@pytest.mark.parametrize("case", ["k1_r0_s3", "tr2d_r0_3steps"])
def test_geometry_files_match_reference(case, tmp_path):
    ref = ref_files(case)
    params, prob = d.load_case(golden_case_path(case))
    dump_geometry(prob, str(tmp_path / "o"))
    for name in ["_dims.txt", "_blocks.txt", "_meshx.txt", "_meshy.txt"] + (["_meshz.txt"] if prob.dims == 3 else []):
        assert read(str(tmp_path / "o") + name) == ref[name], name
    assert "_meshz.txt" in ref or prob.dims == 2
    np.testing.assert_array_equal(cell_centres(np.array([2.0, 1.0, 1.0])), [1.0, 3.0, 4.0])   # the reference's running centre


@pytest.mark.gpu
@pytest.mark.parametrize("fused_si", [False, True], ids=["loop_on_host", "fused_si"])
@pytest.mark.parametrize("case", ["k1_r0_s3", "tr2d_r0_3steps"])
def test_solve_writes_reference_files(case, fused_si, cuda_lib, tmp_path):
    ref = ref_files(case)
    params, prob = d.load_case(golden_case_path(case))
    ctx = abi.Context(prob, cuda_lib)
    base = str(tmp_path / "run")
    s = d.TransportSolver(ctx, params, fused_si=fused_si, output_base=base, problem=prob)
    assert s.solve()
    written = {f[len("run"):]: read(tmp_path / f) for f in os.listdir(tmp_path)}
    assert sorted(written) == sorted(ref), "a different set of files than the reference writes"
    bad = [n for n in ref if written[n] != ref[n]]
    assert not bad, f"bytes differ from the reference's dump: {bad}"
