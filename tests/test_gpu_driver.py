"""The C++ drop-in on the GPU box: doma_b200/bin/doma_b200_driver is the reference's own unmodified
driver loops (compiled from the reference tree, which the box does not have, so the prebuilt binary
travels) bound to libdoma_b200.so through the explicit-specialisation seam of INTEGRATION.md."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden_case_path, load_golden

pytestmark = pytest.mark.gpu
DRIVER = os.path.join(ROOT, "doma_b200", "bin", "doma_b200_driver")


@pytest.mark.parametrize("name", ["k1_r0_s3", "k1_r0_s3_dd", "asm_eig_small_1sweep", "test2d_r0", "tr2d_r0_3steps"])
def test_driver_reproduces_reference(name, tmp_path):
    if not os.path.exists(DRIVER):
        pytest.fail("doma_b200/bin/doma_b200_driver missing: build() compiles it where /root/reference exists")
    g = load_golden(name)
    xml = tmp_path / (name + ".xml")
    xml.write_text(open(golden_case_path(name)).read())
    r = subprocess.run([DRIVER, "-i", str(xml)], capture_output=True, text=True, timeout=600)
    assert "DRIVER impl=b200" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    tail = r.stdout.rsplit("DRIVER", 1)[1]
    assert ("ok=1" in tail) == bool(g["ok"])
    assert int(re.search(r"kernel_launches=(\d+)", tail).group(1)) > 0
    si = [int(x) for x in re.findall(r"SI (?:converged|failed to converge) after (\d+) iterations", r.stdout)]
    assert si == list(g["si_counts"])
    if name.startswith("asm_eig"):
        assert abs(float(re.search(r" k=([-0-9.e+]+)", tail).group(1)) - float(g["k"])) <= 1e-9
    if bool(g["ok"]):
        # the reference's own text dump (6 significant digits) of the GPU-computed flux
        base = str(tmp_path / name)
        suffix = "_t%d" % (len(si) and 2) if name.startswith("tr2d") else ""
        for grp in range(g["phi"].shape[0]):
            f = base + suffix + f"_g{grp}_flux.txt"
            assert os.path.exists(f), f
            phi = np.loadtxt(f)
            ref = g["phi"][grp]
            big = np.abs(ref) > 1e-30
            assert np.max(np.abs(phi[big] - ref[big]) / np.abs(ref[big])) < 2e-5


@pytest.mark.parametrize("name", ["k1_r0_s3", "test2d_r0"])
def test_driver_fast_dump_writes_the_reference_files(name, tmp_path):
    """--fast-dump: the final dump comes from doma_b200_dump_text + the flattened geometry instead of
    BrickMesh::dumpToTextFile; the two runs of the same binary must leave byte-identical files, and for the case
    with a reference text fixture (tests/golden/text) the reference's own bytes."""
    import tarfile
    outs = []
    for flags in ([], ["--fast-dump"]):
        d = tmp_path / ("fast" if flags else "plain")
        d.mkdir()
        xml = d / (name + ".xml")
        xml.write_text(open(golden_case_path(name)).read())
        r = subprocess.run([DRIVER, "-i", str(xml)] + flags, capture_output=True, text=True, timeout=600)
        assert "ok=1" in r.stdout.rsplit("DRIVER", 1)[1], r.stdout[-2000:] + r.stderr[-2000:]
        outs.append({f[len(name):]: open(d / f, "rb").read() for f in os.listdir(d) if f.endswith(".txt")})
    plain, fast = outs
    assert sorted(plain) == sorted(fast) and len(fast) >= 5
    assert not [n for n in plain if plain[n] != fast[n]]
    fixture = os.path.join(ROOT, "tests", "golden", "text", name + ".tar.gz")
    if os.path.exists(fixture):
        with tarfile.open(fixture) as tar:
            ref = {m.name[len("ref"):]: tar.extractfile(m).read() for m in tar.getmembers()}
        assert sorted(ref) == sorted(fast)
        assert not [n for n in ref if ref[n] != fast[n]]
