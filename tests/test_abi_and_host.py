"""CPU-side checks: the C-ABI library loads and exports every symbol include/doma_b200.h declares
(no compute calls without a GPU), the product refuses to run without a device, and the host-side
mirror of the reference's input API / outer loops behaves like the reference."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, golden_case_path, load_golden

import doma_b200 as d
from doma_b200 import abi
from doma_b200.sharding import shard_ordinates, shard_units
from doma_b200.synthetic import random_problem, synthetic_case, synthetic_xml

HEADER = os.path.join(ROOT, "include", "doma_b200.h")


def declared_symbols():
    txt = open(HEADER).read()
    return sorted(set(re.findall(r"\b(doma_b200_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    syms = declared_symbols()
    assert len(syms) >= 25 and "doma_b200_sweep" in syms and "doma_b200_create" in syms
    assert os.path.exists(abi.LIB_PATH), "libdoma_b200.so missing: run __graft_entry__.build()"
    dll = ctypes.CDLL(abi.LIB_PATH)
    missing = [s for s in syms if not hasattr(dll, s)]
    assert not missing, f"declared in the header but not exported: {missing}"


def test_binding_covers_the_header():
    bound = {"doma_b200_" + n for n in list(abi.Library._SIGS) + list(abi.Library._OPTIONAL)} | {"doma_b200_last_error"}
    assert bound - {"doma_b200_set_threads"} <= set(declared_symbols()) | {"doma_b200_set_threads"}
    assert set(declared_symbols()) <= bound, set(declared_symbols()) - bound


def test_no_cpu_fallback():
    """Without a CUDA device create must fail loudly; with one this test is a no-op."""
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    prob = random_problem((4, 4, 4))
    with pytest.raises(abi.DomaError):
        abi.Context(prob, abi.load_library())


def test_product_package_never_references_the_oracle():
    pkg = os.path.join(ROOT, "doma_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cpp")):
                txt = open(os.path.join(dp, f)).read()
                assert "libdoma_oracle" not in txt and "doma_oracle_create" not in txt, f


def test_input_errors_match_reference_behaviour(tmp_path):
    good = open(golden_case_path("k1_r0_s3")).read()
    cases = {
        "scheme": good.replace('scheme="tw_dd"', 'scheme="sc"'),
        "solve_type": good.replace('solve_type="steady"', 'solve_type="fast"'),
        "groups": good.replace('num_groups="1"', 'num_groups="2"'),
        "negative": good.replace('fs_tol="1e-3"', 'fs_tol="-1"'),
        "blocks": good.replace('<ix vals="5,', '<ix vals="5, 1,'),
        "nomat": good.replace('<mat block="2">', '<mat block="7">'),
    }
    for name, xml in cases.items():
        p = tmp_path / f"{name}.xml"
        p.write_text(xml)
        with pytest.raises(d.InputError):
            d.load_case(str(p))


def test_refinement_and_transpose(tmp_path):
    params, prob = d.load_case(golden_case_path("tr2d_r0_3steps"))
    assert prob.dims == 2 and prob.n_ord == 100 and prob.n_delayed_max == 6
    assert abs(params.dt - 0.01) < 1e-15 and prob.mode == abi.MODE_TRANSIENT
    m = int(np.nonzero(prob.block_ids == 1)[0][0])
    # transpose="true": XML row = from-group; stored [to, from] (src/InputParameters.cpp:301-318)
    assert prob.sigma_s[m, 1, 0] == 0.005663016987588101 and prob.sigma_s[m, 0, 1] == 0.0
    assert prob.step_type[int(np.nonzero(prob.block_ids == 2)[0][0])] == abi.STEP_BOTH
    xml = open(golden_case_path("k1_r0_s3")).read().replace('refinement="0"', 'refinement="2"')
    p = tmp_path / "r2.xml"
    p.write_text(xml)
    _, prob2 = d.load_case(str(p))
    assert (prob2.nx, prob2.ny, prob2.nz) == (60, 60, 60) and abs(prob2.dx.sum() - 200.0) < 1e-12


def test_sharding_partitions_the_quadrature():
    _, prob = synthetic_case("/tmp/doma_syn_t.xml", 8, G=2)
    assert prob.n_ord == 512 and len(shard_units(prob)) == 16
    for world in (1, 2, 4, 8, 3):
        parts = [shard_ordinates(prob, r, world) for r in range(world)]
        allidx = np.sort(np.concatenate(parts))
        assert np.array_equal(allidx, np.arange(512))
        if world in (2, 4, 8):
            assert {len(p) for p in parts} == {512 // world}
            # lanes stay full: every shard is made of whole 32-ordinate units of single octants
            for p in parts:
                for o in np.unique(prob.ord_octant[p]):
                    assert np.count_nonzero(prob.ord_octant[p] == o) % 32 == 0


def test_fixed_work_modes_have_deterministic_sweep_counts(oracle_lib, tmp_path):
    """Mode A = exactly K sweeps of group 0, Mode B = one sweep per group."""
    for mode, K, expect in (("A", 3, [3]), ("B", 1, [1, 1])):
        p = tmp_path / f"m{mode}.xml"
        p.write_text(synthetic_xml(8, G=2, npol=1, nazi=1, mode=mode, sweeps=K))
        params, prob = d.load_case(str(p))
        s = d.TransportSolver(abi.Context(prob, oracle_lib), params)
        ok = s.solve()
        assert list(s.trace.si_counts) == expect and s.trace.sweeps == sum(expect)
        assert ok == (mode == "B")


def test_solver_loop_matches_reference_trace_strings(oracle_lib):
    g = load_golden("asm_eig_small_1sweep")
    params, prob = d.load_case(golden_case_path("asm_eig_small_1sweep"))
    lines = []
    s = d.TransportSolver(abi.Context(prob, oracle_lib), params, log=lines.append)
    assert s.solve() is False          # mg_max_it reached -> "PI failed to converge"
    assert any(l.startswith("PI failed to converge after 3 iterations") for l in lines)
    assert np.array_equal(np.array(s.trace.k_history), np.append(g["k_history"], s.k)[: len(s.trace.k_history)])


def test_create_rejects_missing_pointers_with_a_message():
    """Every pointer the create path dereferences is validated first: rc != 0 and a message naming the field, not a fault
    across the C ABI (no GPU needed: validation precedes device selection)."""
    import ctypes as C

    lib = abi.load_library()
    for field, word in [("dz", "dz"), ("mu", "ord_octant"), ("nu_sigma_f", "nu_sigma_f"), ("mat_flags", "mat_flags"), ("step_src", "step_src")]:
        prob = random_problem((4, 3, 2))
        cp = prob.as_c()
        setattr(cp, field, C.cast(None, type(getattr(cp, field))))
        h = C.c_void_p()
        rc = lib.fn["create"](C.byref(cp), C.byref(h))
        assert rc == 1 and not h.value
        assert word in lib.last_error(None), lib.last_error(None)
    cp = random_problem((4, 3, 2)).as_c()
    cp.scheme = 7
    assert lib.fn["create"](C.byref(cp), C.byref(C.c_void_p())) == 1 and "scheme" in lib.last_error(None)
