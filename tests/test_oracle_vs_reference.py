"""Pins the CPU oracle (oracle/doma_oracle.c) and the host-side input mirror against the UNMODIFIED
reference: the golden .npz files were produced by oracle/_ref/ref_harness (tests/golden/make_golden.py).

At one thread the oracle follows the reference's operation order exactly, so fluxes, residuals and k
are required to be BIT-IDENTICAL, and the source-iteration traces identical.
"""
import os

import numpy as np
import pytest

from conftest import golden_case_path, load_golden

import doma_b200 as d
from doma_b200 import abi

# (name, seconds the oracle needs on one core)
FAST = ["k1_r0_s3", "k1_r0_s3_dd", "k3_r0_s2", "asm_eig_small_1sweep", "test2d_r0", "test2d_r0_dd", "tr2d_r0_3steps"]
SLOW = ["asm_fs_small", "asm_eig_small", "tr2d_r0_remove", "asm2d_eig_small"]


def run_case(name, lib, threads=1, sharder=None):
    params, prob = d.load_case(golden_case_path(name))
    ctx = abi.Context(prob, lib)
    if ctx.has("set_threads"):
        ctx._call("set_threads", threads)
    s = d.TransportSolver(ctx, params)
    ok = s.solve()
    dnp = np.stack([ctx.get_array(abi.ARR_DNP_CUR, i) for i in range(prob.n_delayed_max)]) if (
        prob.mode == abi.MODE_TRANSIENT and prob.n_delayed_max) else np.zeros((0, prob.n_cells))
    return params, prob, s, ok, s.flux(), dnp


@pytest.mark.parametrize("name", FAST + SLOW)
def test_input_mirror_matches_reference_objects(name):
    """XML -> mesh widths, block ids, quadrature tables: bit-identical to what the reference built."""
    g = load_golden(name)
    params, prob = d.load_case(golden_case_path(name))
    nx, ny, nz = [int(v) for v in g["shape"]]
    assert (prob.nx, prob.ny, prob.nz) == (nx, ny, nz)
    lx = g["mesh_lx"].reshape(nz, ny, nx)
    ly = g["mesh_ly"].reshape(nz, ny, nx)
    assert np.array_equal(lx[0, 0, :], prob.dx) and np.array_equal(ly[0, :, 0], prob.dy)
    if prob.dims == 3:
        assert np.array_equal(g["mesh_lz"].reshape(nz, ny, nx)[:, 0, 0], prob.dz)
    assert np.array_equal(prob.block_ids[prob.mat_of_cell], g["mesh_block"])
    assert np.array_equal(prob.ord_octant, g["quad_oct"])
    for n, a in (("mu", prob.mu), ("eta", prob.eta), ("xi", prob.xi), ("w", prob.w)):
        assert np.array_equal(a, g["quad_" + n]), n
    total = 4.0 * np.pi if prob.dims == 3 else 2.0 * np.pi
    assert abs(prob.w.sum() - total) < 1e-13


def _check(name, oracle_lib):
    g = load_golden(name)
    params, prob, s, ok, phi, dnp = run_case(name, oracle_lib)
    assert ok == bool(g["ok"])
    assert list(s.trace.si_counts) == list(g["si_counts"]), "source-iteration trace differs from the reference"
    assert np.array_equal(np.array(s.trace.si_residuals), g["si_residuals"])
    assert np.array_equal(phi, g["phi"]), f"max rel diff {np.max(np.abs(phi - g['phi']) / np.maximum(np.abs(g['phi']), 1e-300))}"
    assert np.array_equal(ctx_q(s), g["q"])
    if params.mode == abi.MODE_EIGEN:
        assert s.k == float(g["k"]) and s.k_prev == float(g["k_prev"])
        assert np.array_equal(np.array(s.trace.k_history)[: len(g["k_history"])], g["k_history"][: len(s.trace.k_history)])
    if dnp.size:
        assert np.array_equal(dnp, g["dnp"])


def ctx_q(s):
    return s.ctx.get_array(abi.ARR_Q, 0)


@pytest.mark.parametrize("name", FAST)
def test_oracle_bit_identical_to_reference(name, oracle_lib):
    _check(name, oracle_lib)


@pytest.mark.slow
@pytest.mark.skipif(not os.environ.get("DOMA_SLOW"), reason="minutes of CPU; set DOMA_SLOW=1")
@pytest.mark.parametrize("name", SLOW)
def test_oracle_bit_identical_to_reference_slow(name, oracle_lib):
    _check(name, oracle_lib)


def test_oracle_thread_count_only_reorders_sums(oracle_lib):
    """The reference depends on the OpenMP thread count only through summation order (<= 1e-14)."""
    g = load_golden("k1_r0_s3")
    _, _, s, ok, phi, _ = run_case("k1_r0_s3", oracle_lib, threads=4)
    assert ok and list(s.trace.si_counts) == list(g["si_counts"])
    assert np.max(np.abs(phi - g["phi"]) / np.abs(g["phi"])) < 1e-13
