"""doma_b200_scatter_residual_noalloc (no counterpart in the reference):
the scattering residual with its reduction scratch in a file-scope __device__ array instead of cudaMalloc'ed buffers.

Checked through the C ABI against the existing entry point doma_b200_scatter_residual on the same device state, on two
calls in a row (different problems of different sizes, 3D then 2D, so the second call finds the first call's partial
sums in the shared scratch), within the 1e-12 relative bound the other residual tests use.  The file sorts last
(test_zz_*) so that a failure here cannot hide another test under `pytest -x`; it builds its own contexts and passes
alone in a fresh process as well as after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from doma_b200 import abi
from doma_b200.synthetic import random_problem

pytestmark = pytest.mark.gpu

RES_RTOL = 1e-12          # tests/test_gpu_parity.py: "residuals to 1e-12"
MAX_CELLS = 33554432      # DOMA_NOALLOC_MAX_CELLS (include/doma_b200.h)


def _noalloc(lib, phi_tot_ptr, phi_cur_ptr, dx, dy, dz, dims, shape, sums_dev, stream=0, want_host=True):
    res = C.c_double(-1.0)
    err = C.create_string_buffer(512)
    rc = lib.fn["scatter_residual_noalloc"](phi_tot_ptr, phi_cur_ptr, dx, dy, dz, dims, shape[0], shape[1], shape[2], stream, sums_dev,
                                            C.byref(res) if want_host else None, err, len(err))
    return rc, res.value, err.value.decode()


def _state(lib, shape, dims, seed):
    """A context after one sweep on a random problem with a random phi_tot: (ctx, problem, residual by the existing entry point)."""
    mode = abi.MODE_TRANSIENT if dims == 2 else abi.MODE_FIXED
    prob = random_problem(shape, G=2, dims=dims, scheme=1, npol=3, nazi=3, seed=seed, mode=mode, dt=0.01)
    rng = np.random.default_rng(seed)
    ctx = abi.Context(prob, lib)
    ctx.reset_state(0.0)
    ctx.set_array(abi.ARR_PHI_TOT, 1, 10.0 ** rng.uniform(-6, 1, prob.n_cells))
    ctx.set_array(abi.ARR_Q, 0, rng.uniform(0.0, 1.0, prob.n_cells))
    ctx.sweep(1)
    ref = ctx.scatter_residual(1)       # synchronises; leaves phi_cur / phi_tot untouched
    assert np.isfinite(ref) and ref > 0.0
    return ctx, prob, ref


def test_noalloc_residual_matches_existing_entry_point_twice_in_a_row(cuda_lib):
    import torch

    assert "scatter_residual_noalloc" in cuda_lib.fn, "libdoma_b200.so does not export doma_b200_scatter_residual_noalloc"
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    sums = torch.full((2,), -7.0, dtype=torch.float64, device=dev)      # the caller's output buffer (test-side allocation only)
    seen = []
    # call 1: 3D, 23 blocks of partial sums; call 2: 2D, 1 block - it must not see what call 1 left in the shared scratch
    for shape, dims, seed in (((61, 37, 41), 3, 5), ((45, 38, 1), 2, 9)):
        ctx, prob, ref = _state(cuda_lib, shape, dims, seed)
        w = [torch.tensor(np.asarray(a, dtype=np.float64), device=dev) for a in (prob.dx, prob.dy, prob.dz if dims == 3 else np.ones(1))]
        torch.cuda.synchronize()
        ptrs = [t.data_ptr() for t in w]
        phi_tot, phi_cur = ctx.device_ptr(abi.ARR_PHI_TOT, 1), ctx.device_ptr(abi.ARR_PHI_CUR, 0)
        rc, res, msg = _noalloc(cuda_lib, phi_tot, phi_cur, ptrs[0], ptrs[1], ptrs[2] if dims == 3 else None, dims, shape, sums.data_ptr())
        assert rc == 0, msg
        assert abs(res - ref) <= RES_RTOL * abs(ref), (res, ref)
        # the device-side pair gives the same residual, and the existing entry point still agrees afterwards
        d, t = (float(v) for v in sums.cpu())
        assert t > 1e-8 and abs(np.sqrt(d) / np.sqrt(t) - ref) <= RES_RTOL * abs(ref)
        # the same call again, result left on the device only (no host synchronisation inside the call): identical bits
        sums.fill_(-7.0)
        rc, _, msg = _noalloc(cuda_lib, phi_tot, phi_cur, ptrs[0], ptrs[1], ptrs[2] if dims == 3 else None, dims, shape, sums.data_ptr(),
                              want_host=False)
        assert rc == 0, msg
        torch.cuda.synchronize()
        assert (float(sums[0]), float(sums[1])) == (d, t)
        assert ctx.scatter_residual(1) == ref
        seen.append(res)
        ctx.close()
    assert seen[0] != seen[1]


def test_noalloc_rejects_what_the_static_scratch_cannot_hold(cuda_lib):
    import torch

    torch.cuda.set_device(0)
    buf = torch.zeros(8, dtype=torch.float64, device="cuda")
    p = buf.data_ptr()
    rc, _, msg = _noalloc(cuda_lib, p, p, p, p, p, 3, (512, 512, 129), p)      # 33 816 576 cells > 33 554 432
    assert rc == 40 and "exceed" in msg and str(MAX_CELLS) in msg
    assert 512 * 512 * 129 > MAX_CELLS >= 512 * 512 * 128
    rc, _, msg = _noalloc(cuda_lib, p, p, p, p, p, 4, (2, 2, 2), p)
    assert rc == 1 and "dims" in msg
    rc, _, msg = _noalloc(cuda_lib, p, p, p, p, None, 3, (2, 2, 2), p)
    assert rc == 1 and "null" in msg
    rc, _, msg = _noalloc(cuda_lib, p, p, p, p, p, 3, (2, 0, 2), p)
    assert rc == 1 and "extent" in msg
