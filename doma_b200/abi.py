"""ctypes binding of the C ABI declared in include/doma_b200.h.

The product binds exactly one library, ``doma_b200/libdoma_b200.so`` (hand-written sm_100a CUDA,
built by ``__graft_entry__.build()``), through :func:`load_library`.  There is NO CPU fallback:
if that library is missing, or no CUDA device is usable, loading / context creation raises.

The binding itself is generic in (path, symbol prefix) so that the test-suite can drive its CPU
checker through the very same host code; nothing in this package ever passes another path.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DOMA_B200_LIB") or os.path.join(_HERE, "libdoma_b200.so")  # the override is for kernel experiments only

# enums of include/doma_b200.h (values are the reference's, include/TransportBase.h:20-62)
OCT_PPP, OCT_PPM, OCT_PMP, OCT_PMM, OCT_MPP, OCT_MPM, OCT_MMP, OCT_MMM = range(8)
MODE_FIXED, MODE_EIGEN, MODE_TRANSIENT = 0, 1, 2
SCHEME_DD, SCHEME_TWDD = 0, 1
STEP_NONE, STEP_INSERT, STEP_REMOVE, STEP_BOTH = -1, 0, 1, 2
MAT_HAS_SRC, MAT_HAS_CHI_P, MAT_HAS_NU_SIGMA_F = 1, 2, 4
ARR_PHI_TOT, ARR_PHI_PREV, ARR_PHI_LAST, ARR_DNP_CUR, ARR_DNP_LAST, ARR_Q, ARR_PHI_CUR = range(7)
SWEEP_AUTO, SWEEP_FLAGS, SWEEP_LL = 0, 1, 2
(INFO_KERNEL_LAUNCHES, INFO_SWEEP_VARIANT, INFO_N_CELLS, INFO_N_ORD, INFO_FACE_BYTES, INFO_DEVICE_BYTES, INFO_N_UNITS, INFO_N_TILES,
 INFO_TILE_Y, INFO_TILE_Z, INFO_CELL_LO, INFO_CELL_HI, INFO_ROWS_PER_STEP, INFO_SWEEP_BLOCKS, INFO_TILE_X) = range(15)

_pd = C.POINTER(C.c_double)
_pi = C.POINTER(C.c_int32)


class CProblem(C.Structure):
    """``struct doma_b200_problem`` (include/doma_b200.h)."""

    _fields_ = [
        ("dims", C.c_int32), ("nx", C.c_int32), ("ny", C.c_int32), ("nz", C.c_int32),
        ("dx", _pd), ("dy", _pd), ("dz", _pd), ("mat_of_cell", _pi),
        ("n_groups", C.c_int32), ("n_mat", C.c_int32), ("n_delayed_max", C.c_int32),
        ("mat_flags", _pi), ("n_delayed", _pi),
        ("sigma_t", _pd), ("sigma_s", _pd), ("nu_sigma_f", _pd), ("chi_p", _pd), ("src", _pd), ("inv_v", _pd),
        ("chi_d", _pd), ("beta", _pd), ("lambda_", _pd),
        ("step_type", _pi), ("step_t_insert", _pd), ("step_t_remove", _pd), ("step_src", _pd),
        ("n_ord", C.c_int32), ("ord_octant", _pi), ("mu", _pd), ("eta", _pd), ("xi", _pd), ("w", _pd),
        ("scheme", C.c_int32), ("mode", C.c_int32), ("dt", C.c_double), ("device", C.c_int32),
    ]


@dataclass
class Problem:
    """Flat problem description; the numpy twin of ``doma_b200_problem``.

    Cell order is the reference's (x fastest, src/BrickMesh3D.cpp:86-102).  Ordinates are listed
    octant-major in the reference's generation order (src/GCAngularQuadrature.cpp:46-54) with the
    signs GCAngularQuadrature::direction returns.
    """

    dims: int
    dx: np.ndarray
    dy: np.ndarray
    dz: np.ndarray
    mat_of_cell: np.ndarray          # int32 [nz*ny*nx]
    n_groups: int
    mat_flags: np.ndarray            # int32 [M]
    n_delayed: np.ndarray            # int32 [M]
    sigma_t: np.ndarray              # [M,G]
    sigma_s: np.ndarray              # [M,G,G]  [to, from]
    nu_sigma_f: np.ndarray           # [M,G]
    chi_p: np.ndarray                # [M,G]
    src: np.ndarray                  # [M,G]
    inv_v: np.ndarray                # [M,G]
    chi_d: np.ndarray                # [M,G,D]
    beta: np.ndarray                 # [M,G,D]
    lambda_: np.ndarray              # [M,D]
    step_type: np.ndarray            # int32 [M]
    step_t_insert: np.ndarray        # [M]
    step_t_remove: np.ndarray        # [M]
    step_src: np.ndarray             # [M,G]
    ord_octant: np.ndarray           # int32 [n_ord]
    mu: np.ndarray
    eta: np.ndarray
    xi: np.ndarray
    w: np.ndarray
    scheme: int = SCHEME_TWDD
    mode: int = MODE_FIXED
    dt: float = 0.0
    device: int = 0
    block_ids: Optional[np.ndarray] = None   # material index -> XML block id (informational)
    _keep: Dict[str, np.ndarray] = field(default_factory=dict, repr=False)

    @property
    def nx(self) -> int:
        return int(self.dx.size)

    @property
    def ny(self) -> int:
        return int(self.dy.size)

    @property
    def nz(self) -> int:
        return int(self.dz.size) if self.dims == 3 else 1

    @property
    def n_cells(self) -> int:
        return self.nx * self.ny * self.nz

    @property
    def n_mat(self) -> int:
        return int(self.sigma_t.shape[0])

    @property
    def n_delayed_max(self) -> int:
        return int(self.lambda_.shape[1]) if self.lambda_.ndim == 2 else 0

    @property
    def n_ord(self) -> int:
        return int(self.mu.size)

    def with_ordinates(self, keep: np.ndarray) -> "Problem":
        """Copy of this problem restricted to the ordinates selected by index array / mask ``keep``."""
        import dataclasses

        return dataclasses.replace(
            self, ord_octant=self.ord_octant[keep].copy(), mu=self.mu[keep].copy(), eta=self.eta[keep].copy(),
            xi=self.xi[keep].copy(), w=self.w[keep].copy(), _keep={})

    def as_c(self) -> CProblem:
        """Build the C struct; the converted arrays are kept alive on ``self``."""
        G, M, D = self.n_groups, self.n_mat, self.n_delayed_max
        k = self._keep

        def f64(name, a, shape):
            a = np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(shape))
            k[name] = a
            return a.ctypes.data_as(_pd)

        def i32(name, a, shape):
            a = np.ascontiguousarray(np.asarray(a, dtype=np.int32).reshape(shape))
            k[name] = a
            return a.ctypes.data_as(_pi)

        p = CProblem()
        p.dims, p.nx, p.ny, p.nz = self.dims, self.nx, self.ny, self.nz
        p.dx, p.dy = f64("dx", self.dx, -1), f64("dy", self.dy, -1)
        p.dz = f64("dz", self.dz if self.dims == 3 else np.ones(1), -1)
        p.mat_of_cell = i32("mat_of_cell", self.mat_of_cell, self.n_cells)
        p.n_groups, p.n_mat, p.n_delayed_max = G, M, D
        p.mat_flags, p.n_delayed = i32("mat_flags", self.mat_flags, M), i32("n_delayed", self.n_delayed, M)
        p.sigma_t, p.sigma_s = f64("sigma_t", self.sigma_t, M * G), f64("sigma_s", self.sigma_s, M * G * G)
        p.nu_sigma_f, p.chi_p = f64("nu_sigma_f", self.nu_sigma_f, M * G), f64("chi_p", self.chi_p, M * G)
        p.src, p.inv_v = f64("src", self.src, M * G), f64("inv_v", self.inv_v, M * G)
        p.chi_d = f64("chi_d", self.chi_d if D else np.zeros(1), -1)
        p.beta = f64("beta", self.beta if D else np.zeros(1), -1)
        p.lambda_ = f64("lambda", self.lambda_ if D else np.zeros(1), -1)
        p.step_type = i32("step_type", self.step_type, M)
        p.step_t_insert, p.step_t_remove = f64("sti", self.step_t_insert, M), f64("str", self.step_t_remove, M)
        p.step_src = f64("step_src", self.step_src, M * G)
        p.n_ord = self.n_ord
        p.ord_octant = i32("ord_octant", self.ord_octant, -1)
        p.mu, p.eta = f64("mu", self.mu, -1), f64("eta", self.eta, -1)
        p.xi, p.w = f64("xi", self.xi, -1), f64("w", self.w, -1)
        p.scheme, p.mode, p.dt, p.device = self.scheme, self.mode, float(self.dt), self.device
        return p


class DomaError(RuntimeError):
    pass


class Library:
    """One loaded implementation of the C ABI (symbol ``<prefix><name>``)."""

    _SIGS = {
        "create": [C.POINTER(CProblem), C.POINTER(C.c_void_p)],
        "destroy": [C.c_void_p],
        "set_stream": [C.c_void_p, C.c_void_p],
        "synchronize": [C.c_void_p],
        "reset_state": [C.c_void_p, C.c_double],
        "init_zero_ic": [C.c_void_p],
        "update_step_fluxes": [C.c_void_p],
        "step_dnps": [C.c_void_p],
        "update_mg_source": [C.c_void_p, C.c_int32, C.c_double],
        "update_mg_source_eigen": [C.c_void_p, C.c_int32, C.c_double],
        "update_scatter_source": [C.c_void_p, C.c_int32],
        "sweep": [C.c_void_p, C.c_int32],
        "scatter_residual": [C.c_void_p, C.c_int32, _pd],
        "mg_residual": [C.c_void_p, _pd],
        "update_k": [C.c_void_p, _pd, _pd],
        "get_array": [C.c_void_p, C.c_int32, C.c_int32, _pd],
        "set_array": [C.c_void_p, C.c_int32, C.c_int32, _pd],
        "get_info": [C.c_void_p, C.c_int32, C.POINTER(C.c_int64)],
    }
    # entry points only the CUDA library has (device pointers, variants, timing, fused SI, NCCL)
    _OPTIONAL = {
        "device_ptr": [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)],
        "set_sweep_variant": [C.c_void_p, C.c_int32],
        "set_pairing": [C.c_void_p, C.c_int32],
        "set_rows_per_step": [C.c_void_p, C.c_int32],
        "set_check": [C.c_void_p, C.c_int32],
        "last_sweep_ms": [C.c_void_p, C.POINTER(C.c_float)],
        "set_threads": [C.c_void_p, C.c_int32],
        "sweep_from_host": [C.c_void_p, C.c_int32, _pd, _pd],
        "stage_q_from_host": [C.c_void_p, _pd],
        "sweep_from_host_async": [C.c_void_p, C.c_int32, _pd, _pd],
        "host_wait": [C.c_void_p],
        "dump_text": [C.c_void_p, C.c_char_p, C.c_int32, C.c_int32],
        "dump_text_async": [C.c_void_p, C.c_char_p, C.c_int32, C.c_int32],
        "dump_wait": [C.c_void_p],
        "write_values_text": [_pd, C.c_int64, C.c_char_p, C.c_int32],
        "comm_unique_id": [C.c_char_p, C.c_int32],
        "comm_init": [C.c_void_p, C.c_int32, C.c_int32, C.c_char_p, C.c_int32],
        "allreduce_phi_cur": [C.c_void_p],
        "set_cell_sharding": [C.c_void_p, C.c_int32],
        "set_tile": [C.c_void_p, C.c_int32, C.c_int32, C.c_int32],
        "source_iteration": [C.c_void_p, C.c_int32, C.c_double, C.c_int32, C.POINTER(C.c_int32), _pd, _pd],
        # context-free, allocation-free residual on caller-provided device pointers (static device scratch; see the header)
        "scatter_residual_noalloc": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                     C.c_void_p, C.c_void_p, _pd, C.c_char_p, C.c_int32],
    }

    def __init__(self, path: str = LIB_PATH, prefix: str = "doma_b200_"):
        if not os.path.exists(path):
            raise DomaError(
                f"{path} not found: the CUDA extension has not been built "
                "(run `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
        self.path, self.prefix = path, prefix
        self.dll = C.CDLL(path, mode=C.RTLD_GLOBAL)
        self.fn = {}
        for name, args in self._SIGS.items():
            f = getattr(self.dll, prefix + name)
            f.argtypes, f.restype = args, C.c_int
            self.fn[name] = f
        for name, args in self._OPTIONAL.items():
            try:
                f = getattr(self.dll, prefix + name)
            except AttributeError:
                continue
            f.argtypes, f.restype = args, C.c_int
            self.fn[name] = f
        self._last_error = getattr(self.dll, prefix + "last_error")
        self._last_error.argtypes, self._last_error.restype = [C.c_void_p], C.c_char_p

    def last_error(self, handle) -> str:
        s = self._last_error(handle)
        return s.decode() if s else ""


_default_lib: Optional[Library] = None


def load_library() -> Library:
    """The product library (CUDA).  Raises if it has not been built."""
    global _default_lib
    if _default_lib is None:
        _default_lib = Library(LIB_PATH, "doma_b200_")
    return _default_lib


def write_values_text(values: np.ndarray, path: str, threads: int = 0, lib: Optional[Library] = None):
    """One "%.6g" value per line, the reference's text format; pure host code of the product library."""
    lib = lib or load_library()
    v = np.ascontiguousarray(values, dtype=np.float64).reshape(-1)
    rc = lib.fn["write_values_text"](v.ctypes.data_as(_pd), v.size, os.fsencode(path), int(threads))
    if rc:
        raise DomaError(f"write_values_text failed ({rc}): {lib.last_error(None)}")


class Context:
    """One problem resident in one implementation (on one GPU for the product library)."""

    def __init__(self, problem: Problem, lib: Optional[Library] = None):
        self.lib = lib or load_library()
        self.problem = problem
        self._h = C.c_void_p()
        cp = problem.as_c()
        rc = self.lib.fn["create"](C.byref(cp), C.byref(self._h))
        if rc != 0:
            raise DomaError(f"create failed ({rc}): {self.lib.last_error(None)}")
        self.n_cells = problem.n_cells

    def _call(self, name, *args):
        rc = self.lib.fn[name](self._h, *args)
        if rc != 0:
            raise DomaError(f"{self.lib.prefix}{name} failed ({rc}): {self.lib.last_error(self._h)}")

    def has(self, name: str) -> bool:
        return name in self.lib.fn

    def close(self):
        if self._h:
            self.lib.fn["destroy"](self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # state ---------------------------------------------------------------------------------------
    def reset_state(self, phi_init: float): self._call("reset_state", float(phi_init))
    def init_zero_ic(self): self._call("init_zero_ic")
    def update_step_fluxes(self): self._call("update_step_fluxes")
    def step_dnps(self): self._call("step_dnps")
    def synchronize(self): self._call("synchronize")
    def set_stream(self, stream_ptr: int): self._call("set_stream", C.c_void_p(stream_ptr))
    # source build --------------------------------------------------------------------------------
    def update_mg_source(self, g: int, t: float = 0.0): self._call("update_mg_source", int(g), float(t))
    def update_mg_source_eigen(self, g: int, k: float): self._call("update_mg_source_eigen", int(g), float(k))
    def update_scatter_source(self, g: int): self._call("update_scatter_source", int(g))
    # sweep ---------------------------------------------------------------------------------------
    def sweep(self, g: int): self._call("sweep", int(g))

    def sweep_from_host(self, g: int, q_host: np.ndarray, phi_out: np.ndarray):
        """Host-buffer form of sweep(g): H2D of q, sweep, D2H of phi_cur, all inside the call."""
        assert q_host.dtype == np.float64 and phi_out.dtype == np.float64
        self._call("sweep_from_host", int(g), q_host.ctypes.data_as(_pd), phi_out.ctypes.data_as(_pd))

    def stage_q_from_host(self, q_host: np.ndarray):
        """Start the H2D copy of the next sweep's source (pinned memory) on the copy stream."""
        assert q_host.dtype == np.float64 and q_host.flags.c_contiguous
        self._call("stage_q_from_host", q_host.ctypes.data_as(_pd))

    def sweep_from_host_async(self, g: int, q_host: Optional[np.ndarray], phi_out: np.ndarray):
        """Enqueue H2D of q (or use the staged one when q_host is None), sweep, D2H of phi_cur; see host_wait."""
        assert phi_out.dtype == np.float64 and phi_out.flags.c_contiguous
        qp = q_host.ctypes.data_as(_pd) if q_host is not None else C.cast(None, _pd)
        self._call("sweep_from_host_async", int(g), qp, phi_out.ctypes.data_as(_pd))

    def host_wait(self): self._call("host_wait")

    def dump_text(self, file_base: str, with_dnps: bool = False, threads: int = 0):
        """<base>_g<g>_flux.txt (and <base>_d<d>_dnps.txt): the bytes BrickMesh::dumpToTextFile / dumpDNPsToTextFile write."""
        self._call("dump_text", os.fsencode(file_base), int(bool(with_dnps)), int(threads))

    def dump_text_async(self, file_base: str, with_dnps: bool = False, threads: int = 0):
        """dump_text overlapped with the caller's next work (host snapshot now, files on a worker thread); see dump_wait."""
        self._call("dump_text_async", os.fsencode(file_base), int(bool(with_dnps)), int(threads))

    def dump_wait(self): self._call("dump_wait")

    # reductions ----------------------------------------------------------------------------------
    def source_iteration(self, g: int, tol: float, max_it: int):
        """sourceIteration(g) as one call; returns (iterations, last residual, residual of every iteration)."""
        it, res = C.c_int32(0), C.c_double(0.0)
        hist = np.zeros(max(int(max_it), 1), dtype=np.float64)
        self._call("source_iteration", int(g), float(tol), int(max_it), C.byref(it), C.byref(res), hist.ctypes.data_as(_pd))
        return int(it.value), float(res.value), hist[:int(it.value)]

    def scatter_residual(self, g: int) -> float:
        out = C.c_double()
        self._call("scatter_residual", int(g), C.byref(out))
        return out.value

    def mg_residual(self) -> float:
        out = C.c_double()
        self._call("mg_residual", C.byref(out))
        return out.value

    def update_k(self, k: float):
        """Returns (k_new, k_prev) as updateEigenvalue leaves _k, _k_prev."""
        kk, kp = C.c_double(k), C.c_double()
        self._call("update_k", C.byref(kk), C.byref(kp))
        return kk.value, kp.value

    # arrays --------------------------------------------------------------------------------------
    def get_array(self, which: int, index: int = 0) -> np.ndarray:
        out = np.empty(self.n_cells, dtype=np.float64)
        self._call("get_array", which, index, out.ctypes.data_as(_pd))
        return out

    def set_array(self, which: int, index: int, values: np.ndarray):
        v = np.ascontiguousarray(values, dtype=np.float64)
        assert v.size == self.n_cells
        self._call("set_array", which, index, v.ctypes.data_as(_pd))

    def device_ptr(self, which: int, index: int = 0) -> int:
        out = C.c_void_p()
        self._call("device_ptr", which, index, C.byref(out))
        return int(out.value)

    def info(self, what: int) -> int:
        out = C.c_int64()
        self._call("get_info", what, C.byref(out))
        return out.value

    def set_sweep_variant(self, v: int):
        """SWEEP_AUTO / SWEEP_FLAGS / SWEEP_LL: how tile-boundary faces are handed over (include/doma_b200.h)."""
        self._call("set_sweep_variant", int(v))

    def set_cell_sharding(self, on: bool):
        """Multi-rank contexts: keep / drop the cell sharding of the elementwise phases (collective; see the header)."""
        self._call("set_cell_sharding", int(bool(on)))

    def set_pairing(self, on: bool): self._call("set_pairing", int(bool(on)))
    def set_rows_per_step(self, rows: int): self._call("set_rows_per_step", int(rows))
    def set_check(self, on: bool): self._call("set_check", int(bool(on)))
    def set_tile(self, tx: int, ty: int, tz: int): self._call("set_tile", int(tx), int(ty), int(tz))

    def last_sweep_ms(self) -> float:
        out = C.c_float()
        self._call("last_sweep_ms", C.byref(out))
        return out.value

    def flux(self) -> np.ndarray:
        """phi_tot as [G, N]."""
        return np.stack([self.get_array(ARR_PHI_TOT, g) for g in range(self.problem.n_groups)])
