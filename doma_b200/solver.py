"""Host-side mirror of the reference's outer loops, driving the C ABI.

``TransportSolver`` follows TransportSolver3D<T> / TransportSolver2D<T> member for member
(src/TransportSolver3D.cpp; the 2D file has the same lines): solveFixedSource (:262-316),
solveEigenvalue (:29-85), solveTransient (:88-138), sourceIteration (:458-497), initZeroIC
(:140-156).  Every starred member of SURVEY.md section 8b is ONE call into the C ABI
(include/doma_b200.h); nothing numeric happens in Python.  The loops, the convergence tests and
the order of calls are the reference's, so iteration traces are comparable one to one.

The C++ driver in doma_b200/driver binds the same ABI under the reference's own unmodified loops;
this Python twin exists because the GPU box has no reference tree to compile against, and it is
what the parity tests and bench.py call.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, List, Optional

import numpy as np

from . import abi
from .input_parameters import InputParameters


@dataclass
class Trace:
    """What the reference prints, as data: per-(outer, group) SI counts and residuals, outer residuals, k."""

    si_counts: List[int] = field(default_factory=list)
    si_residuals: List[float] = field(default_factory=list)
    outer_residuals: List[float] = field(default_factory=list)
    k_history: List[float] = field(default_factory=list)
    sweeps: int = 0


class TransportSolver:
    def __init__(self, ctx: abi.Context, params: InputParameters, verbose: bool = False,
                 allreduce: Optional[Callable[[abi.Context], None]] = None, log: Callable[[str], None] = None,
                 fused_si: bool = False, output_base: str = "", problem: Optional[abi.Problem] = None):
        self.ctx = ctx
        # text dumps at the reference's dump points (3D:75,112,127-128,306); "" = none, as in the reference
        self.output_base, self.problem = output_base, problem
        self.num_groups = params.num_groups
        self.mode = params.mode
        self.sit, self.smi = params.fs_tol, params.fs_max_it
        self.mgt, self.mgi = params.mg_tol, params.mg_max_it
        self.k, self.k_prev, self.k_tol = 1.0, 0.0, params.k_tol
        self.t0, self.dt, self.t_steps, self.ic = params.t0, params.dt, params.num_steps, params.ic
        self.verbose = verbose
        self.allreduce = allreduce       # sums phi_cur over ranks after each sweep (angle sharding)
        # fused_si: sourceIteration runs inside the library (one ABI call per group).  Only when the
        # all-reduce is the library's own NCCL one (or there is a single rank): a host-side hook cannot be fused.
        self.fused_si = bool(fused_si) and ctx.has("source_iteration") and (allreduce is None or ctx.has("comm_init"))
        self.trace = Trace()
        self._log = log or (lambda s: None)

    # --- members that are one ABI call each ------------------------------------------------------
    def initialize_solve(self):
        self.ctx.reset_state(1.0 if self.mode == abi.MODE_EIGEN else 0.0)

    def sweep(self, g: int):
        self.ctx.sweep(g)
        if self.allreduce is not None:
            self.allreduce(self.ctx)
        self.trace.sweeps += 1

    # --- sourceIteration (3D:458-497) ------------------------------------------------------------
    def source_iteration(self, g: int) -> bool:
        if self.fused_si:
            return self._source_iteration_fused(g)
        it, res = 0, 0.0
        while True:
            if self.verbose:
                self._log(f"Performing SI {it} for G{g}" + (f", residual = {res!r}" if it else ""))
            self.sweep(g)
            res = self.ctx.scatter_residual(g)
            self.ctx.update_scatter_source(g)
            it += 1
            if not (it < self.smi and self.sit < res):
                break
        self.trace.si_counts.append(it)
        self.trace.si_residuals.append(res)
        ok = it < self.smi
        self._log(f"SI {'converged' if ok else 'failed to converge'} after {it} iterations with a residual of {res!r} for G{g}")
        return ok

    def _source_iteration_fused(self, g: int) -> bool:
        """The same loop owned by the library (doma_b200_source_iteration): sweep, all-reduce, residual and
        scattering-source update without returning to the host in between; prints replayed from the history."""
        it, res, hist = self.ctx.source_iteration(g, self.sit, self.smi)
        if self.verbose:
            for i in range(it):
                self._log(f"Performing SI {i} for G{g}" + (f", residual = {hist[i - 1]!r}" if i else ""))
        self.trace.sweeps += it
        self.trace.si_counts.append(it)
        self.trace.si_residuals.append(res)
        ok = it < self.smi
        self._log(f"SI {'converged' if ok else 'failed to converge'} after {it} iterations with a residual of {res!r} for G{g}")
        return ok

    # --- solveFixedSource (3D:262-316) -----------------------------------------------------------
    def _dump(self, base: str, only_flux: bool = False, dnps: bool = False):
        from .output import dump_to_text_file
        # overlapped with the next time step / the caller's next work; solve() waits for the last one
        dump_to_text_file(self.ctx, self.problem, base, only_flux, with_dnps=dnps, asynchronous=True)

    def solve_fixed_source(self, t: float = 0.0, output_base: Optional[str] = None) -> bool:
        output_base = (self.output_base if self.mode == abi.MODE_FIXED else "") if output_base is None else output_base
        if self.mode != abi.MODE_TRANSIENT:
            self.initialize_solve()
        it, res = 0, 0.0
        while True:
            self._log(f"Performing MGI {it}" + (f", residual = {res!r}" if it else ""))
            for g in range(self.num_groups):
                self.ctx.update_mg_source(g, t)
                if not self.source_iteration(g):
                    return False
            res = self.ctx.mg_residual()
            self.trace.outer_residuals.append(res)
            it += 1
            if not (it < self.mgi and self.mgt < res and self.num_groups > 1):
                break
        ok = it < self.mgi
        if ok and self.num_groups > 1:
            self._log(f"MGI converged after {it} iterations with a residual of {res!r}")
        if ok and output_base:
            self._dump(output_base)                                        # 3D:305-306
        if not ok:
            self._log(f"MGI failed to converge after {it} iterations with a residual of {res!r}")
        return ok

    # --- solveEigenvalue (3D:29-85) --------------------------------------------------------------
    def solve_eigenvalue(self) -> bool:
        self.initialize_solve()
        it, shape_res, k_diff = 0, 0.0, 0.0
        while True:
            self._log(f"Performing PI {it}" + (f", shape residual = {shape_res!r}, k residual = {k_diff!r}\nk = {self.k!r}" if it else ""))
            for g in range(self.num_groups):
                self.ctx.update_mg_source_eigen(g, self.k)
                if not self.source_iteration(g):
                    return False
            self.k, self.k_prev = self.ctx.update_k(self.k)
            shape_res = self.ctx.mg_residual()
            k_diff = abs(self.k - self.k_prev)
            self.trace.outer_residuals.append(shape_res)
            self.trace.k_history.append(self.k)
            it += 1
            if not (it < self.mgi and (self.mgt < shape_res or self.k_tol < k_diff)):
                break
        ok = it < self.mgi
        self._log(f"PI {'converged' if ok else 'failed to converge'} after {it} iterations with a flux shape residual of "
                  f"{shape_res!r} and a k residual of {k_diff!r}." + (f"\nConverged value of k = {self.k!r}" if ok else ""))
        if ok and self.output_base:
            self._dump(self.output_base)                                   # 3D:74-75
        return ok

    # --- solveTransient (3D:88-138) --------------------------------------------------------------
    def solve_transient(self, on_step: Optional[Callable[[int, "TransportSolver"], None]] = None) -> bool:
        if self.ic == "zero":
            self.ctx.init_zero_ic()
        else:
            # ics="steady" indexes empty vectors in the reference (undefined behaviour)
            raise NotImplementedError("ics='steady' is undefined behaviour in the reference; unsupported")
        if self.output_base:
            self._dump(self.output_base)                                   # 3D:112, the initial condition
        t = 0.0
        for step in range(self.t_steps):
            self._log(f"Solving timestep {step}...")
            if not self.solve_fixed_source(t, ""):
                return False
            self.ctx.step_dnps()
            if self.output_base:
                self._dump(f"{self.output_base}_t{step}", only_flux=True, dnps=True)   # 3D:127-128
            if on_step is not None:
                on_step(step, self)
            if step != self.t_steps - 1:
                self.ctx.update_step_fluxes()
            t += self.dt
        return True

    def solve(self) -> bool:
        try:
            if self.mode == abi.MODE_FIXED:
                return self.solve_fixed_source()
            if self.mode == abi.MODE_EIGEN:
                return self.solve_eigenvalue()
            return self.solve_transient()
        finally:
            if self.output_base and self.ctx.has("dump_wait"):
                self.ctx.dump_wait()

    def flux(self) -> np.ndarray:
        return self.ctx.flux()
