"""XML case files -> flat :class:`doma_b200.abi.Problem`.

Host-side mirror of the reference's input API for the 2D/3D hot path: parseInputParameters
(src/InputParameters.cpp:91-524), the interval/block expansion of BrickMesh3D/2D
(src/BrickMesh3D.cpp:33-81, src/BrickMesh2D.cpp) and the quadrature the solvers construct
(src/TransportSolver3D.cpp:10).  Same attribute names, same validation failures (raised as
InputError where the reference prints and exits).  The C++ driver (doma_b200/driver) uses the
reference's own parser instead; this module exists so the Python host layer, tests and bench read
the very same case files.
"""
from __future__ import annotations

import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

from . import abi
from .quadrature import gc_quadrature


class InputError(ValueError):
    pass


def _vec(s: str, conv):
    # parseVecFromString (src/InputParameters.cpp:9-49): strip blanks/newlines, split on ','
    s = s.replace(" ", "").replace("\n", "").replace("\t", "").replace("\r", "")
    try:
        return [conv(t) for t in s.split(",")]
    except ValueError as e:
        raise InputError(f"bad vector '{s[:40]}...': {e}")


def _dbl(node, name) -> float:
    v = node.get(name)
    try:
        x = float(v)
    except (TypeError, ValueError):
        x = -1.0
    if not x >= 0.0:
        raise InputError(f"You must provide a valid param in {name}!")
    return x


def _uint(node, name) -> int:
    v = node.get(name)
    try:
        return int(v)
    except (TypeError, ValueError):
        raise InputError(f"You must provide a valid param in {name}!")


def _str(node, name) -> str:
    v = node.get(name, "")
    if v == "":
        raise InputError(f"You must provide a valid param in {name}!")
    return v


@dataclass
class Material:
    total: List[float] = field(default_factory=list)
    scatter: List[float] = field(default_factory=list)      # [to*G+from] after transpose
    nu_sigma_f: List[float] = field(default_factory=list)
    chi_p: List[float] = field(default_factory=list)
    src: List[float] = field(default_factory=list)
    inv_v: List[float] = field(default_factory=list)
    n_delayed: int = 0
    chi_d: List[float] = field(default_factory=list)         # [g*D+d]
    beta: List[float] = field(default_factory=list)          # [g*D+d]
    lambda_: List[float] = field(default_factory=list)


@dataclass
class SourceStep:
    type: int = abi.STEP_NONE
    insert_time: float = 0.0
    remove_time: float = 0.0
    src: List[float] = field(default_factory=list)


@dataclass
class InputParameters:
    """Field-for-field twin of the reference's InputParameters (include/InputParameters.h:54-117)."""

    mode: int = abi.MODE_FIXED
    num_groups: int = 1
    t0: float = 0.0
    t1: float = 0.0
    num_steps: int = 0
    ic: str = "zero"
    fs_tol: float = 0.0
    mg_tol: float = 0.0
    k_tol: float = 0.0
    fs_max_it: int = 0
    mg_max_it: int = 0
    num_polar: int = 0
    num_azimuthal: int = 0
    scheme: int = abi.SCHEME_TWDD
    dims: int = 3
    refinement: int = 0
    x_intervals: List[int] = field(default_factory=list)
    y_intervals: List[int] = field(default_factory=list)
    z_intervals: List[int] = field(default_factory=list)
    dx: List[float] = field(default_factory=list)
    dy: List[float] = field(default_factory=list)
    dz: List[float] = field(default_factory=list)
    blocks: List[int] = field(default_factory=list)
    materials: Dict[int, Material] = field(default_factory=dict)
    step_sources: Dict[int, SourceStep] = field(default_factory=dict)

    @property
    def dt(self) -> float:
        # TransportSolver3D ctor (src/TransportSolver3D.cpp:20)
        return (self.t1 - self.t0) / float(self.num_steps) if self.mode == abi.MODE_TRANSIENT else 0.0


def parse_input_parameters(path: str) -> InputParameters:
    try:
        root = ET.parse(path).getroot()
    except (ET.ParseError, OSError) as e:
        raise InputError(f"Failed to parse {path}: {e}")
    if root.tag != "doma_inp":
        raise InputError("no <doma_inp> root")
    p = InputParameters()

    s = root.find("solver")
    if s is None:
        raise InputError("A solver block has not been provided!")
    st = s.get("solve_type", "")
    if st not in ("steady", "eigen", "transient"):
        raise InputError(f"Unknown run_mode {st}")
    p.mode = {"steady": abi.MODE_FIXED, "eigen": abi.MODE_EIGEN, "transient": abi.MODE_TRANSIENT}[st]
    sc = s.get("scheme", "")
    if sc not in ("dd", "tw_dd"):
        raise InputError(f"Unknown scheme {sc}")
    p.scheme = abi.SCHEME_DD if sc == "dd" else abi.SCHEME_TWDD
    p.num_groups = _uint(s, "num_groups")
    p.fs_tol, p.fs_max_it = _dbl(s, "fs_tol"), _uint(s, "fs_max_it")
    p.mg_tol, p.mg_max_it = _dbl(s, "mg_tol"), _uint(s, "mg_max_it")
    if p.mode == abi.MODE_EIGEN:
        p.k_tol = _dbl(s, "k_tol")
    if p.mode == abi.MODE_TRANSIENT:
        p.t0, p.t1, p.num_steps = _dbl(s, "t0"), _dbl(s, "t1"), _uint(s, "t_steps")
        p.ic = s.get("ics", "")
        if p.ic not in ("zero", "steady"):
            raise InputError(f"Unknown initial condition {p.ic}")

    aq = root.find("angular_quad")
    if aq is None:
        raise InputError("An angular quadrature block has not been provided!")
    p.num_polar, p.num_azimuthal = _uint(aq, "np"), _uint(aq, "na")

    m = root.find("mesh")
    if m is None:
        raise InputError("A mesh block has not been provided!")
    p.dims, p.refinement = _uint(m, "dims"), _uint(m, "refinement")

    def axis(d, i, what):
        dn, inn = m.find(d), m.find(i)
        if dn is None:
            raise InputError(f"Mesh {what} dimensions have not been provided!")
        if inn is None:
            raise InputError(f"Mesh {what} divisions have not been provided!")
        return _vec(_str(dn, "vals"), float), [v * (p.refinement + 1) for v in _vec(_str(inn, "vals"), int)]

    p.dx, p.x_intervals = axis("dx", "ix", "x")
    if p.dims > 1:
        p.dy, p.y_intervals = axis("dy", "iy", "y")
    if p.dims > 2:
        p.dz, p.z_intervals = axis("dz", "iz", "z")
    b = m.find("blocks")
    if b is None:
        raise InputError("Mesh blocks have not been provided!")
    p.blocks = _vec(_str(b, "vals"), int)

    mats = root.find("materials")
    if mats is None:
        raise InputError("A materials block has not been provided!")
    G = p.num_groups
    for mn in mats:
        blk = _uint(mn, "block")
        if blk in p.materials:
            raise InputError(f"multiple material properties for block {blk}!")
        mat = Material()
        mat.n_delayed = _uint(mn, "delayed_groups") if p.mode == abi.MODE_TRANSIENT else 0
        for rxn in mn:
            t = rxn.get("type", "")
            vals = _vec(_str(rxn, "mgxs"), float)
            if t == "total":
                mat.total = vals
            elif t == "scatter":
                if len(vals) != G * G:
                    raise InputError("Not enough cross sections have been provided! ('scatter')")
                if rxn.get("transpose", "")[:1] in ("1", "t", "T", "y", "Y") and rxn.get("transpose", "") != "":  # pugi as_bool
                    vals = [vals[c * G + r] for r in range(G) for c in range(G)]
                mat.scatter = vals
            elif t == "source":
                mat.src = vals
            elif t == "chi_p":
                mat.chi_p = vals
            elif t == "nu_sigma_f":
                mat.nu_sigma_f = vals
            elif t == "inv_vel":
                mat.inv_v = vals
            elif t == "chi_d":
                mat.chi_d = vals
            elif t == "beta":
                mat.beta = vals
            elif t == "lambda_d":
                mat.lambda_ = vals
            else:
                raise InputError(f"Unsupported reaction type {t}")
        tr = p.mode == abi.MODE_TRANSIENT
        checks = [
            (len(mat.total) != G, "total"), (len(mat.scatter) != G * G, "scatter"),
            (len(mat.src) not in (0, G), "source"), (len(mat.chi_p) not in (0, G), "chi"),
            (len(mat.nu_sigma_f) not in (0, G), "nu_sigma_f"), (tr and len(mat.inv_v) != G, "inv_vel"),
            (tr and len(mat.chi_d) != G * mat.n_delayed, "chi_d"), (tr and len(mat.beta) != G * mat.n_delayed, "beta"),
            (tr and len(mat.lambda_) != mat.n_delayed, "lambda_d"),
        ]
        for bad, what in checks:
            if bad:
                raise InputError(f"Not enough values have been provided for '{what}' in block {blk}")
        # chi_p without nu_sigma_f is undefined behaviour in the reference; reject
        if mat.chi_p and not mat.nu_sigma_f:
            raise InputError(f"block {blk}: chi_p given without nu_sigma_f (undefined in the reference)")
        p.materials[blk] = mat

    trn = root.find("transients")
    if trn is not None and p.mode == abi.MODE_TRANSIENT:
        for tn in trn:
            if tn.tag != "source_step":
                continue
            blk = _uint(tn, "block")
            if blk in p.step_sources:
                raise InputError(f"multiple step sources for block {blk}!")
            ss = SourceStep()
            has_i, has_r = tn.get("insert_at", "") != "", tn.get("remove_at", "") != ""
            if has_i and has_r:
                ss.type, ss.insert_time, ss.remove_time = abi.STEP_BOTH, _dbl(tn, "insert_at"), _dbl(tn, "remove_at")
                if ss.remove_time < ss.insert_time:
                    raise InputError(f"The step source for block {blk} is removed before it is inserted!")
            elif has_i:
                ss.type, ss.insert_time = abi.STEP_INSERT, _dbl(tn, "insert_at")
            elif has_r:
                ss.type, ss.remove_time = abi.STEP_REMOVE, _dbl(tn, "remove_at")
            else:
                raise InputError(f"Invalid step source for block {blk}!")
            ss.src = _vec(_str(tn, "mgxs"), float)
            if len(ss.src) != G:
                raise InputError(f"Not enough external sources for the step source in block {blk}!")
            p.step_sources[blk] = ss
    return p


def expand_axis(lengths: List[float], intervals: List[int]) -> np.ndarray:
    """Per-cell widths: each interval i contributes n_i cells of width d_i / n_i (src/BrickMesh3D.cpp:44-61)."""
    out = []
    for d, n in zip(lengths, intervals):
        out.extend([d / float(n)] * n)
    return np.array(out, dtype=np.float64)


def build_problem(p: InputParameters, device: int = 0) -> abi.Problem:
    """Flatten mesh + materials + quadrature the way the reference's objects hold them.

    Mesh: BrickMesh3D/2D ctor (src/BrickMesh3D.cpp:26-102): interval expansion, block id per cell,
    x-fastest cell order.  Materials: dense index = rank of the XML block id.  Quadrature:
    GCAngularQuadrature(2*na, 2*np, dims) (src/TransportSolver3D.cpp:10).
    """
    if p.dims not in (2, 3):
        raise InputError("only 2D and 3D cases are on the hot path")
    nbx, nby = len(p.x_intervals), len(p.y_intervals)
    nbz = len(p.z_intervals) if p.dims == 3 else 1
    if nbx * nby * nbz != len(p.blocks):
        raise InputError("Error: The blocks vector is not equal to the number of subdivisions!")
    dx, dy = expand_axis(p.dx, p.x_intervals), expand_axis(p.dy, p.y_intervals)
    dz = expand_axis(p.dz, p.z_intervals) if p.dims == 3 else np.ones(1)
    blocks = np.array(p.blocks, dtype=np.int64).reshape(nbz, nby, nbx)
    cell_blocks = np.repeat(np.repeat(np.repeat(blocks, p.z_intervals if p.dims == 3 else [1], axis=0),
                                      p.y_intervals, axis=1), p.x_intervals, axis=2)
    block_ids = np.array(sorted(p.materials.keys()), dtype=np.int64)
    missing = np.setdiff1d(np.unique(cell_blocks), block_ids)
    if missing.size:
        raise InputError(f"Block {int(missing[0])} (defined on the mesh) has no material properties!")
    mat_of_cell = np.searchsorted(block_ids, cell_blocks.reshape(-1)).astype(np.int32)

    G, M = p.num_groups, len(block_ids)
    D = max([p.materials[int(b)].n_delayed for b in block_ids] + [0])
    z = lambda *s: np.zeros(s, dtype=np.float64)
    sigma_t, sigma_s, nsf, chi_p, src, inv_v = z(M, G), z(M, G, G), z(M, G), z(M, G), z(M, G), z(M, G)
    chi_d, beta, lam = z(M, G, max(D, 0)), z(M, G, max(D, 0)), z(M, max(D, 0))
    flags, nd = np.zeros(M, np.int32), np.zeros(M, np.int32)
    st, sti, strm, ssrc = np.full(M, abi.STEP_NONE, np.int32), z(M), z(M), z(M, G)
    for m, b in enumerate(block_ids):
        mat = p.materials[int(b)]
        sigma_t[m] = mat.total
        sigma_s[m] = np.array(mat.scatter).reshape(G, G)
        if mat.src:
            src[m] = mat.src
            flags[m] |= abi.MAT_HAS_SRC
        if mat.chi_p:
            chi_p[m] = mat.chi_p
            flags[m] |= abi.MAT_HAS_CHI_P
        if mat.nu_sigma_f:
            nsf[m] = mat.nu_sigma_f
            flags[m] |= abi.MAT_HAS_NU_SIGMA_F
        if mat.inv_v:
            inv_v[m] = mat.inv_v
        nd[m] = mat.n_delayed
        if mat.n_delayed:
            if not mat.nu_sigma_f:
                raise InputError(f"block {int(b)}: delayed groups without nu_sigma_f")
            chi_d[m, :, :mat.n_delayed] = np.array(mat.chi_d).reshape(G, mat.n_delayed)
            beta[m, :, :mat.n_delayed] = np.array(mat.beta).reshape(G, mat.n_delayed)
            lam[m, :mat.n_delayed] = mat.lambda_
        if int(b) in p.step_sources:
            ss = p.step_sources[int(b)]
            st[m], sti[m], strm[m], ssrc[m] = ss.type, ss.insert_time, ss.remove_time, ss.src
    octs, mu, eta, xi, w = gc_quadrature(2 * p.num_azimuthal, 2 * p.num_polar, p.dims)
    return abi.Problem(
        dims=p.dims, dx=dx, dy=dy, dz=dz, mat_of_cell=mat_of_cell, n_groups=G, mat_flags=flags, n_delayed=nd,
        sigma_t=sigma_t, sigma_s=sigma_s, nu_sigma_f=nsf, chi_p=chi_p, src=src, inv_v=inv_v, chi_d=chi_d, beta=beta,
        lambda_=lam, step_type=st, step_t_insert=sti, step_t_remove=strm, step_src=ssrc, ord_octant=octs, mu=mu,
        eta=eta, xi=xi, w=w, scheme=p.scheme, mode=p.mode, dt=p.dt, device=device, block_ids=block_ids)


def load_case(path: str, device: int = 0):
    """(InputParameters, Problem) for an XML case file."""
    params = parse_input_parameters(path)
    return params, build_problem(params, device)
