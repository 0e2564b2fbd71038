"""Synthetic fixed-work cases of the named shapes (BASELINE.json configs[4]).

``synthetic_xml`` writes a case file in the reference's own input format, so that the reference
arm (oracle/_ref/ref_harness) and this package parse IDENTICAL text: n^3 cells of 1 cm in 4x4x4
blocks, checkerboard of two scattering materials plus one source block, G groups with
down-scatter only, S16 := np=na=8 (512 ordinates, 64 per octant).  Every double is printed with
%.17g.  Two deterministic work modes:
  A: fs_tol=0, fs_max_it=K      -> exactly K sweeps of group 0 (then "SI failed", state comparable)
  B: fs_tol=1e30, mg_tol=1e30, fs_max_it=2, mg_max_it=2 -> exactly one sweep per group, all G groups
"""
from __future__ import annotations

import numpy as np

SEED = 20261104


def synthetic_materials(G: int, seed: int = SEED):
    rng = np.random.default_rng(seed)
    mats = []
    for mid in range(2):
        st = rng.uniform(0.2, 1.0, G)
        c = rng.uniform(0.3, 0.7, G)
        ss = np.zeros((G, G))                      # [to, from]
        for gf in range(G):
            ss[gf, gf] = c[gf] * st[gf]
            for gt in range(gf + 1, G):
                ss[gt, gf] = rng.uniform(0.0, 0.05) * st[gf]
            tot = ss[:, gf].sum()
            if tot > 0.95 * st[gf]:
                ss[:, gf] *= 0.95 * st[gf] / tot
        mats.append((st, ss, None))
    src = np.zeros(G)
    src[0] = 1.0
    mats.append((mats[0][0].copy(), mats[0][1].copy(), src))
    return mats


def synthetic_xml(n: int, G: int = 8, npol: int = 8, nazi: int = 8, mode: str = "A", sweeps: int = 2,
                  nblk: int = 4, scheme: str = "tw_dd", shape=None) -> str:
    nx, ny, nz = shape if shape is not None else (n, n, n)
    for v in (nx, ny, nz):
        assert v % nblk == 0, "cells per axis must be a multiple of the block count"
    if mode == "A":
        solver = f'fs_tol="0.0" fs_max_it="{sweeps}" mg_tol="0.0" mg_max_it="1"'
    else:
        solver = 'fs_tol="1e30" fs_max_it="2" mg_tol="1e30" mg_max_it="2"'
    f = lambda a: ", ".join("%.17g" % v for v in np.asarray(a).reshape(-1))
    ax = lambda m: (", ".join(["%.17g" % (m / nblk)] * nblk), ", ".join([str(m // nblk)] * nblk))
    blocks = []
    for bk in range(nblk):
        for bj in range(nblk):
            for bi in range(nblk):
                blocks.append(2 if (bi, bj, bk) == (0, 0, 0) else (bi + bj + bk) % 2)
    out = ['<doma_inp>',
           f'  <solver solve_type="steady" scheme="{scheme}" num_groups="{G}" {solver}/>',
           f'  <angular_quad type="gc" np="{npol}" na="{nazi}"/>',
           '  <mesh dims="3" refinement="0">']
    for tag, m in (("x", nx), ("y", ny), ("z", nz)):
        d, i = ax(m)
        out.append(f'    <d{tag} vals="{d}"/>')
        out.append(f'    <i{tag} vals="{i}"/>')
    out.append(f'    <blocks vals="{", ".join(str(b) for b in blocks)}"/>')
    out.append('  </mesh>')
    out.append('  <materials>')
    for mid, (st, ss, src) in enumerate(synthetic_materials(G)):
        out.append(f'    <mat block="{mid}">')
        out.append(f'      <reaction type="total" mgxs="{f(st)}"/>')
        out.append(f'      <reaction type="scatter" mgxs="{f(ss)}"/>')
        if src is not None:
            out.append(f'      <reaction type="source" mgxs="{f(src)}"/>')
        out.append('    </mat>')
    out.append('  </materials>')
    out.append('</doma_inp>')
    return "\n".join(out) + "\n"


def synthetic_case(path: str, n: int, **kw):
    """Write the XML to ``path`` and return (InputParameters, Problem) parsed back from it."""
    from .input_parameters import load_case

    with open(path, "w") as fh:
        fh.write(synthetic_xml(n, **kw))
    return load_case(path)


def random_problem(shape, G=2, n_mat=3, npol=2, nazi=2, seed=0, dims=3, scheme=1, mode=0, D=0, dt=0.0, void=True):
    """Seeded irregular problem for kernel-vs-oracle tests: non-uniform widths, random materials
    (optionally one void), random material map.  Not a physical benchmark."""
    from . import abi
    from .quadrature import gc_quadrature

    rng = np.random.default_rng(seed)
    nx, ny = shape[0], shape[1]
    nz = shape[2] if dims == 3 else 1
    dx, dy = rng.uniform(0.2, 2.0, nx), rng.uniform(0.2, 2.0, ny)
    dz = rng.uniform(0.2, 2.0, nz) if dims == 3 else np.ones(1)
    M = n_mat
    sigma_t = rng.uniform(0.05, 2.0, (M, G))
    sigma_s = rng.uniform(0.0, 0.3, (M, G, G)) * sigma_t[:, None, :]
    if void and M > 1:
        sigma_t[M - 1] = 0.0
        sigma_s[M - 1] = 0.0
    nsf = rng.uniform(0.0, 0.2, (M, G))
    chi = rng.uniform(0.0, 1.0, (M, G))
    chi /= chi.sum(axis=1, keepdims=True)
    src = rng.uniform(0.0, 2.0, (M, G))
    flags = np.full(M, abi.MAT_HAS_SRC | abi.MAT_HAS_CHI_P | abi.MAT_HAS_NU_SIGMA_F, np.int32)
    flags[0] = abi.MAT_HAS_SRC
    nsf[0] = 0.0
    chi[0] = 0.0
    nd = np.zeros(M, np.int32)
    if D:
        nd[1:] = D
    octs, mu, eta, xi, w = gc_quadrature(2 * nazi, 2 * npol, dims)
    return abi.Problem(
        dims=dims, dx=dx, dy=dy, dz=dz, mat_of_cell=rng.integers(0, M, nx * ny * nz).astype(np.int32), n_groups=G,
        mat_flags=flags, n_delayed=nd, sigma_t=sigma_t, sigma_s=sigma_s, nu_sigma_f=nsf, chi_p=chi, src=src,
        inv_v=rng.uniform(1e-3, 1e-1, (M, G)), chi_d=rng.uniform(0, 1, (M, G, D)), beta=rng.uniform(0, 1e-2, (M, G, D)),
        lambda_=rng.uniform(0.01, 3.0, (M, D)), step_type=np.full(M, abi.STEP_NONE, np.int32), step_t_insert=np.zeros(M),
        step_t_remove=np.zeros(M), step_src=np.zeros((M, G)), ord_octant=octs, mu=mu, eta=eta, xi=xi, w=w, scheme=scheme,
        mode=mode, dt=dt)
