"""Gauss-Legendre x Chebyshev product quadrature, host-side mirror of the reference's tables.

Mirrors GCAngularQuadrature (src/GCAngularQuadrature.cpp:9-55) and LegendrePolynomial /
ChebyshevPolynomial (src/AngularPolynomials.cpp:8-111): same Newton iteration, same operation order,
same binning into octants in generation order, so that the ordinates the GPU sweeps with are the
ones the reference's driver would hand over (tests pin this against the reference's own tables).
The transport solvers build it as GCAngularQuadrature(2*na, 2*np, dims)
(src/TransportSolver3D.cpp:10).
"""
from __future__ import annotations

import math
from typing import List, Tuple

import numpy as np


def _legendre_eval(degree: int, x: float) -> float:
    # src/AngularPolynomials.cpp:42-62
    if degree == 0:
        return 1.0
    if degree == 1:
        return x
    v, v1, v2 = 1.0, x, 1.0
    for i in range(2, degree + 1):
        v = ((2.0 * float(i) - 1.0) * x * v1 - (float(i) - 1.0) * v2) / float(i)
        v2, v1 = v1, v
    return v


def _legendre_deriv(degree: int, x: float) -> float:
    # src/AngularPolynomials.cpp:66-91
    if degree == 0:
        return 0.0
    if degree == 1:
        return 1.0
    v, d, v1, v2 = 1.0, 0.0, x, 1.0
    f = 1.0 / (x * x - 1.0)
    for i in range(2, degree + 1):
        v = ((2.0 * float(i) - 1.0) * x * v1 - (float(i) - 1.0) * v2) / float(i)
        d = float(i) * f * (x * v - v1)
        v2, v1 = v1, v
    return d


def gauss_legendre(degree: int) -> Tuple[List[float], List[float]]:
    """Roots (descending) and weights; Newton from the cos() guess until |dx| <= 2e-16 (:18-37)."""
    roots, weights = [], []
    for i in range(1, degree + 1):
        x = math.cos(math.pi * (float(i) - 0.25) / (float(degree) + 0.5))
        v, d = _legendre_eval(degree, x), _legendre_deriv(degree, x)
        while True:
            dr = v / d
            x -= dr
            v, d = _legendre_eval(degree, x), _legendre_deriv(degree, x)
            if not abs(dr) > 2e-16:
                break
        roots.append(x)
        weights.append(2.0 / ((1.0 - x * x) * d * d))
    return roots, weights


def chebyshev_angles(degree: int) -> Tuple[List[float], List[float]]:
    """Angular roots omega_j = (2j-1)/(2n) * pi and weights pi/n (:99-111)."""
    om = [(2 * float(i) - 1.0) / (2.0 * float(degree)) * math.pi for i in range(1, degree + 1)]
    return om, [math.pi / float(degree)] * degree


def classify(mu: float, eta: float, xi: float) -> int:
    """Octant enum value (include/TransportBase.h:20-30): bit2 = mu<0, bit1 = eta<0, bit0 = xi<0."""
    if mu == 0.0 or eta == 0.0 or xi == 0.0:
        raise ValueError("ordinate on an octant boundary (the reference exits here)")
    return (4 if mu < 0.0 else 0) | (2 if eta < 0.0 else 0) | (1 if xi < 0.0 else 0)


def gc_quadrature(n_c: int, n_l: int, dims: int):
    """Returns (ord_octant[int32], mu, eta, xi, w) octant-major, generation order inside an octant."""
    roots, wl = gauss_legendre(n_l)
    om, wc = chebyshev_angles(n_c)
    bins = [[] for _ in range(8)]
    for i in range(n_l):
        for j in range(n_c):
            weight = wl[i] * wc[j]
            mu = roots[i]
            s = math.sqrt(1.0 - (mu * mu))
            eta = s * math.cos(om[j])
            xi = s * math.sin(om[j])
            bins[classify(mu, eta, xi)].append((mu, eta, xi, weight))
            if dims > 2:
                nmu, neta, nxi = -1.0 * mu, -1.0 * s * math.cos(om[j]), -1.0 * s * math.sin(om[j])
                bins[classify(nmu, neta, nxi)].append((nmu, neta, nxi, weight))
    octs, rows = [], []
    for o in range(8):
        for r in bins[o]:
            octs.append(o)
            rows.append(r)
    a = np.array(rows, dtype=np.float64).reshape(-1, 4)
    return (np.array(octs, dtype=np.int32), a[:, 0].copy(), a[:, 1].copy(), a[:, 2].copy(), a[:, 3].copy())
