"""Text output in the reference's on-disk format (BrickMesh3D::dumpToTextFile, src/BrickMesh3D.cpp:216-265;
BrickMesh2D::dumpToTextFile, src/BrickMesh2D.cpp:169-221) - the contract of the reference's make_vtk.py and plot scripts.

The solve-dependent files (<base>_g<g>_flux.txt, <base>_d<d>_dnps.txt) are written by the library
(doma_b200_dump_text: device -> host, multi-threaded "%.6g" formatting).  The geometry files written here do not depend
on the solve: <base>_dims.txt, <base>_blocks.txt, <base>_meshx.txt, <base>_meshy.txt (, <base>_meshz.txt).
"""
from __future__ import annotations

import numpy as np

from . import abi


def cell_centres(widths: np.ndarray) -> np.ndarray:
    """The reference's running centre: x_c = 0.5 * dx[0], then x_c += dx[i] after every cell
    (src/BrickMesh3D.cpp:84-101) - exact for uniform widths, reproduced as it is for graded ones."""
    w = np.asarray(widths, dtype=np.float64)
    c = np.empty_like(w)
    acc = 0.5 * w[0]
    for i in range(w.size):          # sequential adds, the reference's rounding
        c[i] = acc
        acc += w[i]
    return c


def dump_geometry(prob: abi.Problem, file_base: str, lib=None):
    nx, ny, nz = prob.nx, prob.ny, (prob.nz if prob.dims == 3 else 1)
    with open(file_base + "_dims.txt", "w") as f:
        f.write(f"num_x: {nx}\nnum_y: {ny}\n" + (f"num_z: {nz}\n" if prob.dims == 3 else "") + f"num_g: {prob.n_groups}\n")
    ids = np.asarray(prob.block_ids if prob.block_ids is not None else np.arange(prob.n_mat))
    blocks = ids[np.asarray(prob.mat_of_cell)]
    np.savetxt(file_base + "_blocks.txt", blocks, fmt="%d")
    xc, yc = cell_centres(prob.dx), cell_centres(prob.dy)
    shape = (nz, ny, nx)
    abi.write_values_text(np.broadcast_to(xc[None, None, :], shape), file_base + "_meshx.txt", lib=lib)
    abi.write_values_text(np.broadcast_to(yc[None, :, None], shape), file_base + "_meshy.txt", lib=lib)
    if prob.dims == 3:
        zc = cell_centres(prob.dz)
        abi.write_values_text(np.broadcast_to(zc[:, None, None], shape), file_base + "_meshz.txt", lib=lib)


def dump_to_text_file(ctx: abi.Context, prob: abi.Problem, file_base: str, only_flux: bool = False, with_dnps: bool = False,
                      asynchronous: bool = False):
    """dumpToTextFile(file_base, only_flux) [+ dumpDNPsToTextFile(file_base)].  asynchronous: the flux / precursor files
    are written by the library's worker thread while the caller goes on (ctx.dump_wait() or the next dump joins it)."""
    if not only_flux:
        dump_geometry(prob, file_base, ctx.lib)
    if asynchronous and ctx.has("dump_text_async"):
        ctx.dump_text_async(file_base, with_dnps=with_dnps)
    else:
        ctx.dump_text(file_base, with_dnps=with_dnps)
