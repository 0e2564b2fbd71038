"""Angle sharding across ranks (one process per GPU) and the phi all-reduce that follows a sweep.

The ordinates of every octant are independent given q (all boundaries are vacuum,
src/CartesianCell3D.cpp:31-38), and phi_cur is a plain sum over ordinates, so each rank sweeps a
subset of the ordinates over the full mesh and the increment is summed once per sweep - the
multi-GPU replacement of the reference's thread-sum (src/TransportSolver3D.cpp:609-621).

Sharding unit = a warp's worth of ordinates of one octant (<= 32 lanes), dealt round-robin, so that
lanes stay full: with S16 (64 per octant) there are 16 units, i.e. 2 whole octants per rank at 4
ranks and one octant (2 units) per rank at 8.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import numpy as np

from . import abi


def shard_units(prob: abi.Problem) -> List[np.ndarray]:
    """Index arrays of the sweep units, in the order libdoma_b200 builds them (api.cu build_units)."""
    units = []
    for oct_ in range(8):
        idx = np.nonzero(prob.ord_octant == oct_)[0]
        if idx.size == 0:
            continue
        nb = (idx.size + 31) // 32
        pos = 0
        for b in range(nb):
            sz = idx.size // nb + (1 if b < idx.size % nb else 0)
            units.append(idx[pos:pos + sz])
            pos += sz
    return units


def shard_ordinates(prob: abi.Problem, rank: int, world: int) -> np.ndarray:
    """Ordinate indices owned by ``rank``: units dealt so that consecutive ranks get different octants."""
    units = shard_units(prob)
    mine = [u for i, u in enumerate(units) if i % world == rank]
    return np.sort(np.concatenate(mine)) if mine else np.zeros(0, dtype=np.int64)


def make_allreduce(ctx: abi.Context, group=None) -> Optional[Callable[[abi.Context], None]]:
    """Returns the per-sweep all-reduce for this process' rank, or None when world_size == 1.

    CUDA contexts: NCCL in-place all-reduce on the context's stream (doma_b200_allreduce_phi_cur),
    communicator bootstrapped by broadcasting the NCCL unique id over torch.distributed.
    Host contexts (CPU tests over gloo): phi_cur goes through torch.distributed.all_reduce.
    """
    import torch.distributed as dist

    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return None
    import torch

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if ctx.has("comm_init"):
        import ctypes as C

        buf = C.create_string_buffer(128)
        if rank == 0:
            rc = ctx.lib.fn["comm_unique_id"](buf, 128)
            if rc:
                raise abi.DomaError(f"comm_unique_id failed: {ctx.lib.last_error(None)}")
        t = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
        if dist.get_backend(group) == "nccl":
            t = t.cuda()
        dist.broadcast(t, src=0, group=group)
        raw = bytes(t.cpu().numpy().tobytes())
        ctx._call("comm_init", rank, world, raw, 128)
        return lambda c: c._call("allreduce_phi_cur")

    def host_allreduce(c: abi.Context):
        v = torch.from_numpy(c.get_array(abi.ARR_PHI_CUR, 0))
        dist.all_reduce(v, group=group)
        c.set_array(abi.ARR_PHI_CUR, 0, v.numpy())

    return host_allreduce
