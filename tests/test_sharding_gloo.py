"""N>1 host path on CPU: two processes over gloo, each owning half of the sweep units, phi_cur summed
with torch.distributed after every sweep (the NCCL all-reduce's stand-in).  Contexts are the CPU
oracle here; the sharding, the solver loop and the all-reduce hook are the product's."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

from conftest import ORACLE_SO, ROOT, golden_case_path, load_golden


def _worker(rank, world, port, name, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    import doma_b200 as d
    from doma_b200 import abi
    from doma_b200.sharding import make_allreduce, shard_ordinates

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lib = abi.Library(ORACLE_SO, "doma_oracle_")
    params, prob = d.load_case(golden_case_path(name))
    shard = prob.with_ordinates(shard_ordinates(prob, rank, world))
    ctx = abi.Context(shard, lib)
    s = d.TransportSolver(ctx, params, allreduce=make_allreduce(ctx))
    ok = s.solve()
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), phi=s.flux(), ok=ok, si=np.array(s.trace.si_counts), k=s.k)
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["k1_r0_s3", "asm_eig_small_1sweep"])
def test_two_ranks_reproduce_the_reference(name, tmp_path, oracle_lib):
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, name, str(tmp_path)), nprocs=2, join=True)
    g = load_golden(name)
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    assert bool(r0["ok"]) == bool(g["ok"]) and list(r0["si"]) == list(g["si_counts"]) == list(r1["si"])
    assert np.array_equal(r0["phi"], r1["phi"])                       # every rank holds the same reduced flux
    rel = np.max(np.abs(r0["phi"] - g["phi"]) / np.maximum(np.abs(g["phi"]), 1e-300))
    assert rel < 1e-12                                                 # only the angle summation order changed
    if name.startswith("asm_eig"):
        assert abs(float(r0["k"]) - float(g["k"])) < 1e-12
