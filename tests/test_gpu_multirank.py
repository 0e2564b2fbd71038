"""N > 1 on real GPUs: two processes, one GPU each, NCCL.  The ordinates are sharded over the ranks; the elementwise
phases run cell-sharded (all-gather q -> sweep -> reduce-scatter phi_cur -> own-cell residual / update / source build,
the default) or replicated behind an all-reduce (the r01 scheme).  Both must reproduce the reference's golden
fixtures - flux 1e-10, k 1e-9, identical iteration traces - and leave every rank with the same gathered flux.
Skipped on a single-GPU box; `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multirank.py -m gpu` runs it.
"""
import os
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from conftest import ROOT, golden_case_path, load_golden

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, name, out_dir, cell_sharded, fused):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    import doma_b200 as d
    from doma_b200 import abi
    from doma_b200.sharding import make_allreduce, shard_ordinates

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    params, prob = d.load_case(golden_case_path(name))
    prob.device = rank
    shard = prob.with_ordinates(shard_ordinates(prob, rank, world))
    ctx = abi.Context(shard)
    allreduce = make_allreduce(ctx)
    if not cell_sharded:
        ctx.set_cell_sharding(False)
    lo, hi = ctx.info(abi.INFO_CELL_LO), ctx.info(abi.INFO_CELL_HI)
    assert (hi - lo < prob.n_cells) == cell_sharded
    s = d.TransportSolver(ctx, params, allreduce=allreduce, fused_si=fused)
    ok = s.solve()
    dnp = np.stack([ctx.get_array(abi.ARR_DNP_CUR, i) for i in range(prob.n_delayed_max)]) if params.mode == abi.MODE_TRANSIENT and prob.n_delayed_max else np.zeros(0)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), phi=s.flux(), ok=ok, si=np.array(s.trace.si_counts), k=s.k, dnp=dnp,
             launches=ctx.info(abi.INFO_KERNEL_LAUNCHES))
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("fused", [True, False], ids=["fused_si", "loop_on_host"])
@pytest.mark.parametrize("cell_sharded", [True, False], ids=["cell_sharded", "replicated"])
@pytest.mark.parametrize("name", ["k1_r0_s3", "asm_eig_small", "tr2d_r0_remove"])
def test_two_gpus_reproduce_the_reference(name, cell_sharded, fused, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    port = 29600 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, name, str(tmp_path), cell_sharded, fused), nprocs=2, join=True)
    g = load_golden(name)
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    assert bool(r0["ok"]) == bool(g["ok"]) and list(r0["si"]) == list(g["si_counts"]) == list(r1["si"])
    assert np.array_equal(r0["phi"], r1["phi"])                       # every rank sees the same flux
    rel = np.max(np.abs(r0["phi"] - g["phi"]) / np.maximum(np.abs(g["phi"]), 1e-300))
    assert rel <= 1e-10
    if "eig" in name:
        assert abs(float(r0["k"]) - float(g["k"])) <= 1e-9 and float(r0["k"]) == float(r1["k"])
    if name.startswith("tr2d"):
        assert np.max(np.abs(r0["dnp"] - g["dnp"]) / np.maximum(np.abs(g["dnp"]), 1e-300)) <= 1e-10
    assert int(r0["launches"]) > 0
