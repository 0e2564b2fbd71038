#!/usr/bin/env python
"""bench.py - sweep throughput of the B200 path on the synthetic 256^3 x 8-group x S16 case.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--n 256] [--impl reference]

One "step" = one Gauss-Seidel pass over all G groups of the synthetic case (Mode B):
for every group: source build -> sweep -> (phi all-reduce when N > 1) -> scattering residual ->
scattering-source update; then the multigroup residual.  That is N_cells * N_ordinates * G
cell-angle-group updates per step.  With N > 1 (launched by torchrun, one rank per GPU) the sweep
units (octant x 32-ordinate batch) are dealt across ranks, each rank sweeps its ordinates over the
full mesh and phi_cur is summed with ncclAllReduce: a fixed problem, "scaling": "strong".

Printed JSON (rank 0, one line): the contract keys, plus
  roofline     - the sweep kernel against the HBM roofline: 48 algorithmic bytes per update
                 (3 face-flux reads + 3 writes, as the reference moves them) / CUDA-event time
  cpu_baseline - the UNMODIFIED reference (oracle/_ref/ref_harness, OpenMP) timed on this box's
                 host cores on a bounded sample of the same case (4 sweeps of group 0 at 128^3)
  e2e          - the same step driven through the host-buffer calls (doma_b200_stage_q_from_host /
                 doma_b200_sweep_from_host_async / doma_b200_host_wait): q pinned host -> device and
                 phi_cur device -> host for every sweep, copies overlapped with the neighbouring kernels
--impl reference times the reference itself (rank 0 only) on the bounded sample, same metric.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "sweep cell-angle-group updates/s"
UNIT = "updates/s"
BYTES_PER_UPDATE = 48.0   # 3 upwind face reads + 3 downwind face writes x 8 B
FP64_OPS_PER_UPDATE = 118.0  # 13 divisions + 105 other FP64 operations as the reference writes them
FP64_INST_PER_UPDATE = 55.7   # FP64-pipe instructions per update in the shipped SASS (368 DFMA + 264 DADD + 259 DMUL per 16 cells)
FP64_WARP_INST_PER_CLK_SM = 2.0  # issue ceiling of the FP64 pipe: 64 lanes per SM (scripts/fp64_peak.cu measures 1.98 with DADD)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm = sorted(float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        mx = max([float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()] or [0.0])
        pw = max([float(r[2]) for r in self.rows if len(r) >= 7 and r[2].replace(".", "").isdigit()] or [0.0])
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "power_w_max": pw or None,
                "samples": len(sm), "reasons": reasons}


def run_reference_sample(n, G, steps, warmup, threads, mode="B", sweeps=1):
    """Times the unmodified reference on a bounded sample; returns (updates/s, seconds per step, description)."""
    from doma_b200.synthetic import synthetic_xml

    harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    if not os.path.exists(harness):
        return None, None, "oracle/_ref/ref_harness missing (build() compiles it where /root/reference exists)"
    tmp = tempfile.mkdtemp(prefix="doma_ref_")
    xml = os.path.join(tmp, "sample.xml")
    with open(xml, "w") as f:
        f.write(synthetic_xml(n, G=G, mode=mode, sweeps=sweeps))
    env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="false")
    times = []
    for i in range(warmup + steps):
        r = subprocess.run([harness, xml, str(threads), os.path.join(tmp, "out"), "--no-dump"], capture_output=True, text=True, env=env)
        if r.returncode != 0:
            return None, None, "ref_harness failed: " + r.stderr[-200:]
        sec = float(r.stdout.rsplit("solve_seconds=", 1)[1].split()[0])
        if i >= warmup:
            times.append(sec)
    upd = float(n) ** 3 * 512 * (G if mode == "B" else sweeps)
    sec = sum(times) / len(times)
    what = f"one sweep of each of the {G} groups (Mode B)" if mode == "B" else f"{sweeps} sweep(s) of group 0 (Mode A)"
    return upd / sec, sec, (f"unmodified reference (ref_harness), synthetic {n}^3 cells x 512 ordinates, {what}, "
                            f"OpenMP {threads} threads, solve time only, mean of {steps} run(s) after {warmup} warm-up")


def parity_check(ctx, n, G, rank, world, emulated):
    """phi_tot of this rank against the reference digest of the same solve; all ranks must hold the same bits."""
    import zlib

    import numpy as np

    from doma_b200.digest import SYNTHETIC_FLOOR, flux_errors, within_bounds

    phi = ctx.flux()                                    # [G, N]; collective when the cell phases are sharded
    crc = zlib.crc32(phi.tobytes())
    crcs = [crc]
    if world > 1:
        import torch.distributed as dist

        crcs = [None] * world
        dist.all_gather_object(crcs, crc)
    out = {"checked": False, "n_ranks": world, "ranks_bit_identical": len(set(crcs)) == 1}
    path = os.path.join(ROOT, "tests", "golden", "full", "syn256_b.npz")
    if emulated > 1:
        out["why"] = "emulated shard: only a subset of the ordinates is swept"
    elif n != 256 or G != 8 or not os.path.exists(path):
        out["why"] = f"the committed reference digest is of the 256^3 x 8-group workload, not {n}^3 x {G}"
    elif rank == 0:
        errs = flux_errors(phi, np.load(path), 3, SYNTHETIC_FLOOR)
        out.update(checked=True, vs="digest of the unmodified reference on this workload (tests/golden/full/syn256_b.npz)",
                   max_err=errs["sample_max"], median_err=errs["sample_median"], errors=errs, tol=1e-10,
                   floor_frac_of_phi_max=SYNTHETIC_FLOOR, ok=bool(within_bounds(errs)) and out["ranks_bit_identical"])
    return out


def host_threads():
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        pass
    # the reference parallelises over the ordinates of ONE octant: 64 for S16, more threads cannot help
    return max(1, min(n, 64))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--n", type=int, default=256, help="cells per axis of the synthetic cube")
    ap.add_argument("--groups", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tile", default="", help="TYxTZ override of the KBA tile, e.g. 4x8")
    ap.add_argument("--variant", default="auto", choices=["auto", "flags", "ll"],
                    help="face hand-over protocol of the sweep kernel (doma_b200_set_sweep_variant); tuning aid")
    ap.add_argument("--ref-n", type=int, default=96, help="cells per axis of the bounded CPU sample")
    ap.add_argument("--cpu-n", type=int, default=128, help="cells per axis of the cpu_baseline sample (4 sweeps of group 0)")
    ap.add_argument("--emulate-world", type=int, default=0,
                    help="tuning aid, 1 GPU only: sweep rank 0's ordinate shard of a W-rank run (no all-reduce); "
                         "the line is marked 'emulated_world' and is not a bench value")
    ap.add_argument("--no-fused-si", action="store_true",
                    help="drive sweep / all-reduce / residual / scattering-source update as separate ABI calls (the drop-in "
                         "driver's sequence) instead of doma_b200_source_iteration")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cell-sharding", action="store_true",
                    help="N > 1: the r01 scheme (all-reduce of phi_cur, elementwise phases replicated on every rank)")
    ap.add_argument("--no-parity", action="store_true", help="skip the comparison with the reference digest after the first warm-up step")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else max(a.warmup, 1)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    G, n = a.groups, a.n

    def workload(cells_per_axis):
        return (f"synthetic 3D {cells_per_axis}^3 cells, {G} groups, S16 GC quadrature (np=na=8, 512 ordinates), TWDD, Mode B "
                f"(one sweep per group, Gauss-Seidel source build between)")

    config = {"workload": workload(n),
              "cells": n ** 3, "ordinates": 512, "groups": G, "parallelism": f"angle-sharded x{world}" if world > 1 else "1 GPU",
              "l2": "working set (face planes + flux arrays, > 1 GB) exceeds the 126 MB L2; no explicit flush",
              # identical in both arms, so that the two config objects compare equal AND say what the CPU arm times
              "reference_arm_sample": f"--impl reference times the unmodified reference on a bounded sample of this workload: "
                                      f"{a.ref_n}^3 cells of the same recipe (same materials, quadrature, Mode B step over all "
                                      f"{G} groups); its rate is flat in the mesh size (DESIGN.md 6)"}

    if a.impl == "reference":
        if rank != 0:
            return 0
        T = host_threads()
        v, sec, desc = run_reference_sample(a.ref_n, G, a.steps, a.warmup, T)
        if v is None:
            print(json.dumps({"impl": "reference", "unavailable": desc}))
            return 0
        # the line names the mesh that was actually timed: the metric is a rate (updates/s), the reference's rate is flat
        # in the mesh size (3.7e8 at 64^3, 3.5e8 at 128^3 on 16 threads, 7.2e7 at 256^3 on this container's 8 vCPU), and
        # a 256^3 Mode-B step costs it ~3 minutes and 14 GB on 16 threads - outside a K-step arm's budget
        timed = {"cells": a.ref_n ** 3, "workload": workload(a.ref_n), "updates_per_step": float(a.ref_n) ** 3 * 512 * G}
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": config, "timed": timed,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": T, "kind": "reference", "sample": desc,
                             "note": "the same unmodified reference on the full 256^3 x 8 groups x 512 ordinates step: 953 s on 8 vCPU "
                                     "(7.2e7 updates/s, tests/golden/full/syn256_b.npz ref_seconds)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return 0

    import numpy as np
    import torch
    import torch.distributed as dist

    from doma_b200 import abi
    from doma_b200.sharding import make_allreduce, shard_ordinates
    from doma_b200.synthetic import synthetic_case

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    xml = os.path.join(tempfile.gettempdir(), f"doma_bench_{n}_{G}_r{rank}.xml")
    params, prob = synthetic_case(xml, n, G=G, mode="B")
    prob.device = local
    full_ord = prob.n_ord
    if world > 1:
        prob = prob.with_ordinates(shard_ordinates(prob, rank, world))
    elif a.emulate_world > 1:
        prob = prob.with_ordinates(shard_ordinates(prob, 0, a.emulate_world))
        full_ord = prob.n_ord
    ctx = abi.Context(prob)
    stream = torch.cuda.Stream(device=local)
    ctx.set_stream(stream.cuda_stream)      # the context launches on this stream, so torch events see its kernels
    ctx.set_sweep_variant({"auto": abi.SWEEP_AUTO, "flags": abi.SWEEP_FLAGS, "ll": abi.SWEEP_LL}[a.variant])
    if a.tile:
        ty, tz = [int(v) for v in a.tile.lower().split("x")]
        ctx.set_tile(8, ty, tz)
    allreduce = make_allreduce(ctx) if world > 1 else None
    if world > 1 and a.no_cell_sharding:
        ctx.set_cell_sharding(False)
    N = prob.n_cells
    updates_per_step = float(N) * full_ord * G

    fused = not a.no_fused_si

    def one_group(g):
        # Mode B of the synthetic case: fs_tol = 1e30, fs_max_it = 2 -> sourceIteration does exactly one sweep
        ctx.update_mg_source(g, 0.0)
        if fused:
            it, _, _ = ctx.source_iteration(g, 1e30, 2)     # sweep -> NCCL all-reduce -> residual + update, in the library
            assert it == 1
        else:
            ctx.sweep(g)
            if allreduce is not None:
                allreduce(ctx)
            ctx.scatter_residual(g)      # synchronises the stream
            ctx.update_scatter_source(g)

    def step():
        for g in range(G):
            one_group(g)
        return ctx.mg_residual()

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    ctx.reset_state(0.0)
    parity = None
    for w_i in range(a.warmup):
        step()
        if w_i == 0 and not a.no_parity:
            # after ONE step from the zero state the context holds exactly the Mode-B solve whose digest the unmodified
            # reference wrote (tests/golden/full/syn256_b.npz): compare, outside the timed region, at every rank count
            parity = parity_check(ctx, n, G, rank, world, a.emulate_world)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ctx.info(abi.INFO_KERNEL_LAUNCHES)
    sweep_ms = []
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record(stream)
    # device time: events on the context's stream bracket every sweep (last_sweep_ms); the whole
    # region is bracketed by stream synchronisation + barrier, so wall == device time here
    for _ in range(a.steps):
        for g in range(G):
            one_group(g)
            sweep_ms.append(ctx.last_sweep_ms())
        ctx.mg_residual()
    ev1.record(stream)
    barrier()
    wall = time.perf_counter() - t0
    dt = ev0.elapsed_time(ev1) * 1e-3       # device time of the K steps (CUDA events on the launching stream)
    launches = ctx.info(abi.INFO_KERNEL_LAUNCHES) - l0
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        sm = torch.tensor([sum(sweep_ms) / len(sweep_ms)], dtype=torch.float64, device="cuda")
        dist.all_reduce(sm, op=dist.ReduceOp.MAX)
        mean_sweep_ms = float(sm.item())
    else:
        mean_sweep_ms = sum(sweep_ms) / len(sweep_ms)
    value = updates_per_step * a.steps / dt

    # ---- end to end through the host-buffer ABI call (q in pinned host memory, phi back to host) ----
    e2e = None
    if not a.no_e2e:
        qh = torch.empty((G, N), dtype=torch.float64).pin_memory()
        ph = torch.empty((G, N), dtype=torch.float64).pin_memory()
        qh.copy_(torch.rand((G, N), dtype=torch.float64, generator=torch.Generator().manual_seed(1)))
        qn, pn = qh.numpy(), ph.numpy()

        staged = [False]

        def e2e_step():
            # the same Gauss-Seidel pass as step(), with every sweep driven through the host-buffer calls: q of the
            # group goes pinned host -> device, the (all-reduced) phi_cur comes back device -> host.  The calls are the
            # asynchronous pair, so the copies of one group run under the kernels of its neighbours; the step ends
            # with host_wait(), i.e. every byte has landed inside the timed region.
            # (the source of the next sweep - of the next step's first group after the last one - is staged as soon
            # as the current sweep is enqueued: every step still uploads its G sources once)
            if not staged[0]:
                ctx.stage_q_from_host(qn[0])
            for g in range(G):
                ctx.update_mg_source(g, 0.0)
                ctx.sweep_from_host_async(g, None, pn[g])
                ctx.stage_q_from_host(qn[(g + 1) % G])
                staged[0] = True
                ctx.scatter_residual(g)
                ctx.update_scatter_source(g)
            r = ctx.mg_residual()
            ctx.host_wait()
            return r + float(pn[G - 1][ctx.info(abi.INFO_CELL_LO)])

        for _ in range(2):
            e2e_step()
        barrier()
        ev0.record(stream)
        for _ in range(a.steps):
            e2e_step()
        ev1.record(stream)
        barrier()
        de = ev0.elapsed_time(ev1) * 1e-3
        if world > 1:
            t = torch.tensor([de], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            de = float(t.item())
        # bytes over PCIe per step, all ranks together: cell-sharded ranks move their own cells only (1x the arrays);
        # replicated ranks each move whole arrays (world x)
        copies = 1 if (world == 1 or not a.no_cell_sharding) else world
        e2e = {"value": updates_per_step * a.steps / de, "unit": UNIT, "h2d_bytes_per_step": 8 * N * G * copies,
               "d2h_bytes_per_step": 8 * N * G * copies, "ms_per_step": de / a.steps * 1e3,
               "api": "doma_b200_stage_q_from_host + doma_b200_sweep_from_host_async + doma_b200_host_wait (C ABI, pinned host buffers, copies overlapped with the neighbouring groups' kernels) per group + the step's source/residual calls"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peaks()
    upd_per_launch = float(N) * prob.n_ord
    achieved = upd_per_launch * BYTES_PER_UPDATE / (mean_sweep_ms * 1e-3) / 1e9
    kernel_rate = upd_per_launch / (mean_sweep_ms * 1e-3)
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    n_sms = torch.cuda.get_device_properties(local).multi_processor_count
    prof = {}
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "sweep_traffic.json")))
    except Exception:
        pass
    per_n = (prof.get("per_n_gpus") or {}).get(str(a.emulate_world if a.emulate_world > 1 else world)) or {}
    fp64_inst = float(prof.get("fp64_inst_per_update") or FP64_INST_PER_UPDATE)
    # The roof that binds is FP64 issue: lane-instructions of the FP64 pipe per second against
    # 64 lanes per SM per clock (an FMA counts once).  The HBM model north_star names (48 B per update) is kept beside it.
    fp64_achieved = kernel_rate * fp64_inst
    fp64_peak = FP64_WARP_INST_PER_CLK_SM * 32.0 * n_sms * sm_mhz * 1e6
    roofline = {"bound": "fp64_issue", "kernel": "sweep_kernel<3,TWDD>", "achieved": fp64_achieved / 1e12, "peak": fp64_peak / 1e12,
                "unit": "T lane-instructions/s of the FP64 pipe (an FMA counts once)", "frac": fp64_achieved / fp64_peak,
                "fp64_pipe_frac": fp64_achieved / fp64_peak, "fp64_inst_per_update": fp64_inst,
                "fp64_inst_source": prof.get("fp64_inst_source", "SASS count of the r01 kernel"),
                "peak_source": f"{n_sms} SMs x 64 FP64 lanes x {sm_mhz:.0f} MHz (scripts/fp64_peak.cu measures 0.99 of it with DADD)",
                "hbm_model_achieved": achieved, "hbm_model_peak": peak, "hbm_model_unit": "GB/s", "hbm_model_frac": achieved / peak,
                "hbm_model_peak_source": peak_src, "algorithmic_bytes_per_update": BYTES_PER_UPDATE,
                "traffic": per_n.get("dram_bytes_per_launch"), "traffic_source": per_n.get("source"),
                "updates_per_launch": upd_per_launch, "launch_ms": mean_sweep_ms, "kernel_updates_per_s": kernel_rate,
                "fp64_gflops_algorithmic": upd_per_launch * FP64_OPS_PER_UPDATE / (mean_sweep_ms * 1e-3) / 1e9,
                "sweep_share_of_step": mean_sweep_ms * G * a.steps / (dt * 1e3)}
    cpu = None
    if not a.no_cpu_baseline and world == 1:
        T = host_threads()
        # bounded sample: 4 sweeps of group 0 at 128^3 = 4.3e9 updates, 10-15 s of work on 16 host threads
        v, sec, desc = run_reference_sample(a.cpu_n, G, 1, 1, T, mode="A", sweeps=4)
        if v is not None:
            cpu = {"value": v, "unit": UNIT, "cores": T, "kind": "reference", "sample": desc}
        else:
            cpu = {"value": None, "unit": UNIT, "cores": T, "kind": "reference", "sample": desc}
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
           "ms_per_step": dt / a.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic", "config": config, "grind_ns": 1e9 / value, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
           "gpu_launches": int(launches), "clocks": clocks, "parity": parity, "wall_ms_per_step": wall / a.steps * 1e3,
           "tile": "%dx%dx%d%s" % (ctx.info(abi.INFO_TILE_X), ctx.info(abi.INFO_TILE_Y), ctx.info(abi.INFO_TILE_Z), "" if a.tile else " (auto)"), "sweep_variant": {abi.SWEEP_FLAGS: "flags", abi.SWEEP_LL: "ll"}[int(ctx.info(abi.INFO_SWEEP_VARIANT))] + ("" if a.variant != "auto" else " (auto)"),
           "units_per_rank": int(ctx.info(abi.INFO_N_UNITS)),
           "inner_loop": "doma_b200_source_iteration (library-owned)" if fused else "separate ABI calls",
           "exchange": ("none (1 GPU)" if world == 1 else
                        "all-reduce(phi_cur), elementwise phases replicated" if a.no_cell_sharding else
                        "all-gather(q) -> sweep -> reduce-scatter(phi_cur), elementwise phases on the rank's own cells")}
    if a.emulate_world > 1:
        out["emulated_world"] = a.emulate_world
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
