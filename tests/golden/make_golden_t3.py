#!/usr/bin/env python
"""Digests of the LONG runs of the unmodified reference (SURVEY.md 8c tier T3), started in the background at the
beginning of round 2 with the shipped budgets unchanged:

    oracle/_ref/ref_harness /root/reference/cases/3D_assembly_eig.xml 4 /tmp/t3/asm_eig --snapshots > /tmp/t3/asm_eig.log
    oracle/_ref/ref_harness /root/reference/cases/3D_assembly_fs.xml  2 /tmp/t3/asm_fs  --snapshots > /tmp/t3/asm_fs.log
    oracle/_ref/ref_harness tr2d_20steps.xml        1 /tmp/t3/tr2d_20steps        > ...   (shipped file, t1=0.2, t_steps=20: same dt)
    oracle/_ref/ref_harness tr2d_20steps_remove.xml 1 /tmp/t3/tr2d_20steps_remove > ...   (+ remove_at=0.105, inside the window)

The assembly cases need tens of hours on this container's 8 vCPU (fs_tol = mg_tol = 1e-4, ~350 sweeps per outer, slowly
converging outers), so the harness's --snapshots mode writes the full-precision state after EVERY outer iteration
(oracle/ref_harness.cpp: a filtering streambuf sees the reference's own "Performing PI n" line; nothing of the
reference is edited or restated).  The state after n outers of the unbounded run is, by construction, the final state
of the same file with mg_max_it="n" - that derived file is what the fixture's XML is, and what the GPU tests run:
real source-iteration budgets (fs_tol 1e-4, up to 163 sweeps per group), real Gauss-Seidel / power iterations, k after
n power iterations to 17 digits.  Re-run this script to pick up however many outers have finished.

    python tests/golden/make_golden_t3.py [/tmp/t3]
"""
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import REF, parse_trace, sub  # noqa: E402
from make_golden_full import N_SAMPLE, digest  # noqa: E402

T3 = sys.argv[1] if len(sys.argv) > 1 else "/tmp/t3"
FULL = os.path.join(HERE, "full")


def write_case(name, xml):
    with open(os.path.join(HERE, "cases", "full_" + name + ".xml"), "w") as f:
        f.write("<!-- shipped DOMA case (iteration budget possibly changed) written by tests/golden/make_golden_t3.py -->\n" + xml)


def transient(name, xml_path, run=None):
    """Digest of a FINISHED run (directory T3/<run>, log T3/<run>.log) as fixture <name> with the XML it ran."""
    run = run or name
    out_dir = os.path.join(T3, run)
    if not os.path.exists(os.path.join(out_dir, "meta.json")):
        print(name, "not finished")
        return
    d = digest(out_dir, open(os.path.join(T3, run + ".log")).read())
    write_case(name, open(xml_path).read())
    np.savez_compressed(os.path.join(FULL, name + ".npz"), **d)
    print(f"{name}: ok={d['ok']} k={d['k']!r} sweeps={int(d['si_counts'].sum())} outers={len(d['outer_residuals'])} ref_seconds={d['ref_seconds']:.0f}")


def snapshots(name, shipped, shape, groups, n_ord, threads, keep):
    """One digest per kept outer count n: the state after n outers of the unbounded run."""
    out_dir = os.path.join(T3, name)
    if not os.path.isdir(out_dir):
        print(name, "no run directory")
        return
    log = open(os.path.join(T3, name + ".log")).read()
    done = sorted(int(d[5:]) for d in os.listdir(out_dir) if d.startswith("snap_") and os.path.exists(os.path.join(out_dir, d, "meta.json")))
    if not done:
        print(name, "no snapshot yet")
        return
    nx, ny, nz = shape
    N = nx * ny * nz
    idx = np.sort(np.random.default_rng(20261104).choice(N, size=min(N_SAMPLE, N), replace=False))
    for n in sorted(set([k for k in keep if k in done] + [done[-1]])):
        snap = os.path.join(out_dir, f"snap_{n}")
        meta = json.load(open(os.path.join(snap, "meta.json")))
        phi = np.stack([np.fromfile(os.path.join(snap, f"phi_g{g}.f64")) for g in range(groups)])
        # the reference's own output up to (and including) the line that announces outer n+1 = state after n outers
        m = re.search(r"^Performing (?:PI|MGI) %d\b.*$" % n, log, flags=re.M)
        head = log[:m.end()]
        nxt = re.match(r"\nk = .*", log[m.end():])          # eigenvalue runs print k on the next line
        if nxt:
            head += nxt.group(0)
        tr = parse_trace(head)
        assert len(tr["outer_residuals"]) == n, (len(tr["outer_residuals"]), n)
        slabs = phi.reshape(groups, nz, ny * nx).sum(axis=2)
        d = dict(idx=idx, phi_s=phi[:, idx], phi_sum=phi.sum(axis=1), phi_l2=np.sqrt((phi * phi).sum(axis=1)), phi_min=phi.min(axis=1),
                 phi_max=phi.max(axis=1), slab_sum=slabs, dnp_s=np.zeros((0, idx.size)), k=meta["k"], k_prev=meta["k_prev"], ok=False, dims=3,
                 shape=np.array([nx, ny, nz]), n_ord=n_ord, ref_seconds=float("nan"), ref_threads=threads, outers=n, **tr)
        fixture = f"{name}_t3_{n}outer"
        np.savez_compressed(os.path.join(FULL, fixture + ".npz"), **d)
        write_case(fixture, sub(open(os.path.join(REF, "cases", shipped)).read(), mg_max_it=str(n)))
        print(f"{fixture}: k={meta['k']!r} sweeps={int(tr['si_counts'].sum())} si_counts[-3:]={tr['si_counts'][-3:].tolist()} "
              f"outer_residual={tr['outer_residuals'][-1]!r} (snapshots available: {done[0]}..{done[-1]})")


def main():
    transient("tr2d_20steps", os.path.join(T3, "tr2d_20steps.xml"))
    transient("tr2d_20steps_remove", os.path.join(T3, "tr2d_20steps_remove.xml"))
    snapshots("asm_eig", "3D_assembly_eig.xml", (152, 152, 38), 3, 128, 4, keep=[2])
    # the UNMODIFIED shipped eigenvalue case run to its own convergence criterion (k_tol 1e-5): north_star's "k_eff within 1e-9"
    transient("asm_eig_shipped", os.path.join(REF, "cases", "3D_assembly_eig.xml"), run="asm_eig")
    snapshots("asm_fs", "3D_assembly_fs.xml", (152, 152, 37), 3, 128, 2, keep=[1])


if __name__ == "__main__":
    main()
