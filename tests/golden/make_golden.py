#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE.

Needs /root/reference (so: this container only) and `make -C oracle ref`.  For each derived case it
writes the XML (tests/golden/cases/<name>.xml: the shipped case with a coarser mesh / quadrature or
a truncated iteration budget, so the reference finishes in seconds), runs oracle/_ref/ref_harness
with ONE thread (canonical summation order) and stores <name>.npz:
   phi [G,N], dnp [D,N], q [N], k, k_prev, ok, quad (oct, mu, eta, xi, w), mesh (lx, ly, lz, block),
   si_counts / si_residuals / outer_residuals / k_history parsed from the reference's own stdout
   (17 significant digits: the harness raises cout's precision, the reference never resets it).
Nothing here is read at test time except the committed .xml/.npz files.
"""
import json
import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("DOMA_REF", "/root/reference")
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def sub(text, **attrs):
    """Replace XML attribute values (first occurrence of name="...")."""
    for k, v in attrs.items():
        text, n = re.subn(r'(\b%s=")[^"]*(")' % k, r"\g<1>%s\g<2>" % v, text, count=1)
        assert n == 1, k
    return text


def sub_vals(text, tag, vals):
    text, n = re.subn(r'(<%s\s+vals=")[^"]*(")' % tag, r"\g<1>%s\g<2>" % vals, text, count=1)
    assert n == 1, tag
    return text


def derive():
    rd = lambda p: open(os.path.join(REF, "cases", p)).read()
    cases = {}
    k1 = rd("3D_kobayashi_1.xml")
    cases["k1_r0_s3"] = sub(k1, refinement="0", np="3", na="3")
    cases["k1_r0_s3_dd"] = sub(k1, refinement="0", np="3", na="3", scheme="dd")
    k3 = rd("3D_kobayashi_3.xml")
    cases["k3_r0_s2"] = sub(k3, refinement="0", np="2", na="2")
    thirteen = lambda a, b: ", ".join([str(a)] + [str(b)] * 11 + [str(a)])
    fs = rd("3D_assembly_fs.xml")
    fs = sub_vals(sub_vals(sub_vals(fs, "ix", thirteen(5, 2)), "iy", thirteen(5, 2)), "iz", "4, 1, 4")
    cases["asm_fs_small"] = sub(fs, fs_max_it="500", mg_max_it="4", mg_tol="0.0")
    eig = rd("3D_assembly_eig.xml")
    eig = sub_vals(sub_vals(sub_vals(eig, "ix", thirteen(5, 2)), "iy", thirteen(5, 2)), "iz", "9")
    cases["asm_eig_small"] = sub(eig, mg_max_it="5", mg_tol="0.0")
    cases["asm_eig_small_1sweep"] = sub(eig, fs_tol="1e30", fs_max_it="2", mg_tol="0.0", mg_max_it="3")
    tr = rd("2D_assembly_transient/2D_assembly_tr.xml")
    cases["tr2d_r0_3steps"] = sub(tr, refinement="0", t1="0.03", t_steps="3", fs_tol="1e-6", mg_tol="1e-6")
    cases["tr2d_r0_remove"] = sub(tr, refinement="0", t1="0.04", t_steps="4", fs_tol="1e-6", mg_tol="1e-6", remove_at="0.015")
    cases["test2d_r0"] = sub(rd("2D_test.xml"), refinement="0", np="4", na="4")
    cases["test2d_r0_dd"] = sub(rd("2D_test.xml"), refinement="0", np="4", na="4", scheme="dd")
    cases["asm2d_eig_small"] = sub(sub_vals(sub_vals(rd("2D_assembly_eig.xml"), "ix", thirteen(6, 2)), "iy", thirteen(6, 2)),
                                   mg_max_it="6", mg_tol="0.0")
    return cases


def parse_trace(out):
    f = r"([-+0-9.eE]+|nan|inf)"
    si = re.findall(r"SI (?:converged|failed to converge) after (\d+) iterations with a residual of %s for G(\d+)" % f, out)
    outer = re.findall(r"Performing (?:MGI|PI) \d+, (?:shape )?residual = %s" % f, out)
    last = re.findall(r"(?:MGI|PI) (?:converged|failed to converge) after \d+ iterations with a (?:flux shape )?residual of %s" % f, out)
    ks = re.findall(r"^k = %s" % f, out, flags=re.M) + re.findall(r"Converged value of k = %s" % f, out)
    return dict(si_counts=np.array([int(s[0]) for s in si], dtype=np.int64),
                si_residuals=np.array([float(s[1]) for s in si]),
                outer_residuals=np.array([float(x) for x in outer] + [float(x) for x in last]),
                k_history=np.array([float(x) for x in ks]))


def main():
    os.makedirs(os.path.join(HERE, "cases"), exist_ok=True)
    only = set(sys.argv[1:])
    for name, xml in derive().items():
        if only and name not in only:
            continue
        path = os.path.join(HERE, "cases", name + ".xml")
        with open(path, "w") as f:
            f.write("<!-- derived from the DOMA case of the same family by tests/golden/make_golden.py -->\n" + xml)
        out_dir = os.path.join("/tmp", "doma_golden", name)
        r = subprocess.run([HARNESS, path, "1", out_dir], capture_output=True, text=True, check=True)
        meta = json.load(open(os.path.join(out_dir, "meta.json")))
        G, N = meta["groups"], meta["nx"] * meta["ny"] * meta["nz"]
        ld = lambda n, dt=np.float64: np.fromfile(os.path.join(out_dir, n), dtype=dt)
        phi = np.stack([ld("phi_g%d.f64" % g) for g in range(G)])
        dn = sorted(f for f in os.listdir(out_dir) if f.startswith("dnp_d"))
        dnp = np.stack([ld(f) for f in dn]) if dn else np.zeros((0, N))
        tr = parse_trace(r.stdout)
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"), phi=phi, dnp=dnp, q=ld("q.f64"), k=meta["k"], k_prev=meta["k_prev"],
            ok=meta["ok"], dims=meta["dims"], shape=np.array([meta["nx"], meta["ny"], meta["nz"]]),
            quad_oct=ld("quad_oct.i32", np.int32), quad_mu=ld("quad_mu.f64"), quad_eta=ld("quad_eta.f64"),
            quad_xi=ld("quad_xi.f64"), quad_w=ld("quad_w.f64"), mesh_lx=ld("mesh_lx.f64"), mesh_ly=ld("mesh_ly.f64"),
            mesh_lz=ld("mesh_lz.f64") if meta["dims"] == 3 else np.zeros(0), mesh_block=ld("mesh_block.i32", np.int32), **tr)
        print(f"{name}: N={N} G={G} ord={meta['n_ord']} ok={meta['ok']} k={meta['k']!r} sweeps={int(tr['si_counts'].sum())} "
              f"ref_seconds={meta['solve_seconds']:.2f}")


if __name__ == "__main__":
    main()
