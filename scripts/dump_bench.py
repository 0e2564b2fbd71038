#!/usr/bin/env python
"""Output path: time to write the flux text files of the 256^3 x 8-group case with
doma_b200_dump_text, beside the reference's own dumpToTextFile timed on a 64^3 sample through ref_harness --text."""
import os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from doma_b200 import abi
from doma_b200.synthetic import synthetic_case, synthetic_xml

n = int(os.environ.get("N", 256))
tmp = tempfile.mkdtemp(prefix="doma_dump_")
params, prob = synthetic_case(os.path.join(tmp, "c.xml"), n, G=8, mode="B")
ctx = abi.Context(prob)
ctx.reset_state(0.0)
for g in range(8):
    ctx.update_mg_source(g, 0.0); ctx.source_iteration(g, 1e30, 2)
ctx.synchronize()
for threads in (0, 1):
    t0 = time.perf_counter(); ctx.dump_text(os.path.join(tmp, "gpu"), False, threads); dt = time.perf_counter() - t0
    size = sum(os.path.getsize(os.path.join(tmp, f)) for f in os.listdir(tmp) if f.startswith("gpu_g"))
    print(f"doma_b200_dump_text {n}^3 x 8 groups: {dt:.2f} s, {size / 1e6:.0f} MB, {8 * n ** 3 / dt / 1e6:.1f} M values/s, threads {threads or os.cpu_count()}")
def step():
    for g in range(8):
        ctx.update_mg_source(g, 0.0); ctx.source_iteration(g, 1e30, 2)
    ctx.mg_residual()
t0 = time.perf_counter(); step(); ctx.synchronize(); t_step = time.perf_counter() - t0
t0 = time.perf_counter(); ctx.dump_text_async(os.path.join(tmp, "gpu"), False, 0); t_call = time.perf_counter() - t0
k = 0
while time.perf_counter() - t0 < 1.8:
    step(); k += 1
ctx.synchronize(); t_steps = time.perf_counter() - t0 - t_call
t1 = time.perf_counter(); ctx.dump_wait(); t_wait = time.perf_counter() - t1
print(f"doma_b200_dump_text_async: call returns after {t_call:.2f} s (host snapshot), {k} steps of {t_step * 1e3:.0f} ms ran under the formatting "
      f"({t_steps / max(k, 1) * 1e3:.0f} ms each), dump_wait then took {t_wait:.2f} s")
harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
if os.path.exists(harness):
    m = 64
    xml = os.path.join(tmp, "r.xml"); open(xml, "w").write(synthetic_xml(m, G=8, mode="B"))
    T = str(os.cpu_count() or 1)
    def run(extra):
        t0 = time.perf_counter(); subprocess.run([harness, xml, T, os.path.join(tmp, "o"), "--no-dump"] + extra, capture_output=True, check=True); return time.perf_counter() - t0
    a = min(run([]) for _ in range(2)); b = min(run(["--text", os.path.join(tmp, "ref")]) for _ in range(2))
    vals = 8 * m ** 3 + 4 * m ** 3   # 8 flux files + blocks and three mesh files
    print(f"reference dumpToTextFile {m}^3 x 8 groups (+ geometry files): {b - a:.2f} s for {vals / 1e6:.1f} M values = {vals / (b - a) / 1e6:.2f} M values/s")
import shutil; shutil.rmtree(tmp, ignore_errors=True)
