#!/usr/bin/env python
"""Host-side timeline of one end-to-end step (asynchronous host-buffer calls): which call blocks for how long."""
import os, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from doma_b200 import abi
from doma_b200.synthetic import synthetic_case
n, G = int(os.environ.get("N", 256)), 8
params, prob = synthetic_case(os.path.join(tempfile.gettempdir(), "probe.xml"), n, G=G, mode="B")
ctx = abi.Context(prob); N = prob.n_cells
qh = torch.rand((G, N), dtype=torch.float64).pin_memory(); ph = torch.empty((G, N), dtype=torch.float64).pin_memory()
qn, pn = qh.numpy(), ph.numpy()
ctx.reset_state(0.0)
def step(log=None, mode="async"):
    t0 = time.perf_counter(); marks = []
    def mark(name): marks.append((name, (time.perf_counter() - t0) * 1e3))
    if mode == "async": ctx.stage_q_from_host(qn[0]); mark("stage0")
    sw = []
    for g in range(G):
        ctx.update_mg_source(g, 0.0); mark(f"mg{g}")
        if mode == "async":
            ctx.sweep_from_host_async(g, None, pn[g]); mark(f"async{g}")
            if g + 1 < G: ctx.stage_q_from_host(qn[g + 1]); mark(f"stage{g+1}")
        elif mode == "sync":
            ctx.sweep_from_host(g, qn[g], pn[g]); mark(f"sync{g}")
        else:
            ctx.sweep(g); mark(f"sweep{g}")
        ctx.scatter_residual(g); mark(f"res{g}"); sw.append(ctx.last_sweep_ms())
        ctx.update_scatter_source(g); mark(f"upd{g}")
    ctx.mg_residual(); mark("mgres")
    if mode == "async": ctx.host_wait(); mark("wait")
    if log: print(mode, " ".join(f"{k}:{v:.1f}" for k, v in marks)); print("sweep ms", " ".join(f"{v:.2f}" for v in sw))
    return (time.perf_counter() - t0) * 1e3
for mode in ("resident", "sync", "async"):
    for _ in range(3): step(mode=mode)
    ts = [step(mode=mode) for _ in range(3)]
    print(mode, "ms per step", " ".join(f"{t:.1f}" for t in ts))
    step(log=True, mode=mode)
