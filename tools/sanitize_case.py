#!/usr/bin/env python
"""Small runs of every SOR engine and of the slab path for compute-sanitizer (memcheck / racecheck):
   compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
m = importlib.import_module("navierstokes-parallel_b200")
for eng in (m.SOR_STREAM, m.SOR_STREAM_LAUNCH, m.SOR_RESIDENT, m.SOR_SIMPLE, m.SOR_SERIAL_ORDER):
    for (ni, nj) in ((130, 70), (257, 64)):
        P = m.make_params(i_max=ni, j_max=nj, T=1.0, max_it=40)
        with m.Solver(P, sor_engine=eng, chunk_rows=16) as s:
            log, _ = s.run(max_steps=2)
            print("engine", eng, ni, nj, [x["sor_iters"] for x in log], s.center(), flush=True)
os.environ["NSB_RES_MODE"] = "grid"
P = m.make_params(i_max=200, j_max=150, T=1.0, max_it=40)
with m.Solver(P, sor_engine=m.SOR_RESIDENT) as s:
    log, _ = s.run(max_steps=2); print("resident grid", [x["sor_iters"] for x in log], flush=True)
del os.environ["NSB_RES_MODE"]
P = m.make_params(i_max=150, j_max=96, T=1.0, max_it=40)
with m.Group(P, [0, 0, 0], chunk_rows=16) as g:
    log, _ = g.run(max_steps=2); print("3 slabs lock step", [x["sor_iters"] for x in log], g.center(), flush=True)
u = np.zeros((P.i_max + 2, P.j_max + 2))
print("ok")
