#!/usr/bin/env python
"""Small runs of every SOR engine, of the two-iteration pass and of the slab path for compute-sanitizer:
   NSB_SPIN_TIMEOUT_MS=600000 compute-sanitizer --tool racecheck python tools/sanitize_case.py
(the in-kernel waits have a 2 s deadline by default; under the sanitizer a kernel runs orders of magnitude slower)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
m = importlib.import_module("navierstokes-parallel_b200")
its = int(os.environ.get("SANITIZE_ITS", "12"))
for eng, k in ((m.SOR_STREAM, 1), (m.SOR_STREAM, 2), (m.SOR_STREAM_LAUNCH, 1), (m.SOR_RESIDENT, 1), (m.SOR_SIMPLE, 1), (m.SOR_SERIAL_ORDER, 1)):
    for (ni, nj) in ((130, 70), (257, 64)):
        P = m.make_params(i_max=ni, j_max=nj, T=1.0, max_it=its)
        with m.Solver(P, sor_engine=eng, chunk_rows=16, pass_k=k, spin_timeout_ms=600000) as s:
            log, _ = s.run(max_steps=2)
            print("engine", eng, "k", k, ni, nj, [x["sor_iters"] for x in log], s.center(), flush=True)
os.environ["NSB_RES_MODE"] = "grid"
P = m.make_params(i_max=200, j_max=150, T=1.0, max_it=its)
with m.Solver(P, sor_engine=m.SOR_RESIDENT, spin_timeout_ms=600000) as s:
    log, _ = s.run(max_steps=2); print("resident grid", [x["sor_iters"] for x in log], flush=True)
del os.environ["NSB_RES_MODE"]
P = m.make_params(i_max=150, j_max=96, T=1.0, max_it=its)
with m.Group(P, [0, 0, 0], chunk_rows=16, spin_timeout_ms=600000) as g:
    log, _ = g.run(max_steps=2); print("3 slabs lock step", [x["sor_iters"] for x in log], g.center(), flush=True)
print("ok")
