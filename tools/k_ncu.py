#!/usr/bin/env python
"""Developer probe for ncu: a few persistent-sweep launches at one size (NSB_SOR_K from the environment)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
m = importlib.import_module("navierstokes-parallel_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
its = int(sys.argv[2]) if len(sys.argv) > 2 else 20
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 0
P = m.make_params(i_max=n, j_max=n, T=1e9, max_it=4)
s = m.Solver(P, device=0, sor_engine=m.SOR_STREAM, chunk_rows=chunk)
s.step(0.0)
s.upload(m.P, np.random.default_rng(5).uniform(-1.0, 1.0, s.shape))
for _ in range(3):
    s.sor_iterations(its)
s.close()
