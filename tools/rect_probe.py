#!/usr/bin/env python
"""Developer probe: persistent sweep on a rectangular grid (the shape of one y-slab), us per SOR iteration.
usage: tools/rect_probe.py ni nj [k,chunk ...]   (k = 0 / chunk = 0: the planner's choice)"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
m = importlib.import_module("navierstokes-parallel_b200")
ni, nj = int(sys.argv[1]), int(sys.argv[2])
cases = [tuple(int(x) for x in a.split(",")) for a in sys.argv[3:]] or [(0, 0)]
for k, chunk in cases:
    P = m.make_params(i_max=ni, j_max=nj, a=1.0, b=float(nj) / ni, T=1e9, max_it=100)
    s = m.Solver(P, device=0, sor_engine=m.SOR_STREAM, chunk_rows=chunk, pass_k=k)
    plan = s.sor_plan()
    st = torch.cuda.ExternalStream(s.stream)
    s.step(0.0)
    s.upload(m.P, np.random.default_rng(5).uniform(-1.0, 1.0, s.shape))
    s.sor_iterations(20)
    best = 1e30
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st); s.sor_iterations(200); e1.record(st); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1000.0 / 200.0)
    print(f"{ni} x {nj}  asked k {k} chunk {chunk}  plan {plan}  {best:8.2f} us/iteration  {ni * nj / best / 1e3:7.1f} GLUPS", flush=True)
    s.close()
