#!/usr/bin/env python
"""Developer probe: time per SOR iteration on small / mid-size grids (latency-bound regime)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
m = importlib.import_module("navierstokes-parallel_b200")
eng = int(sys.argv[1]) if len(sys.argv) > 1 else m.SOR_STREAM
for n in (48, 256, 512, 1024, 2048, 4096, 8192):
    P = m.make_params(i_max=n, j_max=n, T=1e9, max_it=100)
    s = m.Solver(P, device=0, sor_engine=eng)
    st = torch.cuda.ExternalStream(s.stream)
    s.step(0.0); s.sor_iterations(10)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st); s.sor_iterations(200); e1.record(st); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1000 / 200
    print(f"engine {eng} n {n:5d}: {us:9.2f} us/iteration  {n * n / us / 1e3:9.1f} GLUPS", flush=True)
    s.close()
