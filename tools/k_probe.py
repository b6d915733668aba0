#!/usr/bin/env python
"""Developer probe: persistent sweep with one and with two iterations per pass (NSB_SOR_K), us per SOR iteration."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
m = importlib.import_module("navierstokes-parallel_b200")
sizes = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [2048, 4096, 8192]
ks = sys.argv[2].split(",") if len(sys.argv) > 2 else ["1", "2"]
chunks = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0]
div = int(sys.argv[4]) if len(sys.argv) > 4 else 0
for n in sizes:
    for k in ks:
        for chunk in chunks:
            os.environ["NSB_SOR_K"] = k   # overrides nsb_options.sor_pass_k
            P = m.make_params(i_max=n, j_max=n, T=1e9, max_it=100)
            s = m.Solver(P, device=0, sor_engine=m.SOR_STREAM, chunk_rows=chunk, div_mode=div)
            st = torch.cuda.ExternalStream(s.stream)
            s.step(0.0)
            rng = np.random.default_rng(5)
            s.upload(m.P, rng.uniform(-1.0, 1.0, s.shape))
            s.sor_iterations(20)
            best = 1e30
            for rep in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st); s.sor_iterations(200); e1.record(st); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1) * 1000.0 / 200.0)
            print(f"n {n:5d} K {k} chunk {chunk:4d}  {best:8.2f} us/iteration  {24.0 * n * n / best / 1e3:7.1f} GB/s(24B/LUP)  {n * n / best / 1e3:7.1f} GLUPS", flush=True)
            s.close()
