#!/usr/bin/env python
"""Developer probe: persistent sweep, microseconds per SOR iteration against the chunk height at mid-size grids."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
m = importlib.import_module("navierstokes-parallel_b200")
div = int(sys.argv[3]) if len(sys.argv) > 3 else 0
sizes = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1536, 2048, 3072, 4096, 6144]
chunks = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0, 16, 24, 32, 48, 64, 96, 120, 240]
for n in sizes:
    for chunk in chunks:
        P = m.make_params(i_max=n, j_max=n, T=1e9, max_it=100)
        s = m.Solver(P, device=0, sor_engine=m.SOR_STREAM, chunk_rows=chunk, div_mode=div)
        st = torch.cuda.ExternalStream(s.stream)
        s.step(0.0)
        if os.environ.get("PROBE_FILL", "1") == "1":      # a developed field: no exact zeros in p
            rng = np.random.default_rng(5)
            s.upload(m.P, rng.uniform(-1.0, 1.0, s.shape))
        s.sor_iterations(20)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st); s.sor_iterations(400); e1.record(st); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1000.0 / 400.0
        print(f"n {n:5d} chunk {chunk:4d}  {us:8.2f} us/iteration  {24.0 * n * n / us / 1e3:7.1f} GB/s  {n * n / us / 1e3:7.1f} GLUPS", flush=True)
        s.close()
