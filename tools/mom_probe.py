#!/usr/bin/env python
"""Developer probe: the momentum kernels alone (F and G, RHS, velocity update + reductions), us per launch, CUDA events
on the context's stream.  usage: tools/mom_probe.py [sizes] ; bytes: F,G 40 B per cell, RHS 24, update 40."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
m = importlib.import_module("navierstokes-parallel_b200")
sizes = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [8192, 8191]
for n in sizes:
    P = m.make_params(i_max=n, j_max=n, T=1e9, max_it=4)
    with m.Solver(P, device=0) as s:
        st = torch.cuda.ExternalStream(s.stream)
        rng = np.random.default_rng(7)
        s.upload(m.U, rng.uniform(-1, 1, s.shape)); s.upload(m.V, rng.uniform(-1, 1, s.shape))
        s.upload(m.P, rng.uniform(-1, 1, s.shape))
        for name, fn, byts in (("fg", lambda: s.FG(1e-3, 0.7), 40), ("rhs", lambda: s.RHS(1e-3), 24),
                               ("update_uv", lambda: s.update_uv(1e-9), 40)):
            fn(); fn()
            best = 1e30
            for rep in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(10):
                    fn()
                e1.record(st); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1) * 100.0)
            print(f"n {n:5d} {name:10s} {best:8.1f} us  {byts * n * n / best / 1e3:7.1f} GB/s", flush=True)
