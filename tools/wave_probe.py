#!/usr/bin/env python
"""Developer probe: SOR launch time against the number of block rounds (i_max = 8192 fixed, j_max = 120 * k)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
m = importlib.import_module("navierstokes-parallel_b200")
chunk = int(sys.argv[1]) if len(sys.argv) > 1 else 120
ks = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [13, 25, 26, 38, 51, 52, 64, 69, 77, 78, 103]
for k in ks:
    nj = chunk * k
    P = m.make_params(i_max=8192, j_max=nj, a=1.0, b=nj / 8192.0, T=1e9, max_it=100)
    s = m.Solver(P, device=0, sor_engine=m.SOR_STREAM, chunk_rows=chunk)
    st = torch.cuda.ExternalStream(s.stream)
    s.step(0.0)
    s.sor_iterations(10)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st); s.sor_iterations(100); e1.record(st); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 10.0
    tasks = 69 * k
    print(f"chunk {chunk} k {k:4d} nj {nj:6d} tasks {tasks:6d} rounds {tasks / 1776.0:5.2f}  {us:8.1f} us/launch  {us / k:6.3f} us/chunk-row-band  "
          f"{24.0 * 8192 * nj / us / 1e3:7.1f} GB/s", flush=True)
    s.close()
