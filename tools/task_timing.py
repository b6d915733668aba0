#!/usr/bin/env python
"""Developer probe (needs a library built with -DSOR_TASK_TIMING, see tools/build_variants.sh): where a task of the
persistent sweep spends its time.  usage: NSB200_LIB=build/variants/lib_tt.so tools/task_timing.py ni,nj,k,chunk ..."""
import ctypes as C, importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
m = importlib.import_module("navierstokes-parallel_b200")
L = m.lib()
names = ["claim", "ctl word", "3x3 flags", "setup", "march", "publish", "close", "tasks"]
for spec in sys.argv[1:]:
    ni, nj, k, chunk = (int(x) for x in spec.split(","))
    P = m.make_params(i_max=ni, j_max=nj, a=1.0, b=float(nj) / ni, T=1e9, max_it=100)
    s = m.Solver(P, device=0, sor_engine=m.SOR_STREAM, chunk_rows=chunk, pass_k=k)
    plan = s.sor_plan()
    st = torch.cuda.ExternalStream(s.stream)
    s.step(0.0)
    s.upload(m.P, np.random.default_rng(5).uniform(-1.0, 1.0, s.shape))
    s.sor_iterations(20)
    out = (C.c_ulonglong * 16)()
    L.nsb_debug_task_timing(out)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_it = 200
    e0.record(st); s.sor_iterations(n_it); e1.record(st); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1000.0 / n_it
    L.nsb_debug_task_timing(out)
    n = max(1, out[7])
    kk = plan["pass_k"]
    rows = plan["chunk_rows"]
    steps = rows + (11 if kk == 2 else 6)
    print(f"{ni} x {nj} plan {plan}: {us:.1f} us/iteration, {n} tasks in {n_it} iterations, per task (us): " +
          ", ".join(f"{names[i]} {out[i] / n / 1e3:.2f}" for i in range(6)) + f", close per pass {out[6] / 1e3 / max(1, n_it // kk):.1f}" +
          f" | march per step {out[4] / n / 1e3 / steps:.3f} us ({steps} steps)" +
          " | march by kind (us): " + ", ".join(f"{nm} {out[8 + i] / max(1, out[11 + i]) / 1e3:.2f} (x{out[11 + i] // max(1, n_it // kk)})" for i, nm in enumerate(("interior", "wall chunk", "edge strip"))), flush=True)
    s.close()
