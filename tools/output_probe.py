#!/usr/bin/env python
"""Developer probe: field dumps in the reference's text format at 8192^2.
   - time of a synchronous dump (3 files, ~1.7 GB of text);
   - time steps with an overlapped dump in flight against time steps without."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
m = importlib.import_module("navierstokes-parallel_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
out = sys.argv[2] if len(sys.argv) > 2 else "/tmp/nsb_out"
os.makedirs(out, exist_ok=True)
P = m.make_params(i_max=n, j_max=n, T=1e9, max_it=100)
s = m.Solver(P, device=0)
t = 0.0
for _ in range(2):
    t += s.step(t)["dt"]
s.upload(m.P, np.random.default_rng(1).uniform(-1, 1, s.shape))
def steps(k):
    global t
    t0 = time.perf_counter()
    for _ in range(k):
        t += s.step(t)["dt"]
    return (time.perf_counter() - t0) / k * 1e3
s.output(t, os.path.join(out, "warm"))                 # allocates the staging buffers
base = steps(6)
t0 = time.perf_counter(); s.output(t, os.path.join(out, "sync")); sync_s = time.perf_counter() - t0
size = sum(os.path.getsize(os.path.join(out, "sync" + x)) for x in ("_u.txt", "_v.txt", "_p.txt"))
t0 = time.perf_counter(); s.output(t, os.path.join(out, "async"), wait=False); enq_ms = (time.perf_counter() - t0) * 1e3
during = steps(6)
t0 = time.perf_counter(); s.output_wait(); tail_s = time.perf_counter() - t0
print(f"grid {n}x{n}: synchronous dump {sync_s:.2f} s for {size / 1e9:.2f} GB of text ({3 * (n + 2) ** 2 / sync_s / 1e6:.0f} M values/s); "
      f"async enqueue {enq_ms:.1f} ms; time step {base:.1f} ms alone, {during:.1f} ms with a dump in flight "
      f"({(during / base - 1) * 100:+.1f} %); writer finished {tail_s:.2f} s after those 6 steps; host cores {os.cpu_count()}")
s.close()
