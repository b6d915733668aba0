"""Developer diagnostic: where does the streaming SOR engine differ from the launch-per-phase engine?"""
import importlib, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
m = importlib.import_module("navierstokes-parallel_b200")
ni, nj, its = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
chunk = int(sys.argv[4]) if len(sys.argv) > 4 else 0
rng = np.random.default_rng(31)
outs = []
reps = int(os.environ.get('REPS', '6'))
for eng in (m.SOR_SIMPLE,) + (m.SOR_STREAM,) * reps:
    P = m.make_params(i_max=ni, j_max=nj, a=1.0, b=1.0, T=1.0, Re=1000.0, max_it=100, epsilon=1e-4)
    s = m.Solver(P, device=0, sor_engine=eng, chunk_rows=chunk)
    rng = np.random.default_rng(31)
    p0 = rng.uniform(-1, 1, (ni + 2, nj + 2)); r0 = rng.uniform(-1, 1, (ni + 2, nj + 2))
    s.upload(m.P, p0); s.upload(m.RHS, r0)
    res = s.sor_iterations(its)
    outs.append((s.download(m.P)[1:-1, 1:-1], res))
    s.close()
b = outs[0][0]
bad = 0
for k in range(1, len(outs)):
    a = outs[k][0]
    d = a != b
    if d.any():
        bad += 1
        ii, jj = np.nonzero(d)
        print(f"run {k}: {int(d.sum())} mismatches, i {ii.min()+1}..{ii.max()+1}, j {jj.min()+1}..{jj.max()+1}, res {outs[k][1]!r} vs {outs[0][1]!r}")
print(f"{ni}x{nj} its={its} chunk={chunk}: {bad} bad runs of {len(outs)-1}")
