"""Developer diagnostic: classify wrong cells of ONE streaming SOR iteration (host numpy restatement)."""
import importlib, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
m = importlib.import_module("navierstokes-parallel_b200")
ni, nj = int(sys.argv[1]), int(sys.argv[2])
reps = int(os.environ.get("REPS", "30"))
rng = np.random.default_rng(31)
p0 = rng.uniform(-1, 1, (ni + 2, nj + 2)); r0 = rng.uniform(-1, 1, (ni + 2, nj + 2))
om = 1.7; dx = 1.0 / ni; dy = 1.0 / nj; dx2 = dx * dx; dy2 = dy * dy
w = om / (2.0 * (1.0 / dx2 + 1.0 / dy2)); om1 = 1.0 - om
def ghosts(p):
    p[0, :] = p[1, :]; p[-1, :] = p[-2, :]; p[:, 0] = p[:, 1]; p[:, -1] = p[:, -2]
def sweep(p, c):
    ghosts(p)
    I, J = np.meshgrid(np.arange(1, ni + 1), np.arange(1, nj + 1), indexing="ij")
    mask = ((I + J) % 2 == c)
    new = om1 * p[1:-1, 1:-1] + w * (((p[2:, 1:-1] + p[:-2, 1:-1]) / dx2 + (p[1:-1, 2:] + p[1:-1, :-2]) / dy2) - r0[1:-1, 1:-1])
    q = p.copy(); q[1:-1, 1:-1] = np.where(mask, new, p[1:-1, 1:-1]); return q
pr = sweep(p0.copy(), 0); pf = sweep(pr.copy(), 1)
P = m.make_params(i_max=ni, j_max=nj, a=1.0, b=1.0, T=1.0, Re=1000.0, omega=om, max_it=100, epsilon=1e-4)
bad = 0
for k in range(reps):
    s = m.Solver(P, device=0, sor_engine=m.SOR_STREAM)
    s.upload(m.P, p0); s.upload(m.RHS, r0)
    s.sor_iterations(1)
    a = s.download(m.P); s.close()
    d = a[1:-1, 1:-1] != pf[1:-1, 1:-1]
    if k == 0: print("first run mismatches vs numpy:", int(d.sum()))
    if d.any():
        bad += 1
        ii, jj = np.nonzero(d)
        print(f"run {k}: {len(ii)} wrong cells")
        for i, j in list(zip(ii + 1, jj + 1))[:12]:
            v = a[i, j]
            tag = "==p0" if v == p0[i, j] else ("==p_red" if v == pr[i, j] else "other")
            print(f"   (i={i}, j={j}) colour {(i+j)%2} got {v:+.6e} want {pf[i,j]:+.6e} p0 {p0[i,j]:+.6e} p_red {pr[i,j]:+.6e} {tag}")
print("bad runs", bad, "of", reps)
