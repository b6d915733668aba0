"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle.

Bars (DESIGN.md "parity"):
  * F, G, RHS, u, v, p fields: BIT-EXACT against the oracle with the same sweep order (the kernels
    use non-contracting FP64 intrinsics in the reference's association order);
  * SOR iteration counts and the dt sequence: EQUAL;
  * residual norm / norm_p (order-dependent sums): relative 1e-12;
  * against the serial (lexicographic) order at tight eps: u, v and mean-free p within 1e-8
    relative (north_star's tolerance).
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

from oracle import binding as ob
from util import ROOT, nsb

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _solver(m, engine=None, chunk_rows=0, **kw):
    P = m.make_params(**kw)
    eng = m.SOR_STREAM if engine is None else engine
    return m.Solver(P, sor_engine=eng, chunk_rows=chunk_rows), P


def _orc_params(P):
    return ob.make_params(**{k: getattr(P, k) for k, _ in ob.Params._fields_})


def _rand(shape, seed):
    return np.ascontiguousarray(np.random.default_rng(seed).uniform(-1, 1, shape))


# ------------------------------------------------------------------------------- phases

@pytest.mark.parametrize("ni,nj,a,b", [(16, 16, 1.0, 1.0), (37, 23, 1.3, 0.7), (130, 5, 1.0, 1.0), (5, 250, 2.0, 1.0)])
def test_bc_fg_rhs_update_bit_exact(ni, nj, a, b):
    m = nsb()
    s, P = _solver(m, i_max=ni, j_max=nj, a=a, b=b, Re=77.0, g_x=0.05, g_y=-0.03)
    L = ob.lib()
    u, v, p = _rand(s.shape, 1), _rand(s.shape, 2), _rand(s.shape, 3)
    s.upload(m.U, u); s.upload(m.V, v); s.upload(m.P, p)
    dx, dy = a / ni, b / nj
    dt, gamma = 0.0123, 0.41
    # max_mat (signed, seeded with m[0][0])
    um, vm = s.max_mat()
    assert um == L.orc_max_mat(ni, nj, ob._ptr(u)) and vm == L.orc_max_mat(ni, nj, ob._ptr(v))
    # dt / gamma: host arithmetic in the reference's order
    dtg = s.dt_gamma(um, vm)
    odt, og = C.c_double(), C.c_double()
    L.orc_dt_gamma(C.byref(_orc_params(P)), dx, dy, um, vm, C.byref(odt), C.byref(og))
    assert dtg == (odt.value, og.value)
    # every side, non-trivial inflow
    for side in (m.LEFT, m.RIGHT, m.BOTTOM, m.TOP):
        s.set_inflow(side, 0.75, -0.25)
        L.orc_set_inflow(ni, nj, ob._ptr(u), ob._ptr(v), side, 0.75, -0.25)
    assert np.array_equal(s.download(m.U), u) and np.array_equal(s.download(m.V), v)
    # F, G  (untouched cells keep their previous content, here zero, like the reference)
    s.FG(dt, gamma)
    F, G = np.zeros_like(u), np.zeros_like(u)
    L.orc_FG(ob._ptr(F), ob._ptr(G), ob._ptr(u), ob._ptr(v), ni, nj, P.Re, P.g_x, P.g_y, dt, dx, dy, gamma)
    assert np.array_equal(s.download(m.F), F)
    assert np.array_equal(s.download(m.G), G)
    # RHS
    s.RHS(dt)
    R = np.zeros_like(u)
    L.orc_RHS(ob._ptr(R), ob._ptr(F), ob._ptr(G), ni, nj, dt, dx, dy)
    assert np.array_equal(s.download(m.RHS), R)
    # velocity update + the reductions it refreshes
    s.update_uv(dt)
    L.orc_update_uv(ob._ptr(u), ob._ptr(v), ob._ptr(F), ob._ptr(G), ob._ptr(p), ni, nj, dt, dx, dy)
    assert np.array_equal(s.download(m.U), u) and np.array_equal(s.download(m.V), v)
    um, vm = s.max_mat()
    assert um == L.orc_max_mat(ni, nj, ob._ptr(u)) and vm == L.orc_max_mat(ni, nj, ob._ptr(v))
    s.close()


@pytest.mark.parametrize("ni,nj,a,b", [(130, 9, 1.0, 1.0), (300, 301, 1.3, 0.7), (257, 257, 1.0, 1.0), (64, 2, 1.0, 0.3), (33, 1, 1.0, 1.0)])
def test_fg_row_march_and_ieee_quotient_window(ni, nj, a, b):
    """fg_kernel marches 3 rows per thread and carries the y-face fluxes from row to row (momentum.cu): row counts of
    every remainder mod 3, one and two rows; with non power-of-two spacing the Laplacians divide by dx^2, dy^2 through
    the refined reciprocal and fall back to the real division outside its window -- exact zeros, -0.0, denormals and
    large values in u and v.  F and G bit for bit (the sign of zero included)."""
    m = nsb()
    s, P = _solver(m, i_max=ni, j_max=nj, a=a, b=b, Re=313.0, g_x=0.0, g_y=-0.5)
    u, v = _rand(s.shape, 61), _rand(s.shape, 62)
    i1, i2 = ni // 4, ni // 2
    u[i1:i2, :] = 0.0; v[i1 + 2:i2 + 2, :] = 0.0
    u[i1 + 1, :] = -0.0
    u[i2 + 3, 1:] = 1e-310; v[i2 + 5, :] = -3e-320; u[i2 + 7, 1] = 4.9e-324
    u[2, :] = 1e120; v[3, 1] = -1e140
    s.upload(m.U, u); s.upload(m.V, v)
    dx, dy = a / ni, b / nj
    with np.errstate(all="ignore"):
        for dt, gamma in ((0.0123, 0.41), (1e-3, 0.0)):
            s.FG(dt, gamma)
            F, G = np.zeros_like(u), np.zeros_like(u)
            ob.lib().orc_FG(ob._ptr(F), ob._ptr(G), ob._ptr(u), ob._ptr(v), ni, nj, P.Re, P.g_x, P.g_y, dt, dx, dy, gamma)
            assert np.all(np.isfinite(F)) and np.all(np.isfinite(G))
            assert np.array_equal(s.download(m.F).view(np.int64), F.view(np.int64))
            assert np.array_equal(s.download(m.G).view(np.int64), G.view(np.int64))
    s.close()


def test_layout_roundtrip_both_layouts():
    m = nsb()
    s, P = _solver(m, i_max=45, j_max=70)
    a = _rand(s.shape, 9)
    s.upload(m.P, a, m.LAYOUT_REF)
    assert np.array_equal(s.download(m.P, m.LAYOUT_REF), a)
    assert np.array_equal(s.download(m.P, m.LAYOUT_XFAST), a.T)
    s.upload(m.U, np.ascontiguousarray(a.T), m.LAYOUT_XFAST)
    assert np.array_equal(s.download(m.U, m.LAYOUT_REF), a)
    s.close()


# ------------------------------------------------------------------------------- SOR

SOR_SHAPES = [(8, 8), (16, 16), (31, 17), (120, 40), (121, 33), (257, 64), (64, 300), (500, 7), (3, 3), (1, 9), (9, 1)]


@pytest.mark.parametrize("engine", ["stream", "launch", "simple", "resident"])
@pytest.mark.parametrize("ni,nj", SOR_SHAPES)
def test_sor_per_iteration_bit_exact(engine, ni, nj):
    """p after 1, 2, 3, +5 iterations equals the oracle's red-black iterate bit for bit, residual
    norm to 1e-12 relative; sizes include non-multiples of every tile dimension (the reference's
    CUDA kernel is wrong for those)."""
    m = nsb()
    eng = dict(stream=m.SOR_STREAM, launch=m.SOR_STREAM_LAUNCH, simple=m.SOR_SIMPLE, resident=m.SOR_RESIDENT)[engine]
    a, b = 1.0, 0.9
    s, P = _solver(m, engine=eng, i_max=ni, j_max=nj, a=a, b=b, omega=1.7)
    L = ob.lib()
    p = _rand(s.shape, 11) * 0.5
    rhs = _rand(s.shape, 12) * 3.0
    s.upload(m.P, p); s.upload(m.RHS, rhs)
    dx, dy = a / ni, b / nj
    for n in (1, 1, 1, 5):
        r_gpu = s.sor_iterations(n)
        for _ in range(n):
            r_cpu = L.orc_SOR_rb_iter(ob._ptr(p), ni, nj, dx, dy, ob._ptr(rhs), P.omega)
        got = s.download(m.P)
        assert np.array_equal(got[1:-1, 1:-1], p[1:-1, 1:-1])
        # ghosts (non-corner) hold what the refresh before the last black sweep copied
        assert np.array_equal(got[0, 1:-1], p[0, 1:-1]) and np.array_equal(got[-1, 1:-1], p[-1, 1:-1])
        assert np.array_equal(got[1:-1, 0], p[1:-1, 0]) and np.array_equal(got[1:-1, -1], p[1:-1, -1])
        assert abs(r_gpu - r_cpu) <= 1e-12 * abs(r_cpu)
    s.close()


@pytest.mark.parametrize("ni,nj", [(31, 17), (257, 64), (64, 300), (600, 480)])
def test_resident_grid_variant(ni, nj, monkeypatch):
    """The cooperative-grid variant of the resident engine (tiles over all SMs, halos through L2), forced at
    sizes the oracle reaches: same bits per iteration, same converged iteration count as the simple engine."""
    m = nsb()
    monkeypatch.setenv("NSB_RES_MODE", "grid")
    s, P = _solver(m, engine=m.SOR_RESIDENT, i_max=ni, j_max=nj, a=1.0, b=0.9, omega=1.7, epsilon=1e-3, max_it=400)
    L = ob.lib()
    p = _rand(s.shape, 21) * 0.5
    rhs = _rand(s.shape, 22) * 3.0
    rhs -= rhs[1:-1, 1:-1].mean()
    p0 = p.copy()
    s.upload(m.P, p); s.upload(m.RHS, rhs)
    dx, dy = 1.0 / ni, 0.9 / nj
    for n in (1, 2, 5):
        r_gpu = s.sor_iterations(n)
        for _ in range(n):
            r_cpu = L.orc_SOR_rb_iter(ob._ptr(p), ni, nj, dx, dy, ob._ptr(rhs), P.omega)
        got = s.download(m.P)
        assert np.array_equal(got[1:-1, 1:-1], p[1:-1, 1:-1])
        assert np.array_equal(got[0, 1:-1], p[0, 1:-1]) and np.array_equal(got[-1, 1:-1], p[-1, 1:-1])
        assert np.array_equal(got[1:-1, 0], p[1:-1, 0]) and np.array_equal(got[1:-1, -1], p[1:-1, -1])
        assert abs(r_gpu - r_cpu) <= 1e-12 * abs(r_cpu)
    # converged solve with the roll-back of the speculative first colour: against the simple engine
    s.upload(m.P, p0)
    ret, its, res, n0 = s.SOR()
    got = s.download(m.P)
    s.close()
    monkeypatch.delenv("NSB_RES_MODE")
    s2, _ = _solver(m, engine=m.SOR_SIMPLE, i_max=ni, j_max=nj, a=1.0, b=0.9, omega=1.7, epsilon=1e-3, max_it=400)
    s2.upload(m.P, p0); s2.upload(m.RHS, rhs)
    ret2, its2, res2, n02 = s2.SOR()
    want = s2.download(m.P)
    s2.close()
    assert (ret, its) == (ret2, its2)
    assert np.array_equal(got[1:-1, 1:-1], want[1:-1, 1:-1])
    assert np.array_equal(got[0, 1:-1], want[0, 1:-1]) and np.array_equal(got[1:-1, 0], want[1:-1, 0])


@pytest.mark.parametrize("engine", ["stream", "launch", "simple", "resident"])
@pytest.mark.parametrize("ni,nj", [(150, 40), (400, 36)])
def test_sor_outside_division_window(engine, ni, nj):
    """IEEE division by dx^2 / dy^2 is done with a refined reciprocal and a residual correction that is
    only valid inside an exponent window (nsb_math.cuh); operands outside it -- denormals, values next
    to overflow -- and exact zeros of either sign must still give the reference's bits (the streaming
    sweep repeats the task with real divisions, the other engines branch per cell)."""
    m = nsb()
    eng = dict(stream=m.SOR_STREAM, launch=m.SOR_STREAM_LAUNCH, simple=m.SOR_SIMPLE, resident=m.SOR_RESIDENT)[engine]
    a, b = 1.0, 0.9
    s, P = _solver(m, engine=eng, chunk_rows=16, i_max=ni, j_max=nj, a=a, b=b, omega=1.7)
    p = _rand(s.shape, 31) * 0.5
    rhs = _rand(s.shape, 32) * 3.0
    # reference layout: [i][j].  A block of exact zeros (+0 and -0) with denormals inside, a block of huge values.
    p[20:60, 5:30] = 0.0; p[30:40, 10:20] = -0.0; rhs[18:62, 3:32] = 0.0
    p[25, 12] = 1e-310; p[45, 20] = -3e-320; p[50, 8] = 4.9e-324
    big = 4e306 * (a / ni) ** 2          # quotients by dx^2 land between 2^1017 and overflow
    p[100:110, 10:20] *= big
    if ni > 300:
        p[200:240, 4:34] = 0.0; rhs[198:242, 2:36] = 0.0; p[220, 20] = 1e-312; p[300:305, 6:12] *= big
    s.upload(m.P, p); s.upload(m.RHS, rhs)
    dx, dy = a / ni, b / nj
    L = ob.lib()
    with np.errstate(all="ignore"):
        for n in (1, 2):
            s.sor_iterations(n)
            for _ in range(n):
                L.orc_SOR_rb_iter(ob._ptr(p), ni, nj, dx, dy, ob._ptr(rhs), P.omega)
            got = s.download(m.P)
            assert np.all(np.isfinite(p[1:-1, 1:-1]))
            assert np.array_equal(got[1:-1, 1:-1].view(np.int64), p[1:-1, 1:-1].view(np.int64))   # the sign of zero counts
    s.close()


@pytest.mark.parametrize("div_mode", ["exact", "recip"])
def test_sor_division_modes(div_mode):
    """Non power-of-two spacing: NSB_DIV_EXACT (and AUTO) is bit-exact; NSB_DIV_RECIP stays within
    the stated 1e-8 relative tolerance."""
    m = nsb()
    ni, nj, a, b = 50, 30, 1.0, 1.0
    P = m.make_params(i_max=ni, j_max=nj, a=a, b=b)
    s = m.Solver(P, sor_engine=m.SOR_STREAM, div_mode=dict(exact=m.DIV_EXACT, recip=m.DIV_RECIP)[div_mode])
    p, rhs = _rand(s.shape, 21), _rand(s.shape, 22)
    s.upload(m.P, p); s.upload(m.RHS, rhs)
    s.sor_iterations(20)
    for _ in range(20):
        ob.lib().orc_SOR_rb_iter(ob._ptr(p), ni, nj, a / ni, b / nj, ob._ptr(rhs), P.omega)
    got = s.download(m.P)[1:-1, 1:-1]
    if div_mode == "exact":
        assert np.array_equal(got, p[1:-1, 1:-1])
    else:
        assert np.max(np.abs(got - p[1:-1, 1:-1])) <= 1e-8 * np.max(np.abs(p))
    s.close()


@pytest.mark.parametrize("engine", ["stream", "launch", "simple", "resident"])
def test_sor_convergence_iteration_count(engine):
    """Converged solve: same iteration count, same return value (it+1), bit-equal p; then the
    not-converged contract (-1, p = last iterate)."""
    m = nsb()
    eng = dict(stream=m.SOR_STREAM, launch=m.SOR_STREAM_LAUNCH, simple=m.SOR_SIMPLE, resident=m.SOR_RESIDENT)[engine]
    ni, nj = 48, 40
    s, P = _solver(m, engine=eng, i_max=ni, j_max=nj, epsilon=1e-5, max_it=5000)
    p = np.zeros(s.shape)
    x = (np.arange(ni + 2) - 0.5) / ni
    y = (np.arange(nj + 2) - 0.5) / nj
    rhs = np.ascontiguousarray(np.cos(np.pi * x)[:, None] * np.cos(np.pi * y)[None, :])
    s.upload(m.P, p); s.upload(m.RHS, rhs)
    ret, its, res, n0 = s.SOR()
    oret, oits, ores, on0 = ob.sor_rb(p, rhs, 1.0 / ni, 1.0 / nj, P.omega, P.epsilon, P.max_it)
    assert (ret, its) == (oret, oits) and ret > 0
    assert np.array_equal(s.download(m.P)[1:-1, 1:-1], p[1:-1, 1:-1])
    assert abs(res - ores) <= 1e-12 * ores and n0 == on0
    s.close()
    # max_it too small
    s, P = _solver(m, engine=eng, i_max=ni, j_max=nj, epsilon=1e-9, max_it=37)
    p = np.zeros(s.shape)
    s.upload(m.P, p); s.upload(m.RHS, rhs)
    ret, its, res, n0 = s.SOR()
    oret, oits, ores, on0 = ob.sor_rb(p, rhs, 1.0 / ni, 1.0 / nj, P.omega, P.epsilon, P.max_it)
    assert ret == oret == -1 and its == oits == 37
    assert np.array_equal(s.download(m.P)[1:-1, 1:-1], p[1:-1, 1:-1])
    s.close()


# ------------------------------------------------------------------------------- whole runs

def _compare_run(m, c, engine):
    P = m.make_params(**c["params"])
    s = m.Solver(P, sor_engine=engine)
    log, t_end = s.run()
    assert len(log) == c["steps"]
    assert [x["sor_iters"] for x in log] == c["sor_iters"]
    if c["sor_order"] == 0:      # serial SOR returns 0 / -1 (integration.h:165); the ABI reports it+1 / -1
        assert [0 if x["sor_ret"] > 0 else -1 for x in log] == c["sor_ret"]
    else:
        assert [x["sor_ret"] for x in log] == c["sor_ret"]
    assert [float(x["dt"]).hex() for x in log] == c["dt_hex"]
    ref = np.load(os.path.join(GOLD, c["fields"]))
    for name, fid in (("u", m.U), ("v", m.V), ("p", m.P)):
        got = s.download(fid)
        want = ref[name]
        if name == "p":       # corner ghosts are unspecified in the reference (racy copies)
            got, want = got.copy(), want.copy()
            for arr in (got, want):
                arr[0, 0] = arr[0, -1] = arr[-1, 0] = arr[-1, -1] = 0.0
        assert np.array_equal(got, want), (c["name"], name)
    uc, vc = s.center()
    assert float(uc).hex() == c["u_center_hex"] and float(vc).hex() == c["v_center_hex"]
    s.close()


@pytest.mark.parametrize("engine", ["stream", "launch", "simple", "resident"])
def test_whole_run_matches_committed_redblack_fixtures(engine):
    m = nsb()
    eng = dict(stream=m.SOR_STREAM, launch=m.SOR_STREAM_LAUNCH, simple=m.SOR_SIMPLE, resident=m.SOR_RESIDENT)[engine]
    cases = [c for c in json.load(open(os.path.join(GOLD, "runs.json"))) if c["sor_order"] == 1]
    assert len(cases) >= 4
    for c in cases:
        _compare_run(m, c, eng)


def test_whole_run_vs_live_oracle_nonsquare_gravity():
    m = nsb()
    kw = dict(i_max=70, j_max=45, a=1.4, b=0.9, T=0.3, Re=200.0, g_x=0.02, g_y=-0.05, tau=0.8, max_it=400)
    s, P = _solver(m, **kw)
    log, _ = s.run()
    o = ob.run(ob.make_params(**kw), sor_order=1)
    assert [x["sor_iters"] for x in log] == [x["sor_iters"] for x in o["log"]]
    assert [x["dt"] for x in log] == [x["dt"] for x in o["log"]]
    for name, fid in (("u", m.U), ("v", m.V)):
        assert np.array_equal(s.download(fid), o[name])
    assert np.array_equal(s.download(m.P)[1:-1, 1:-1], o["p"][1:-1, 1:-1])
    s.close()


def test_serial_order_agreement_at_tight_eps():
    """north_star: <= 1e-8 relative max-abs difference to the serial C solver after convergence
    to the same (tight) eps.  p is compared mean-free (pure Neumann problem)."""
    m = nsb()
    kw = dict(i_max=32, j_max=32, T=0.6, epsilon=1e-10, max_it=200000)
    s, P = _solver(m, **kw)
    log, _ = s.run()
    o = ob.run(ob.make_params(**kw), sor_order=0)
    assert len(log) == o["steps"]
    for name, fid in (("u", m.U), ("v", m.V)):
        a, b = s.download(fid), o[name]
        assert np.max(np.abs(a - b)) <= 1e-8 * np.max(np.abs(b))
    a, b = s.download(m.P)[1:-1, 1:-1], o["p"][1:-1, 1:-1]
    a, b = a - a.mean(), b - b.mean()
    assert np.max(np.abs(a - b)) <= 1e-8 * np.max(np.abs(b))
    s.close()


def test_unknown_problem_type_is_an_error():
    m = nsb()
    s, P = _solver(m, i_max=8, j_max=8, problem=3)
    with pytest.raises(m.NsbError):
        s.step(0.0)
    s.close()


# ------------------------------------------------------------------------------- size-independent properties

def test_stream_engine_equals_simple_engine_at_scale():
    """At a size the CPU oracle cannot reach quickly: the fused streaming pass and the literal
    launch-per-phase engine must produce the same bits."""
    m = nsb()
    ni, nj = 2048, 1536
    outs = []
    for eng in (m.SOR_STREAM, m.SOR_STREAM_LAUNCH, m.SOR_SIMPLE):
        s, P = _solver(m, engine=eng, i_max=ni, j_max=nj)
        s.upload(m.P, _rand(s.shape, 31)); s.upload(m.RHS, _rand(s.shape, 32))
        r = s.sor_iterations(7)
        outs.append((s.download(m.P)[1:-1, 1:-1], r))
        s.close()
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][0], outs[2][0])
    assert outs[0][1] == outs[1][1], "persistent and per-launch sweeps must sum the residual in the same order"
    assert abs(outs[0][1] - outs[2][1]) <= 1e-12 * outs[2][1]


def test_stream_engine_is_race_free_over_repeated_runs():
    """Regression for a shared-memory reuse race (TMA re-arming a stage before the last LDS of the
    previous row pair had landed): it showed as sporadically wrong cells next to strip edges in about
    half of the runs at this shape.  20 runs must all produce the bits of the launch-per-phase engine."""
    m = nsb()
    ni, nj = 2048, 1536
    p0, r0 = _rand((ni + 2, nj + 2), 31), _rand((ni + 2, nj + 2), 32)
    s, P = _solver(m, engine=m.SOR_SIMPLE, i_max=ni, j_max=nj)
    s.upload(m.P, p0); s.upload(m.RHS, r0)
    s.sor_iterations(3)
    want = s.download(m.P)[1:-1, 1:-1]
    s.close()
    for eng in (m.SOR_STREAM, m.SOR_STREAM_LAUNCH):
        for chunk in (0, 16):
            for rep in range(10):
                s, P = _solver(m, engine=eng, i_max=ni, j_max=nj, chunk_rows=chunk)
                s.upload(m.P, p0); s.upload(m.RHS, r0)
                s.sor_iterations(3)
                got = s.download(m.P)[1:-1, 1:-1]
                s.close()
                assert np.array_equal(got, want), f"engine {eng} run {rep} (chunk_rows={chunk}) differs in {(got != want).sum()} cells"


def test_sor_linearity_and_constant_nullspace():
    """SOR is affine in (p, RHS): with RHS = 0, iterating a*p gives a*iterate(p) for a power of
    two (exact in FP64); a constant field is a fixed point up to rounding (pure Neumann nullspace)."""
    m = nsb()
    ni, nj = 300, 200
    s, P = _solver(m, i_max=ni, j_max=nj)
    p = _rand(s.shape, 41)
    zero = np.zeros(s.shape)
    s.upload(m.RHS, zero)
    s.upload(m.P, p); s.sor_iterations(6); a = s.download(m.P)[1:-1, 1:-1]
    s.upload(m.P, 4.0 * p); s.sor_iterations(6); b = s.download(m.P)[1:-1, 1:-1]
    assert np.array_equal(4.0 * a, b)
    s.upload(m.P, np.full(s.shape, 3.25)); r = s.sor_iterations(3)
    assert np.max(np.abs(s.download(m.P)[1:-1, 1:-1] - 3.25)) <= 1e-13 and r <= 1e-8
    s.close()


# ------------------------------------------------------------------------------- host-grid API, driver, output

def test_host_grid_api_SOR_and_FG():
    m = nsb()
    L = m.lib()
    ni, nj = 40, 24
    dx, dy = 1.0 / ni, 1.0 / nj
    p, rhs = _rand((ni + 2, nj + 2), 51) * 0.1, _rand((ni + 2, nj + 2), 52)
    rhs[1:-1, 1:-1] -= rhs[1:-1, 1:-1].mean()
    q = p.copy()
    res = np.zeros_like(p)
    it, lr = C.c_int(), C.c_double()
    rc = L.nsb_SOR_iters(m.rows(q), ni, nj, dx, dy, m.rows(res), m.rows(rhs), 1.7, 1e-6, 3000, C.byref(it), C.byref(lr))
    oret, oits, ores, _ = ob.sor_rb(p, rhs, dx, dy, 1.7, 1e-6, 3000)
    assert rc == oret and it.value == oits
    assert np.array_equal(q[1:-1, 1:-1], p[1:-1, 1:-1])
    # serial-style return code
    q2 = np.zeros_like(p)
    assert L.nsb_SOR(m.rows(q2), ni, nj, dx, dy, m.rows(res), m.rows(rhs), 1.7, 1e-6, 3000) == 0
    assert L.nsb_SOR(m.rows(q2), ni, nj, dx, dy, m.rows(res), m.rows(rhs), 1.7, 1e-14, 3) == -1
    # FG on jagged grids: u has i_max+1 rows, v rows have j_max+1 entries
    u = _rand((ni + 1, nj + 2), 53)
    v = _rand((ni + 2, nj + 1), 54)
    F, G = np.zeros((ni + 2, nj + 2)), np.zeros((ni + 2, nj + 2))
    assert L.nsb_FG(m.rows(F), m.rows(G), m.rows(u), m.rows(v), ni, nj, 100.0, 0.0, 0.0, 0.01, dx, dy, 0.3) == 0
    uf, vf = np.zeros((ni + 2, nj + 2)), np.zeros((ni + 2, nj + 2))
    uf[:ni + 1] = u; vf[:, :nj + 1] = v
    F2, G2 = np.zeros_like(F), np.zeros_like(F)
    ob.lib().orc_FG(ob._ptr(F2), ob._ptr(G2), ob._ptr(uf), ob._ptr(vf), ni, nj, 100.0, 0.0, 0.0, 0.01, dx, dy, 0.3)
    assert np.array_equal(F, F2) and np.array_equal(G, G2)
    assert L.nsb_max_mat(ni, nj, m.rows(u)) == ob.lib().orc_max_mat(ni, nj, ob._ptr(uf))
    # ... and on the v shape (rows of j_max+1 doubles: nothing past a row's end may be read -- the last row of v ends
    # the allocation here, so an over-read would run off the numpy buffer)
    assert L.nsb_max_mat(ni, nj, m.rows(v)) == ob.lib().orc_max_mat(ni, nj, ob._ptr(vf))
    L.nsb_host_api_reset()


def test_host_grid_timestep_equals_context_run():
    m = nsb()
    L = m.lib()
    kw = dict(i_max=32, j_max=32, T=1.5)
    P = m.make_params(**kw)
    ni, nj = P.i_max, P.j_max
    u, v = np.zeros((ni + 1, nj + 2)), np.zeros((ni + 2, nj + 1))
    p, F, G, R = (np.zeros((ni + 2, nj + 2)) for _ in range(4))
    t, its = 0.0, []
    while t < P.T:
        info = m.StepInfo()
        assert L.nsb_timestep(C.byref(P), t, m.rows(u), m.rows(v), m.rows(p), m.rows(F), m.rows(G), m.rows(R), C.byref(info)) == 0
        its.append(info.sor_iters)
        t += info.dt
    L.nsb_host_api_reset()
    c = [c for c in json.load(open(os.path.join(GOLD, "runs.json"))) if c["name"] == "rb_32"][0]
    assert its == c["sor_iters"]
    ref = np.load(os.path.join(GOLD, c["fields"]))
    assert np.array_equal(u, ref["u"][:ni + 1]) and np.array_equal(v, ref["v"][:, :nj + 1])
    assert np.array_equal(p[1:-1, 1:-1], ref["p"][1:-1, 1:-1])


def test_output_files_match_reference_format(tmp_path):
    """io.c:61-120 text format incl. its quirks, byte for byte against the oracle's writer (which is
    pinned to the reference's output() in test_oracle_output)."""
    m = nsb()
    kw = dict(i_max=12, j_max=9, T=0.5, a=1.0, b=0.75)
    s, P = _solver(m, **kw)
    log, t_end = s.run()
    (tmp_path / "g").mkdir(); (tmp_path / "o").mkdir()
    s.output(t_end, str(tmp_path / "g" / "0"))
    u, v, p = s.download(m.U), s.download(m.V), s.download(m.P)
    ob.lib().orc_output(P.i_max, P.j_max, ob._ptr(u), ob._ptr(v), ob._ptr(p), t_end, P.a, P.b, str(tmp_path / "o" / "0").encode())
    for suf in ("_u.txt", "_v.txt", "_p.txt"):
        assert (tmp_path / "g" / ("0" + suf)).read_bytes() == (tmp_path / "o" / ("0" + suf)).read_bytes()
    txt = (tmp_path / "g" / "0_u.txt").read_text().splitlines()
    assert len(txt) == 3 + P.j_max + 2 and len(txt[3].split()) == P.i_max + 1
    assert txt[-1].split()[-1] == "2.00000"          # u[i_max][j_max+1] = 2 quirk
    s.close()


def test_output_async_and_slabs_write_the_same_bytes(tmp_path):
    """nsb_output_async (device snapshot, side-stream copy, host formatting threads) while the time loop goes on,
    and nsb_group_output over 3 y-slabs: the same bytes as the oracle's writer (pinned to io.c:output) for the
    fields AT THE TIME OF THE CALL."""
    m = nsb()
    kw = dict(i_max=70, j_max=48, T=1e9, a=1.0, b=0.75, max_it=60)
    P = m.make_params(**kw)
    (tmp_path / "a").mkdir(); (tmp_path / "o").mkdir(); (tmp_path / "g").mkdir()
    with m.Solver(P, sor_engine=m.SOR_STREAM, chunk_rows=16) as s:
        t = 0.0
        for k in range(3):
            t += s.step(t)["dt"]
        u, v, p = s.download(m.U), s.download(m.V), s.download(m.P)
        s.output(t, str(tmp_path / "a" / "0"), wait=False)
        t2 = t
        for k in range(3):                                   # the solver moves on while the dump is written
            t2 += s.step(t2)["dt"]
        s.output_wait()
        assert not np.array_equal(s.download(m.U), u)
    ob.lib().orc_output(P.i_max, P.j_max, ob._ptr(u), ob._ptr(v), ob._ptr(p), t, P.a, P.b, str(tmp_path / "o" / "0").encode())
    with m.Group(P, [0, 0, 0], chunk_rows=16) as g:
        tg = 0.0
        for k in range(3):
            tg += g.step(tg)["dt"]
        assert tg == t
        g.output(tg, str(tmp_path / "g" / "0"))
    for suf in ("_u.txt", "_v.txt", "_p.txt"):
        want = (tmp_path / "o" / ("0" + suf)).read_bytes()
        assert (tmp_path / "a" / ("0" + suf)).read_bytes() == want
        assert (tmp_path / "g" / ("0" + suf)).read_bytes() == want


def test_c_driver_process_contract(tmp_path):
    """build/parallel: same stdout lines as the serial reference binary at %.6f, stderr ends with
    the SOR seconds and no newline."""
    exe = os.path.join(ROOT, "build", "parallel")
    assert os.path.exists(exe)
    f = tmp_path / "case.in"
    f.write_text("1\n1\n64\n64\n1.0\n1.0\n0.1\n1000.0\n0.0\n0.0\n1.0\n1.7\n0.0001\n20000\n1\n")
    out = subprocess.run([exe, str(f), "16"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    o = ob.run(ob.make_params(i_max=64, j_max=64, T=0.1), sor_order=1)
    assert out.stdout == "U-CENTER: %.6f\nV-CENTER: %.6f\n" % (o["u"][32, 32], o["v"][32, 32])
    last = out.stderr.splitlines()[-1]
    assert not out.stderr.endswith("\n") and float(last) > 0.0
    if ob.have_ref():
        ref = subprocess.run([ob.REF_SERIAL, str(f)], capture_output=True, text=True)
        # the reference's own acceptance rule (colab-runner.ipynb cell 4): tokens within 1e-4
        for a, b in zip(out.stdout.split(), ref.stdout.split()):
            try:
                assert abs(float(a) - float(b)) <= 1e-4
            except ValueError:
                assert a == b


# ------------------------------------------------------------------------------- BASELINE sizes, reference binary

def _oracle_first_step(P, n_it):
    """Oracle: BC, F, G, RHS of the first cavity step, then n_it red-black iterations on all host cores."""
    L = ob.lib()
    oP = _orc_params(P)
    ni, nj = P.i_max, P.j_max
    dx, dy = P.a / ni, P.b / nj
    u, v, F, G, R, p = (ob.zeros(oP) for _ in range(6))
    dt, gamma = C.c_double(), C.c_double()
    L.orc_dt_gamma(C.byref(oP), dx, dy, 0.0, 0.0, C.byref(dt), C.byref(gamma))
    L.orc_apply_bc(C.byref(oP), 0.0, ob._ptr(u), ob._ptr(v))
    L.orc_FG(ob._ptr(F), ob._ptr(G), ob._ptr(u), ob._ptr(v), ni, nj, P.Re, P.g_x, P.g_y, dt.value, dx, dy, gamma.value)
    L.orc_RHS(ob._ptr(R), ob._ptr(F), ob._ptr(G), ni, nj, dt.value, dx, dy)
    del u, v, F, G
    _, its, res = ob.sor_rb_omp(p, R, dx, dy, P.omega, 0.0, n_it)
    assert its == n_it
    return p, R, res


def test_full_size_8192_against_oracle_and_nullspace():
    """BASELINE.json configs[2] size (8192^2) at the PRODUCTION task shape (planned chunk height, 69 strips x 64 chunks,
    persistent sweep) against the CPU oracle (orc_SOR_rb_omp on all host cores): RHS and p of the cavity's first step
    after 3 iterations, then 4 more iterations on a developed (random, zero-free) field -- bit for bit, residual 1e-12.
    Size-independent property on top: adding a constant to p shifts the iterate by exactly that constant (pure
    Neumann problem: constants are the null space; 0.5 is exact in binary)."""
    m = nsb()
    n = 8192
    P = m.make_params(i_max=n, j_max=n, T=1e9, max_it=3)
    po, Ro, ro = _oracle_first_step(P, 3)
    with m.Solver(P, sor_engine=m.SOR_STREAM) as s:
        info = s.step(0.0)                            # BC, F, G, RHS of the first cavity step + 3 iterations
        assert info["sor_iters"] == 3 and abs(info["res"] - ro) <= 1e-12 * ro
        assert np.array_equal(s.download(m.RHS)[1:-1, 1:-1], Ro[1:-1, 1:-1])
        assert np.array_equal(s.download(m.P)[1:-1, 1:-1].view(np.int64), po[1:-1, 1:-1].view(np.int64))
        del po
        q = np.random.default_rng(5).uniform(-1.0, 1.0, (n + 2, n + 2))
        s.upload(m.P, q)
        r1 = s.sor_iterations(4)
        a = s.download(m.P)[1:-1, 1:-1].copy()
        _, _, r1o = ob.sor_rb_omp(q, Ro, 1.0 / n, 1.0 / n, P.omega, 0.0, 4)
        assert np.array_equal(a, q[1:-1, 1:-1]) and abs(r1 - r1o) <= 1e-12 * r1o
        # null space
        q = np.random.default_rng(5).uniform(-1.0, 1.0, (n + 2, n + 2))
        q += 0.5
        s.upload(m.P, q)
        del q
        r2 = s.sor_iterations(4)
        d = s.download(m.P)[1:-1, 1:-1] - a
        assert np.abs(d - 0.5).max() <= 64 * np.finfo(float).eps * max(1.0, np.abs(a).max())
        assert abs(r2 - r1) <= 1e-9 * r1


def test_large_grid_ieee_division_against_oracle():
    """A large grid whose spacing is not a power of two (4100 x 4090: IEEE division by dx^2 / dy^2, done with the refined
    reciprocal and the window test of nsb_math.cuh; many chunks of the planned height, partial last strip and chunk)
    against the CPU oracle, which really divides: the first cavity step (p exactly zero away from the lid: the +-0
    paths) and a developed field without zeros, bit for bit; the per-launch and validation engines must agree too."""
    m = nsb()
    ni, nj = 4100, 4090
    P = m.make_params(i_max=ni, j_max=nj, T=1e9, max_it=3)
    po, Ro, ro = _oracle_first_step(P, 5)             # the step's 3 iterations + 2 more
    q0 = np.random.default_rng(77).uniform(-1.0, 1.0, (ni + 2, nj + 2))
    qo = q0.copy()
    _, _, r1o = ob.sor_rb_omp(qo, Ro, P.a / ni, P.b / nj, P.omega, 0.0, 3)
    for eng in (m.SOR_STREAM, m.SOR_STREAM_LAUNCH, m.SOR_SIMPLE):
        with m.Solver(P, sor_engine=eng) as s:
            s.step(0.0)                               # zeros almost everywhere
            r0 = s.sor_iterations(2)
            assert np.array_equal(s.download(m.RHS)[1:-1, 1:-1], Ro[1:-1, 1:-1])
            assert np.array_equal(s.download(m.P)[1:-1, 1:-1].view(np.int64), po[1:-1, 1:-1].view(np.int64))
            s.upload(m.P, q0)                         # a field without zeros
            r1 = s.sor_iterations(3)
            assert np.array_equal(s.download(m.P)[1:-1, 1:-1], qo[1:-1, 1:-1])
            assert abs(r0 - ro) <= 1e-12 * abs(ro) and abs(r1 - r1o) <= 1e-12 * abs(r1o)


def test_reference_cuda_binary_same_stdout(tmp_path):
    """The reference's own CUDA build (src/parallel/*.cu compiled unmodified for sm_100 into
    oracle/_ref/parallel by `make -C oracle refcuda`) on this GPU against build/parallel, with the
    reference's acceptance rule (colab-runner.ipynb cell 4: stdout tokens within 1e-4).  64^2 and 128^2
    are multiples of the block size 16 -- the only sizes the reference kernel is valid for."""
    ref = os.path.join(ROOT, "oracle", "_ref", "parallel")
    exe = os.path.join(ROOT, "build", "parallel")
    if not os.path.exists(ref):
        pytest.skip("oracle/_ref/parallel not built (make -C oracle refcuda)")
    import importlib.util
    spec = importlib.util.spec_from_file_location("compare_outputs", os.path.join(ROOT, "tools", "compare_outputs.py"))
    cmp = importlib.util.module_from_spec(spec); spec.loader.exec_module(cmp)
    for n, T in ((64, 0.5), (128, 0.05)):
        f = tmp_path / f"c{n}.in"
        f.write_text(f"1\n1\n{n}\n{n}\n1.0\n1.0\n{T}\n1000.0\n0.0\n0.0\n1.0\n1.7\n0.0001\n2000\n1\n")
        a = subprocess.run([ref, str(f), "16"], capture_output=True, text=True, timeout=300)
        b = subprocess.run([exe, str(f), "16"], capture_output=True, text=True, timeout=300)
        assert a.returncode == 0 and b.returncode == 0, (a.stderr, b.stderr)
        (tmp_path / "a.txt").write_text(a.stdout); (tmp_path / "b.txt").write_text(b.stdout)
        ok, msg = cmp.equivalent(str(tmp_path / "a.txt"), str(tmp_path / "b.txt"))
        assert ok, msg
        assert a.stdout.startswith("U-CENTER: ") and b.stdout.count("\n") == 2


# ------------------------------------------------------------------------------- serial (lexicographic) order

@pytest.mark.parametrize("ni,nj", [(8, 8), (31, 17), (64, 64), (120, 40), (33, 130)])
def test_serial_order_sor_bit_exact(ni, nj):
    """NSB_SOR_SERIAL_ORDER: the in-place lexicographic sweep of the SERIAL reference (integration.c:129-173)
    run as an anti-diagonal wavefront: p bit for bit, same number of sweeps, same 0 / -1 outcome."""
    m = nsb()
    a, b = 1.0, 0.9
    for eps, max_it in ((1e-3, 3000), (1e-12, 23)):
        s, P = _solver(m, engine=m.SOR_SERIAL_ORDER, i_max=ni, j_max=nj, a=a, b=b, omega=1.7, epsilon=eps, max_it=max_it)
        p = _rand(s.shape, 41) * 0.5
        rhs = _rand(s.shape, 42)
        rhs -= rhs[1:-1, 1:-1].mean()
        s.upload(m.P, p); s.upload(m.RHS, rhs)
        ret, its, res, n0 = s.SOR()
        oret, oits, ores, on0 = ob.sor_lex(p, rhs, a / ni, b / nj, P.omega, eps, max_it)
        assert (0 if ret > 0 else -1, its) == (oret, oits)
        got = s.download(m.P)
        assert np.array_equal(got[1:-1, 1:-1], p[1:-1, 1:-1])
        assert np.array_equal(got[0, 1:-1], p[0, 1:-1]) and np.array_equal(got[1:-1, -1], p[1:-1, -1])
        assert abs(res - ores) <= 1e-12 * ores and abs(n0 - on0) <= 1e-14 * on0
        s.close()


def test_whole_run_matches_committed_serial_fixtures():
    """Whole runs in the serial order against fixtures generated from the UNMODIFIED reference serial
    sources (oracle/_ref/libnsref.so): fields, dt sequence and iteration counts, bit for bit."""
    m = nsb()
    runs = [c for c in json.load(open(os.path.join(GOLD, "runs.json"))) if c["sor_order"] == 0]
    assert runs
    for c in runs:
        _compare_run(m, c, m.SOR_SERIAL_ORDER)


# ------------------------------------------------------------------------------- IEEE division by dx^2, dy^2
@pytest.mark.parametrize("n", [130, 200, 1000, 1536, 3000, 6144, 10000])
def test_division_by_constant_matches_ddiv(n):
    """The sweep's x / dx^2 (refined reciprocal + residual correction, nsb_math.cuh) against __ddiv_rn, bit for bit:
    2^27 pairs of field-like operands, 2^26 pairs of random bit patterns and 2^22 edge patterns (zeros, denormals,
    the edges of the exponent window, Inf / NaN) per divisor pair."""
    L = nsb().lib()
    f = L.nsb_debug_divcheck
    f.restype = C.c_int
    f.argtypes = [C.c_double, C.c_double, C.c_longlong, C.c_ulonglong, C.c_int, C.POINTER(C.c_longlong),
                  C.POINTER(C.c_double), C.POINTER(C.c_double)]
    dx, dy = 1.0 / n, 0.75 / (n + 7)
    for mode, count in ((0, 1 << 27), (1, 1 << 26), (2, 1 << 22)):
        bad, fx, fd = C.c_longlong(-1), C.c_double(), C.c_double()
        assert f(dx * dx, dy * dy, count, 1234567 + n, mode, C.byref(bad), C.byref(fx), C.byref(fd)) == 0
        assert bad.value == 0, f"mode {mode}: {bad.value} quotients differ, first x = {fx.value.hex()} d = {fd.value.hex()}"


def test_resident_engine_refused_falls_back_within_the_same_solve(monkeypatch):
    """when the device refuses the cluster / cooperative launch (MIG slice, busy GPC, other SKU) the FIRST
    step must not fail: the same solve runs on the streaming sweep, and the refusal is remembered per context."""
    m = nsb()
    c = [c for c in json.load(open(os.path.join(GOLD, "runs.json"))) if c["name"] == "rb_64"][0]
    monkeypatch.setenv("NSB_DEBUG_RES_REFUSE", "1")
    _compare_run(m, c, m.SOR_AUTO)
    _compare_run(m, c, m.SOR_RESIDENT)
