"""Temporal blocking of the persistent sweep (two SOR iterations per pass over HBM, sor_stream.cu) against the oracle.

What is specific to the two-iteration pass and therefore tested here on top of tests/test_gpu_parity.py (whose
stream-engine cases already run it whenever a solve has >= 2 iterations):
  * per-sub-iteration residuals: the solve must stop at exactly the oracle's iteration, whether that is the first
    or the second sub-iteration of a pass (the first needs the replay from the intact input buffer), and return it+1;
  * odd and even iteration budgets (the odd one ends with a single-iteration pass), three rotating buffers over
    many solves in one context;
  * chunk heights from the smallest the 5-row halo allows to one chunk per strip; edge strips, partial last strip;
  * IEEE-division arithmetic (non power-of-two spacing) incl. the out-of-window fallback;
  * one iteration per pass (the default, nsb_options.sor_pass_k = 0 / 1) gives the same bits, iteration counts and
    residual sums as two (sor_pass_k = 2, what every test here asks for).
Bars: p bit-exact, iteration counts equal, residual norms 1e-12 relative (order-dependent sums).
"""
import numpy as np
import pytest

from oracle import binding as ob
from util import nsb

pytestmark = pytest.mark.gpu


def _rand(shape, seed, scale=1.0):
    return np.ascontiguousarray(np.random.default_rng(seed).uniform(-1, 1, shape) * scale)


@pytest.mark.parametrize("ni,nj,chunk", [(64, 64, 0), (250, 131, 10), (250, 131, 16), (121, 300, 64), (500, 40, 0), (37, 11, 0)])
def test_iteration_budgets_odd_and_even(ni, nj, chunk):
    """n = 2, 3, 4, 7, 8, 21 iterations in one context (the buffers rotate through all three positions): p and the
    non-corner ghosts equal the oracle's after every call."""
    m = nsb()
    a, b = 1.0, 0.9
    P = m.make_params(i_max=ni, j_max=nj, a=a, b=b)
    L = ob.lib()
    with m.Solver(P, pass_k=2, sor_engine=m.SOR_STREAM, chunk_rows=chunk) as s:
        p, rhs = _rand(s.shape, 1, 0.5), _rand(s.shape, 2, 3.0)
        s.upload(m.P, p); s.upload(m.RHS, rhs)
        for n in (2, 3, 4, 7, 8, 21):
            r_gpu = s.sor_iterations(n)
            for _ in range(n):
                r_cpu = L.orc_SOR_rb_iter(ob._ptr(p), ni, nj, a / ni, b / nj, ob._ptr(rhs), P.omega)
            got = s.download(m.P)
            assert np.array_equal(got[1:-1, 1:-1], p[1:-1, 1:-1]), n
            assert np.array_equal(got[0, 1:-1], p[0, 1:-1]) and np.array_equal(got[-1, 1:-1], p[-1, 1:-1])
            assert np.array_equal(got[1:-1, 0], p[1:-1, 0]) and np.array_equal(got[1:-1, -1], p[1:-1, -1])
            assert abs(r_gpu - r_cpu) <= 1e-12 * abs(r_cpu)


def test_convergence_in_either_sub_iteration():
    """Solves that converge after an odd and after an even number of iterations (first / second sub-iteration of a
    pass): same return value it+1, same iteration count, bit-equal p -- and the context keeps working afterwards."""
    m = nsb()
    ni, nj = 96, 80
    x = (np.arange(ni + 2) - 0.5) / ni
    y = (np.arange(nj + 2) - 0.5) / nj
    rhs = np.ascontiguousarray(np.cos(np.pi * x)[:, None] * np.cos(np.pi * y)[None, :])
    seen = set()
    for eps in (3e-3, 2e-3, 1e-3, 7e-4, 5e-4, 3e-4, 2e-4, 1e-4, 5e-5, 3e-5):
        po = np.zeros((ni + 2, nj + 2))
        oret, oits, ores, on0 = ob.sor_rb(po, rhs, 1.0 / ni, 1.0 / nj, 1.7, eps, 5000)
        assert oret == oits > 0
        seen.add(oits & 1)
        for max_it in (5000, oits, oits + 1):
            P = m.make_params(i_max=ni, j_max=nj, epsilon=eps, max_it=max_it)
            with m.Solver(P, pass_k=2, sor_engine=m.SOR_STREAM) as s:
                s.upload(m.P, np.zeros((ni + 2, nj + 2))); s.upload(m.RHS, rhs)
                ret, its, res, n0 = s.SOR()
                assert (ret, its) == (oret, oits), (eps, max_it)
                assert np.array_equal(s.download(m.P)[1:-1, 1:-1], po[1:-1, 1:-1])
                assert abs(res - ores) <= 1e-12 * ores and n0 == on0
                # the context is still consistent: two more iterations from the converged iterate
                q = po.copy()
                s.sor_iterations(2)
                for _ in range(2):
                    ob.lib().orc_SOR_rb_iter(ob._ptr(q), ni, nj, 1.0 / ni, 1.0 / nj, ob._ptr(rhs), 1.7)
                assert np.array_equal(s.download(m.P)[1:-1, 1:-1], q[1:-1, 1:-1])
        if oits < 3:
            continue
        # budget one short of convergence: -1, p = last iterate
        P = m.make_params(i_max=ni, j_max=nj, epsilon=eps, max_it=oits - 1)
        with m.Solver(P, pass_k=2, sor_engine=m.SOR_STREAM) as s:
            s.upload(m.P, np.zeros((ni + 2, nj + 2))); s.upload(m.RHS, rhs)
            ret, its, res, n0 = s.SOR()
            po = np.zeros((ni + 2, nj + 2))
            oret2, oits2, _, _ = ob.sor_rb(po, rhs, 1.0 / ni, 1.0 / nj, 1.7, eps, oits - 1)
            assert ret == oret2 == -1 and its == oits2 == oits - 1
            assert np.array_equal(s.download(m.P)[1:-1, 1:-1], po[1:-1, 1:-1])
    assert seen == {0, 1}, "the eps list must produce both an odd and an even iteration count"


@pytest.mark.parametrize("ni,nj", [(150, 40), (400, 36)])
def test_ieee_division_and_window_fallback(ni, nj):
    """Non power-of-two spacing (IEEE division by dx^2, dy^2 through the refined reciprocal) with operands outside the
    quotient window (denormals, +-0 blocks, near-overflow): the two-iteration pass repeats the task with real
    divisions; bits, the sign of zero included, equal the oracle's."""
    m = nsb()
    a, b = 1.0, 0.9
    P = m.make_params(i_max=ni, j_max=nj, a=a, b=b)
    with m.Solver(P, pass_k=2, sor_engine=m.SOR_STREAM, chunk_rows=16) as s:
        p, rhs = _rand(s.shape, 31, 0.5), _rand(s.shape, 32, 3.0)
        p[20:60, 5:30] = 0.0; p[30:40, 10:20] = -0.0; rhs[18:62, 3:32] = 0.0
        p[25, 12] = 1e-310; p[45, 20] = -3e-320; p[50, 8] = 4.9e-324
        p[100:110, 10:20] *= 4e306 * (a / ni) ** 2
        s.upload(m.P, p); s.upload(m.RHS, rhs)
        with np.errstate(all="ignore"):
            for n in (2, 4, 3):
                s.sor_iterations(n)
                for _ in range(n):
                    ob.lib().orc_SOR_rb_iter(ob._ptr(p), ni, nj, a / ni, b / nj, ob._ptr(rhs), P.omega)
                got = s.download(m.P)
                assert np.all(np.isfinite(p[1:-1, 1:-1]))
                assert np.array_equal(got[1:-1, 1:-1].view(np.int64), p[1:-1, 1:-1].view(np.int64))


@pytest.mark.parametrize("ni,nj", [(1000, 700), (2048, 1536)])
def test_one_and_two_iterations_per_pass_agree(ni, nj):
    """Beyond the oracle's comfortable size: one iteration per pass (the default) and the two-iteration pass give the same p bits, the same residual sum bits, and the same converged iteration count."""
    m = nsb()
    outs = []
    for k in (1, 2):
        P = m.make_params(i_max=ni, j_max=nj, epsilon=0.2, max_it=400)
        with m.Solver(P, pass_k=k, sor_engine=m.SOR_STREAM, chunk_rows=48) as s:
            x = (np.arange(ni + 2) - 0.5) / ni
            y = (np.arange(nj + 2) - 0.5) / nj
            p0 = _rand(s.shape, 41, 0.5)
            r0 = np.ascontiguousarray(np.cos(np.pi * x)[:, None] * np.cos(2 * np.pi * y)[None, :])
            s.upload(m.P, p0); s.upload(m.RHS, r0)
            r9 = s.sor_iterations(9)
            a = s.download(m.P)[1:-1, 1:-1].copy()
            s.upload(m.P, np.zeros(s.shape))          # threshold 0.2 * 1.5 against an initial residual of ~0.5: converges
            ret, its, res, n0 = s.SOR()
            outs.append((a, r9, ret, its, res, s.download(m.P)[1:-1, 1:-1].copy()))
    assert np.array_equal(outs[0][0], outs[1][0]) and outs[0][1] == outs[1][1]
    assert outs[0][2:5] == outs[1][2:5]
    assert np.array_equal(outs[0][5], outs[1][5])
