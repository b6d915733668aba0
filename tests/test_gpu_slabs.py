"""y-slab decomposition on ONE GPU: the slabs of a group share device 0 and run in lock step
on one stream, so every peer store, flag and decomposition-independent reduction of the multi-GPU path
is exercised without a second device.  The bar: 2, 3 and 4 slabs give the bits of the undivided run
(fields, residuals, iteration counts, dt sequence) and of the CPU oracle."""
import numpy as np
import pytest

from oracle import binding as ob
from util import nsb

pytestmark = pytest.mark.gpu


def _rand(shape, seed):
    return np.ascontiguousarray(np.random.default_rng(seed).uniform(-1, 1, shape))


def _params(m, **kw):
    base = dict(a=1.0, b=1.0, T=1.0, Re=1000.0, max_it=300, epsilon=1e-4)
    base.update(kw)
    return m.make_params(**base)


@pytest.mark.parametrize("world", [2, 3, 4])
@pytest.mark.parametrize("ni,nj,chunk", [(150, 96, 16), (64, 130, 16), (300, 64, 16)])
def test_sor_iterations_equal_single_slab(world, ni, nj, chunk):
    m = nsb()
    if (nj + chunk - 1) // chunk < world:
        pytest.skip("fewer bands than slabs")
    P = _params(m, i_max=ni, j_max=nj)
    p0, r0 = _rand((ni + 2, nj + 2), 5), _rand((ni + 2, nj + 2), 6)
    with m.Solver(P, sor_engine=m.SOR_STREAM, chunk_rows=chunk) as s:
        s.upload(m.P, p0); s.upload(m.RHS, r0)
        res1 = [s.sor_iterations(1) for _ in range(3)] + [s.sor_iterations(4)]
        want = s.download(m.P)
    with m.Group(P, [0] * world, chunk_rows=chunk) as g:
        assert sum(x["nj"] for x in g.slabs()) == nj
        g.upload(m.P, p0); g.upload(m.RHS, r0)
        resg = [g.sor_iterations(1) for _ in range(3)] + [g.sor_iterations(4)]
        got = g.download(m.P)
    assert np.array_equal(got[1:-1, 1:-1], want[1:-1, 1:-1])
    assert resg == res1                      # same bands, same order: the residual has the same bits


@pytest.mark.parametrize("world", [2, 4])
def test_whole_run_equals_single_slab_and_oracle(world):
    m = nsb()
    kw = dict(i_max=96, j_max=80, T=1.0, max_it=400)
    P = _params(m, **kw)
    with m.Solver(P, sor_engine=m.SOR_STREAM, chunk_rows=16) as s:
        log1, _ = s.run(max_steps=4)
        f1 = [s.download(f) for f in (m.U, m.V, m.P)]
        c1 = s.center()
    with m.Group(P, [0] * world, chunk_rows=16) as g:
        logg, _ = g.run(max_steps=4)
        fg = [g.download(f) for f in (m.U, m.V, m.P)]
        cg = g.center()
    assert [x["sor_iters"] for x in logg] == [x["sor_iters"] for x in log1]
    assert [x["dt"] for x in logg] == [x["dt"] for x in log1]
    assert [x["res"] for x in logg] == [x["res"] for x in log1]
    assert [x["norm_p"] for x in logg] == [x["norm_p"] for x in log1]
    for a, b in zip(fg, f1):
        assert np.array_equal(a[1:-1, 1:-1], b[1:-1, 1:-1])
    assert np.array_equal(fg[0], f1[0]) and np.array_equal(fg[1], f1[1])      # u, v incl. boundary values
    assert cg == c1
    o = ob.run(ob.make_params(**{k: getattr(P, k) for k, _ in ob.Params._fields_}), sor_order=1, max_steps=4)
    assert [x["sor_iters"] for x in logg] == [x["sor_iters"] for x in o["log"]]
    assert [x["dt"] for x in logg] == [x["dt"] for x in o["log"]]
    assert np.array_equal(fg[0], o["u"]) and np.array_equal(fg[1], o["v"])
    assert np.array_equal(fg[2][1:-1, 1:-1], o["p"][1:-1, 1:-1])


def test_slabs_at_default_chunk_and_nonsquare_box():
    """124-row bands (what the planner picks for 8192-row slabs), ragged last band, dx != dy, gravity, oscillating lid;
    then the band height the library plans for THIS decomposition (nsb_sor_preplan_slabs): other bands, so the residual
    is summed in another order -- the fields are the same bits as long as the iteration counts agree."""
    m = nsb()
    P = _params(m, i_max=130, j_max=500, a=1.0, b=2.5, T=0.2, max_it=60, problem=2, f=3.0, g_x=0.3, g_y=-0.2)
    with m.Solver(P, sor_engine=m.SOR_STREAM, chunk_rows=124) as s:
        log1, _ = s.run(max_steps=3)
        f1 = [s.download(f) for f in (m.U, m.V, m.P)]
    with m.Group(P, [0, 0, 0], chunk_rows=124) as g:
        assert [x["nj"] for x in g.slabs()] == [248, 124, 128]
        logg, _ = g.run(max_steps=3)
        fg = [g.download(f) for f in (m.U, m.V, m.P)]
    assert [(x["sor_iters"], x["dt"], x["res"]) for x in logg] == [(x["sor_iters"], x["dt"], x["res"]) for x in log1]
    for a, b in zip(fg, f1):
        assert np.array_equal(a[1:-1, 1:-1], b[1:-1, 1:-1])
    with m.Group(P, [0, 0, 0]) as g:
        nj = [x["nj"] for x in g.slabs()]
        assert sum(nj) == 500 and min(nj) >= 16
        logd, _ = g.run(max_steps=3)
        fd = [g.download(f) for f in (m.U, m.V, m.P)]
    assert [x["sor_iters"] for x in logd] == [x["sor_iters"] for x in log1]
    assert np.allclose([x["res"] for x in logd], [x["res"] for x in log1], rtol=1e-12, atol=0.0)
    for a, b in zip(fd, f1):
        assert np.array_equal(a[1:-1, 1:-1], b[1:-1, 1:-1])


def test_too_many_slabs_is_an_error():
    m = nsb()
    P = _params(m, i_max=32, j_max=32)
    with pytest.raises(m.NsbError):
        m.Group(P, [0] * 4, chunk_rows=16)       # 2 bands, 4 slabs
