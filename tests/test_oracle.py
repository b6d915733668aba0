"""Pins the CPU oracle (oracle/ns_oracle.c) to the reference.

 1. the reference's own known-answer vectors tests/der{,2,3,4}.txt (reference tests/stencils.c:33-71),
    committed gzip'd under tests/golden/;
 2. the UNMODIFIED reference serial functions compiled in place into oracle/_ref/libnsref.so
    (bit-for-bit, function by function and end to end) -- skipped where oracle/_ref is absent;
 3. committed fixtures generated from that reference build (tests/golden/make_golden.py), so the
    pin survives on machines without /root/reference or oracle/_ref.
"""
import ctypes as C
import gzip
import json
import math
import os
import subprocess

import numpy as np
import pytest

from oracle import binding as ob

GOLD = os.path.join(os.path.dirname(__file__), "golden")
needs_ref = pytest.mark.skipif(not ob.have_ref(), reason="oracle/_ref not built")


def _stencil_tables():
    """Restates reference tests/stencils.c:33-49 with the oracle's stencils."""
    L = ob.lib()
    n = 100
    ii = np.arange(n, dtype=np.float64)
    u = np.ascontiguousarray(np.repeat(np.sin(0.1 * ii)[:, None], n, axis=1))
    v = np.ascontiguousarray(np.repeat(np.cos(0.1 * ii)[None, :], n, axis=0))
    out = [np.zeros((n, n)) for _ in range(4)]
    up, vp = ob._ptr(u), ob._ptr(v)
    for i in range(1, n - 1):
        for j in range(1, n - 1):
            out[0][i, j] = L.orc_du2_dx(up, n, i, j, 0.1, 1.0)
            out[1][i, j] = L.orc_dv2_dy(vp, n, i, j, 0.1, 1.0)
            out[2][i, j] = L.orc_duv_dx(up, vp, n, i, j, 0.1, 1.0)
            out[3][i, j] = L.orc_duv_dy(up, vp, n, i, j, 0.1, 1.0)
    return out


def _fmt(tab):
    # reference tests/stencils.c:60-71: rows j, columns i, "%.5f " and a newline per row
    n = tab.shape[0]
    return "".join("".join("%.5f " % tab[i, j] for i in range(n)) + "\n" for j in range(n))


@pytest.mark.parametrize("k,name", [(0, "der.txt"), (1, "der2.txt"), (2, "der3.txt"), (3, "der4.txt")])
def test_stencils_match_reference_golden_vectors(k, name):
    with gzip.open(os.path.join(GOLD, name + ".gz"), "rt") as fh:
        want = fh.read()
    got = _fmt(_stencil_tables()[k])
    assert got == want      # byte-identical text


@needs_ref
def test_reference_stencil_program_reproduces_golden(tmp_path):
    subprocess.check_call([ob.REF_STENCILS], cwd=tmp_path)
    for name in ("der.txt", "der2.txt", "der3.txt", "der4.txt"):
        with gzip.open(os.path.join(GOLD, name + ".gz"), "rb") as fh:
            assert fh.read() == (tmp_path / name).read_bytes()


def _rand_fields(i_max, j_max, seed):
    rng = np.random.default_rng(seed)
    return [np.ascontiguousarray(rng.uniform(-1, 1, (i_max + 2, j_max + 2))) for _ in range(3)]


@needs_ref
@pytest.mark.parametrize("i_max,j_max", [(8, 8), (17, 9), (5, 24)])
def test_functions_bit_exact_vs_reference(i_max, j_max):
    L, R = ob.lib(), ob.ref()
    u, v, p = _rand_fields(i_max, j_max, 1)
    dx, dy, dt, gamma, Re = 1.0 / i_max, 0.7 / j_max, 0.0123, 0.37, 123.0
    # max_mat
    assert L.orc_max_mat(i_max, j_max, ob._ptr(u)) == R.max_mat(i_max, j_max, ob.rows(u))
    # boundary conditions, every side, inflow values non-trivial
    for side in (ob.LEFT, ob.RIGHT, ob.BOTTOM, ob.TOP):
        u1, v1, u2, v2 = u.copy(), v.copy(), u.copy(), v.copy()
        L.orc_set_inflow(i_max, j_max, ob._ptr(u1), ob._ptr(v1), side, 0.75, -0.25)
        R.set_inflow(i_max, j_max, ob.rows(u2), ob.rows(v2), side, 0.75, -0.25)
        assert np.array_equal(u1, u2) and np.array_equal(v1, v2)
    # F, G
    F1, G1, F2, G2 = (np.zeros_like(u) for _ in range(4))
    L.orc_FG(ob._ptr(F1), ob._ptr(G1), ob._ptr(u), ob._ptr(v), i_max, j_max, Re, 0.1, -0.2, dt, dx, dy, gamma)
    R.FG(ob.rows(F2), ob.rows(G2), ob.rows(u), ob.rows(v), i_max, j_max, Re, 0.1, -0.2, dt, dx, dy, gamma)
    assert np.array_equal(F1, F2) and np.array_equal(G1, G2)
    # RHS (main.c:116-120 is inline in the reference main: restated here in numpy, same order)
    RHS1 = np.zeros_like(u)
    L.orc_RHS(ob._ptr(RHS1), ob._ptr(F1), ob._ptr(G1), i_max, j_max, dt, dx, dy)
    RHS2 = np.zeros_like(u)
    RHS2[1:-1, 1:-1] = 1.0 / dt * ((F1[1:-1, 1:-1] - F1[:-2, 1:-1]) / dx + (G1[1:-1, 1:-1] - G1[1:-1, :-2]) / dy)
    assert np.array_equal(RHS1, RHS2)
    # lexicographic SOR: same p, same return code
    pa, pb = p.copy(), p.copy()
    resb = np.zeros_like(p)
    ret, its, r, n0 = ob.sor_lex(pa, RHS1 * 1e-3, dx, dy, 1.7, 1e-6, 300)
    retb = R.SOR(ob.rows(pb), i_max, j_max, dx, dy, ob.rows(resb), ob.rows(np.ascontiguousarray(RHS1 * 1e-3)), 1.7, 1e-6, 300)
    assert ret == retb and np.array_equal(pa, pb)
    assert L.orc_L2(ob._ptr(pa), i_max, j_max) == R.L2(ob.rows(pb), i_max, j_max)
    # velocity update via the reference's dp_dx/dp_dy
    ua, va = u.copy(), v.copy()
    L.orc_update_uv(ob._ptr(ua), ob._ptr(va), ob._ptr(F1), ob._ptr(G1), ob._ptr(pa), i_max, j_max, dt, dx, dy)
    ub, vb = u.copy(), v.copy()
    prow = ob.rows(pb)
    for i in range(1, i_max + 1):
        for j in range(1, j_max + 1):
            if i <= i_max - 1:
                ub[i, j] = F1[i, j] - dt * R.dp_dx(prow, i, j, dx)
            if j <= j_max - 1:
                vb[i, j] = G1[i, j] - dt * R.dp_dy(prow, i, j, dy)
    assert np.array_equal(ua, ub) and np.array_equal(va, vb)


def _write_param_file(path, P):
    vals = [P.problem, P.f, P.i_max, P.j_max, P.a, P.b, P.T, P.Re, P.g_x, P.g_y, P.tau, P.omega,
            P.epsilon, P.max_it, P.n_print]
    with open(path, "w") as fh:
        for x in vals:
            fh.write(("%d" % x if isinstance(x, int) else repr(float(x))) + "   # comment\n")


@needs_ref
@pytest.mark.parametrize("kw", [dict(i_max=32, j_max=32, T=1.5), dict(i_max=24, j_max=40, T=1.2, problem=2, f=3.0, b=1.5)])
def test_whole_run_matches_reference_binary(tmp_path, kw):
    P = ob.make_params(**kw)
    pf = tmp_path / "case.in"
    _write_param_file(pf, P)
    out = subprocess.run([ob.REF_SERIAL, str(pf)], capture_output=True, text=True, check=True)
    got = ob.run(P, sor_order=0)
    want = "U-CENTER: %.6f\nV-CENTER: %.6f\n" % (got["u"][P.i_max // 2, P.j_max // 2], got["v"][P.i_max // 2, P.j_max // 2])
    assert out.stdout == want
    # parameter parser agrees with the reference's init()
    P2 = ob.read_params(str(pf))
    for k, _ in ob.Params._fields_:
        assert getattr(P2, k) == getattr(P, k), k


def test_committed_fixture_serial_and_redblack():
    """Fixtures were written by tests/golden/make_golden.py from oracle/_ref (serial order) and, for
    the red-black order, from the oracle after it passed the checks above."""
    with open(os.path.join(GOLD, "runs.json")) as fh:
        cases = json.load(fh)
    assert cases
    for c in cases:
        P = ob.make_params(**c["params"])
        got = ob.run(P, sor_order=c["sor_order"])
        assert got["steps"] == c["steps"]
        assert [s["sor_iters"] for s in got["log"]] == c["sor_iters"]
        assert [s["dt"].hex() for s in got["log"]] == c["dt_hex"]
        ic, jc = P.i_max // 2, P.j_max // 2
        assert float(got["u"][ic, jc]).hex() == c["u_center_hex"]
        assert float(got["v"][ic, jc]).hex() == c["v_center_hex"]
        ref = np.load(os.path.join(GOLD, c["fields"]))
        for k in ("u", "v", "p"):
            assert np.array_equal(got[k], ref[k]), (c["name"], k)


def test_redblack_ghost_free_statement():
    """the red-black iteration can be stated without ghost storage.  Checks the
    statement the CUDA kernels implement against the ghost-array oracle, incl. odd sizes."""
    L = ob.lib()
    for (ni, nj) in [(6, 6), (7, 5), (9, 12)]:
        rng = np.random.default_rng(5)
        p = rng.uniform(-1, 1, (ni + 2, nj + 2))
        rhs = rng.uniform(-1, 1, (ni + 2, nj + 2))
        dx, dy, om = 1.0 / ni, 1.3 / nj, 1.7
        dx2, dy2 = dx * dx, dy * dy
        q = p.copy()
        r_or = L.orc_SOR_rb_iter(ob._ptr(q), ni, nj, dx, dy, ob._ptr(rhs), om)
        w = p.copy()
        c = om / (2.0 * (1.0 / dx2 + 1.0 / dy2))
        lvl1 = None
        for color in (0, 1):
            src = w.copy()
            for i in range(1, ni + 1):
                for j in range(1, nj + 1):
                    if (i + j) % 2 != color:
                        continue
                    e = src[i + 1, j] if i < ni else src[i, j]
                    ww = src[i - 1, j] if i > 1 else src[i, j]
                    n = src[i, j + 1] if j < nj else src[i, j]
                    s = src[i, j - 1] if j > 1 else src[i, j]
                    w[i, j] = (1.0 - om) * src[i, j] + c * ((e + ww) / dx2 + (n + s) / dy2 - rhs[i, j])
            if color == 0:
                lvl1 = w.copy()
        assert np.array_equal(w[1:-1, 1:-1], q[1:-1, 1:-1])
        acc = 0.0
        for i in range(1, ni + 1):
            for j in range(1, nj + 1):
                e = w[i + 1, j] if i < ni else lvl1[i, j]
                ww = w[i - 1, j] if i > 1 else lvl1[i, j]
                n = w[i, j + 1] if j < nj else lvl1[i, j]
                s = w[i, j - 1] if j > 1 else lvl1[i, j]
                r = (e - 2.0 * w[i, j] + ww) / dx2 + (n - 2.0 * w[i, j] + s) / dy2 - rhs[i, j]
                acc += r * r
        assert math.sqrt(acc / (ni * nj)) == r_or


@needs_ref
def test_oracle_output_matches_reference_output(tmp_path, monkeypatch):
    """io.c:61-120: byte-identical files, on jagged grids shaped like the reference allocates."""
    R = ob.ref()
    ni, nj = 7, 5
    rng = np.random.default_rng(3)
    u, v, p = (np.ascontiguousarray(rng.uniform(-2, 2, (ni + 2, nj + 2))) for _ in range(3))
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    monkeypatch.chdir(tmp_path)     # the reference builds file names in char[64] (io.c:65-76): keep the prefix short
    assert ob.lib().orc_output(ni, nj, ob._ptr(u), ob._ptr(v), ob._ptr(p), 0.123456, 1.0, 2.0, b"a/x") == 0
    uj = np.ascontiguousarray(u[:ni + 1])
    R.output(ni, nj, ob.rows(uj), ob.rows(v), ob.rows(p), 0.123456, 1.0, 2.0, b"b/x")
    for suf in ("_u.txt", "_v.txt", "_p.txt"):
        assert (tmp_path / "a" / ("x" + suf)).read_bytes() == (tmp_path / "b" / ("x" + suf)).read_bytes()
    assert ob.lib().orc_output(ni, nj, ob._ptr(u), ob._ptr(v), ob._ptr(p), 0.0, 1.0, 1.0, b"/nonexistent_dir/x") == -1


def test_redblack_omp_variant_equals_sequential():
    ni, nj = 40, 33
    rng = np.random.default_rng(8)
    p = np.ascontiguousarray(rng.uniform(-1, 1, (ni + 2, nj + 2)))
    rhs = np.ascontiguousarray(rng.uniform(-1, 1, (ni + 2, nj + 2)))
    q = p.copy()
    a = ob.sor_rb(p, rhs, 1.0 / ni, 1.0 / nj, 1.7, 1e-30, 25)
    b = ob.sor_rb_omp(q, rhs, 1.0 / ni, 1.0 / nj, 1.7, 1e-30, 25, 4)
    assert a[0] == b[0] == -1 and np.array_equal(p, q)
    assert abs(a[2] - b[2]) <= 1e-12 * a[2]
