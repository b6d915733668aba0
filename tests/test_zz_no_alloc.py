"""nsb_max_mat_noalloc (include/nsb200.h section C; csrc/noalloc.cu): the allocation-free max_mat against the existing
entry point nsb_max_mat and the oracle, through the C-ABI.

No counterpart in the reference.  The file name sorts last on purpose:
a failure here must not hide the parity tests.

Bar: max is an exact, order-independent operation -- the two entry points must return the SAME double (==), which is
also what tests/test_gpu_parity.py asks of nsb_max_mat against the oracle.
Device buffers are made by torch in the TEST (the caller owns inputs and outputs); the entry point itself allocates
nothing.  Two calls in a row share the static scratch array: the second call uses a smaller grid with a smaller maximum,
so a stale partial or a ticket left over from the first call would show.
"""
import ctypes as C

import numpy as np
import pytest

from oracle import binding as ob
from util import nsb

pytestmark = pytest.mark.gpu

ERR_ARG = -2          # NSB_ERR_ARG, include/nsb200.h


def _grid(ni, nj, seed, scale, shift):
    rng = np.random.default_rng(seed)
    a = np.ascontiguousarray(rng.uniform(-1.0, 1.0, (ni + 2, nj + 2)) * scale + shift)
    # ghost cells must not count (only m[0][0] does): make them larger than any interior value
    a[0, 1:] = a[-1, :] = 9e9
    a[1:, 0] = a[:, -1] = 9e9
    return a


def _noalloc(L, torch, a, ni, nj, out, stride=None):
    d = torch.from_numpy(a).cuda()
    stride = a.shape[1] if stride is None else stride
    rc = L.nsb_max_mat_noalloc(C.c_void_p(d.data_ptr()), stride, ni, nj, C.c_void_p(out.data_ptr()), None)
    assert rc == 0, L.nsb_last_error(None)
    torch.cuda.synchronize()
    return float(out.cpu()[0])


def test_no_alloc_max_mat_equals_existing_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    m = nsb()
    L = m.lib()
    torch.cuda.set_device(0)
    out = torch.full((1,), -7.0, dtype=torch.float64, device="cuda")
    # (grid, seed, scale, shift): call 2 is smaller and everywhere BELOW call 1's maximum; call 3 has m[0][0] on top
    cases = [(1000, 700, 1, 5.0, 0.0), (37, 211, 2, 1.0, -3.0), (64, 64, 3, 1.0, 0.0), (1, 1, 4, 1.0, 0.0)]
    seen = []
    for k, (ni, nj, seed, scale, shift) in enumerate(cases):
        a = _grid(ni, nj, seed, scale, shift)
        if k == 2:
            a[0, 0] = 2.5                                          # the seed of the scan wins (io.c:124)
        else:
            a[0, 0] = -1e9
        want = L.nsb_max_mat(ni, nj, m.rows(a))                    # the existing entry point (host grids, its own context)
        got = _noalloc(L, torch, a, ni, nj, out)
        assert got == want, (k, got, want)
        assert got == ob.lib().orc_max_mat(ni, nj, ob._ptr(a))
        assert got == max(a[0, 0], a[1:-1, 1:-1].max())
        got2 = _noalloc(L, torch, a, ni, nj, out)                  # same call again: same value
        assert got2 == want
        seen.append(got)
    assert seen[1] < seen[0]                                       # the second grid's maximum is below the first's
    L.nsb_host_api_reset()


def test_no_alloc_row_stride_and_back_to_back_launches_on_one_stream():
    """Padded rows (row_stride > j_max + 2) and the v shape (row_stride = j_max + 1); several launches queued on one
    stream without a host synchronisation in between (ordered calls share the scratch safely)."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    m = nsb()
    L = m.lib()
    torch.cuda.set_device(0)
    ni, nj = 300, 129
    a = _grid(ni, nj, 11, 2.0, 0.5)
    a[0, 0] = -1e9
    wide = np.full((ni + 2, nj + 40), 8e9)
    wide[:, :nj + 2] = a
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    assert _noalloc(L, torch, wide, ni, nj, out, stride=nj + 40) == L.nsb_max_mat(ni, nj, m.rows(a))
    v = np.ascontiguousarray(a[:, :nj + 1])                        # rows of j_max + 1 doubles (memory.c:15)
    assert _noalloc(L, torch, v, ni, nj, out, stride=nj + 1) == L.nsb_max_mat(ni, nj, m.rows(v))
    # queued launches: three grids, three result slots, one synchronisation
    grids = [_grid(200 + 50 * k, 90 + 7 * k, 20 + k, 1.0 + k, -float(k)) for k in range(3)]
    for g in grids:
        g[0, 0] = -1e9
    dev = [torch.from_numpy(g).cuda() for g in grids]
    outs = torch.zeros(3, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    for k, g in enumerate(grids):
        rc = L.nsb_max_mat_noalloc(C.c_void_p(dev[k].data_ptr()), g.shape[1], g.shape[0] - 2, g.shape[1] - 2,
                                   C.c_void_p(outs.data_ptr() + 8 * k), None)
        assert rc == 0
    torch.cuda.synchronize()
    got = outs.cpu().numpy()
    for k, g in enumerate(grids):
        assert got[k] == L.nsb_max_mat(g.shape[0] - 2, g.shape[1] - 2, m.rows(g))
    L.nsb_host_api_reset()


def test_no_alloc_rejects_sizes_beyond_the_static_scratch():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    m = nsb()
    L = m.lib()
    buf = torch.zeros(16, dtype=torch.float64, device="cuda")
    p = C.c_void_p(buf.data_ptr())
    assert L.nsb_max_mat_noalloc(p, 16386, 16385, 8, p, None) == ERR_ARG
    assert b"exceeds the static scratch limit" in L.nsb_last_error(None)
    assert L.nsb_max_mat_noalloc(p, 16386, 8, 16385, p, None) == ERR_ARG
    assert L.nsb_max_mat_noalloc(p, 4, 2, 4, p, None) == ERR_ARG          # row_stride < j_max + 1
    assert L.nsb_max_mat_noalloc(None, 4, 2, 2, p, None) == ERR_ARG
    assert L.nsb_max_mat_noalloc(p, 4, 0, 2, p, None) == ERR_ARG
    torch.cuda.synchronize()
