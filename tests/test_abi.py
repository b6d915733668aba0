"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/nsb200.h declares, binds with the signatures the python mirror uses, and refuses to work
without a GPU instead of falling back to a CPU path.  No compute calls here."""
import ctypes as C
import os
import re
import subprocess

import pytest

from util import ROOT, nsb


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "nsb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nsb_[A-Za-z0-9_]+)\s*\(", text)))


def test_header_declares_expected_surface():
    syms = _declared_symbols()
    for want in ("nsb_create", "nsb_destroy", "nsb_max_uv", "nsb_set_inflow", "nsb_fg", "nsb_rhs",
                 "nsb_sor", "nsb_update_uv", "nsb_step", "nsb_output", "nsb_SOR", "nsb_FG", "nsb_timestep"):
        assert want in syms


def test_library_exports_every_declared_symbol():
    m = nsb()
    assert os.path.exists(m.LIB_PATH), "libnsb200.so not built (run __graft_entry__.build())"
    L = C.CDLL(m.LIB_PATH)
    missing = [s for s in _declared_symbols() if not hasattr(L, s)]
    assert not missing, missing
    assert L.nsb_abi_version() == 1


def test_python_mirror_covers_the_header():
    m = nsb()
    assert sorted(m.ABI) == _declared_symbols()


def test_struct_layouts_match_header_order():
    m = nsb()
    assert [f for f, _ in m.Params._fields_] == ["problem", "f", "i_max", "j_max", "a", "b", "T", "Re", "g_x",
                                                 "g_y", "tau", "omega", "epsilon", "max_it", "n_print"]
    assert C.sizeof(m.Params) == 104
    assert C.sizeof(m.StepInfo) == 64


def test_param_parser_matches_reference_grammar(tmp_path):
    m = nsb()
    f = tmp_path / "p.in"
    f.write_text("2   # problem\n0.5 # f\n33\t# i_max\n17 #j\n1.5 # a\n2.0\n0.25 # T comes BEFORE Re\n400.0\n"
                 "0.1\n-0.2\n0.9\n1.6\n1e-5\n1234\n7 # n_print")
    P = m.read_params(str(f))
    assert (P.problem, P.f, P.i_max, P.j_max, P.a, P.b, P.T, P.Re) == (2, 0.5, 33, 17, 1.5, 2.0, 0.25, 400.0)
    assert (P.g_x, P.g_y, P.tau, P.omega, P.epsilon, P.max_it, P.n_print) == (0.1, -0.2, 0.9, 1.6, 1e-5, 1234, 7)
    with pytest.raises(FileNotFoundError):
        m.read_params(str(tmp_path / "nope.in"))


def test_dt_gamma_is_host_arithmetic_in_reference_order():
    # nsb_dt_gamma needs a context; without a GPU creation must fail loudly (next test), so the
    # formula is checked on the GPU box in test_gpu_parity.  Here: the oracle's formula vs numpy.
    from oracle import binding as ob
    P = ob.make_params(i_max=48, j_max=40, a=1.0, b=0.8, Re=321.0, tau=0.7)
    dx, dy = P.a / P.i_max, P.b / P.j_max
    dt, g = C.c_double(), C.c_double()
    ob.lib().orc_dt_gamma(C.byref(P), dx, dy, 0.31, -0.12, C.byref(dt), C.byref(g))
    want = 0.7 * min(321.0 / 2.0 / (1.0 / dx / dx + 1.0 / dy / dy), dx / 0.31, dy / 0.12)
    assert dt.value == want
    assert g.value == max(0.31 * want / dx, -0.12 * want / dy)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    m = nsb()
    with pytest.raises(m.NsbError) as e:
        m.Solver(m.make_params(i_max=8, j_max=8))
    assert "no CPU path" in str(e.value) or "CUDA" in str(e.value)


def test_driver_binary_fails_loudly_without_gpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    exe = os.path.join(ROOT, "build", "parallel")
    if not os.path.exists(exe):
        pytest.skip("driver not built")
    f = tmp_path / "p.in"
    f.write_text("1\n1\n8\n8\n1\n1\n0.01\n100\n0\n0\n1\n1.7\n1e-4\n10\n1\n")
    out = subprocess.run([exe, str(f)], capture_output=True, text=True)
    assert out.returncode != 0 and "CUDA" in out.stderr


def test_product_never_touches_the_oracle():
    """The product path must not import, link or call anything under oracle/."""
    pkg = os.path.join(ROOT, "navierstokes-parallel_b200")
    for base, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".c", "Makefile")):
                text = open(os.path.join(base, fn), errors="ignore").read()
                for pat in ("import oracle", "from oracle", "liboracle", "ns_oracle", "oracle/", "orc_", "libnsref"):
                    assert pat not in text, (os.path.join(base, fn), pat)
    text = open(os.path.join(ROOT, "driver", "main.c")).read()
    assert "oracle" not in text and "orc_" not in text
    out = subprocess.run(["ldd", os.path.join(pkg, "libnsb200.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out and "nsref" not in out


def test_chunk_height_model_of_the_persistent_sweep():
    """nsb_plan_chunk_rows (host arithmetic, sor_stream.cu): the chunk heights the plan picks on a B200 (148 SMs x 3 blocks
    x 4 warps = 1776 resident warps) are, within ~3 % of the time per iteration, the optima measured in
    profiles/r2j_chunk_sweep.txt (reciprocal product) and profiles/r1h_chunk_sweep2.txt (IEEE division); heights are
    even, 10..128, and grow with the grid (up to the quantisation of whole chunks and blocks of 8 steps)."""
    L = nsb().lib()
    f = L.nsb_plan_chunk_rows
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_int, C.c_longlong, C.c_int]
    sizes = (2048, 3072, 4096, 5120, 6144, 7168, 8192, 16384)
    assert [f(n, n, 1776, 1) for n in sizes] == [22, 26, 50, 74, 106, 122, 128, 122]      # reciprocal product
    assert [f(n, n, 1776, 0) for n in sizes[:6]] == [10, 18, 32, 48, 66, 90]              # IEEE division: more tasks per warp
    for recip in (0, 1):
        prev = 0
        for n in range(64, 20000, 448):
            r = f(n, n, 1776, recip)
            assert 10 <= r <= 128 and r % 2 == 0 and r >= prev - 12
            prev = r
    assert f(8192, 8, 1776, 1) == 10 and f(5, 250, 1776, 1) == 10 and f(100, 100, 0, 1) >= 10


def test_joint_plan_of_iterations_per_pass_and_chunk_height():
    """nsb_plan_sor (host arithmetic): iterations per pass AND chunk height for a grid -- also what a y-slab decomposition
    plans its band height with, applied to one slab of ceil(j_max / world) rows (nsb_sor_preplan_slabs).  B200: 1776 / 1184
    resident warps with one / two iterations per pass.  Measured counterparts: profiles/r2n_k2_sweep.txt (squares),
    profiles/r3n_rect_probe.txt (slab shapes: 16384 x 2048 planned 109 us per iteration, fixed 124-row bands 134-138)."""
    L = nsb().lib()
    f = L.nsb_plan_sor
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_int, C.c_longlong, C.c_longlong, C.c_int, C.POINTER(C.c_int)]

    def plan(ni, nj, recip=1):
        rows = C.c_int()
        k = f(ni, nj, 1776, 1184, recip, C.byref(rows))
        return k, rows.value

    assert plan(8192, 8192) == (2, 124)             # the headline grid; also one slab of the weak-scaling runs
    assert plan(16384, 16384) == (2, 124)
    assert plan(16384, 2048) == (2, 68)             # 16384^2 on 8 GPUs
    assert plan(16384, 4096) == (2, 122)            # ... on 4
    assert plan(2048, 2048) == (1, 22) and plan(4096, 4096)[0] == 1      # one iteration per pass below ~5000^2
    assert plan(8191, 8191, recip=0)[0] == 1        # IEEE division: instruction bound, one iteration per pass
    for ni, nj in ((300, 8), (5, 5000), (20000, 64)):
        k, rows = plan(ni, nj)
        assert k in (1, 2) and 10 <= rows <= 128 and rows % 2 == 0


def test_output_formatter_equals_printf():
    """output.cu formats "%.5f" in exact integer arithmetic (host code of libnsb200.so, no GPU needed): it must produce
    printf's bytes -- round half to even on the exact binary value -- for field-like values, exact ties (k/64),
    zeros of both signs, denormals, huge values and infinities (the reference writes with fprintf, io.c:103-112)."""
    import ctypes as C
    import random
    import struct
    L = nsb().lib()
    f = L.nsb_debug_fmt5
    f.restype = C.c_int
    f.argtypes = [C.c_double, C.c_char_p]
    buf = C.create_string_buffer(512)
    rnd = random.Random(7)
    vals = [0.0, -0.0, 1.0, -1.0, 0.015625, 0.046875, -0.015625, 1e-9, -1e-9, 123456789.123456, 0.999995, 0.9999949999,
            0.000005, 5e-6, 4.9e-324, 1e-310, 7e13, 7.1e13, 1e15, 1e300, float("inf"), float("-inf"), 2.0 ** 46 - 0.001, 2.0]
    vals += [k / 64 for k in range(-200, 200)] + [k / 128 + 1000 for k in range(100)]
    vals += [rnd.uniform(-1, 1) * 10.0 ** rnd.choice([0, 0, 0, 1, 2, 3, -3, -5, -8, 5, 10]) for _ in range(100000)]
    for _ in range(50000):
        x = struct.unpack("<d", struct.pack("<Q", rnd.getrandbits(64)))[0]
        if x == x:                       # glibc prints the sign of a NaN, python does not: NaNs go through snprintf itself
            vals.append(x)
    for x in vals:
        n = f(x, buf)
        assert buf.value.decode() == "%.5f" % x and n == len(buf.value), repr(x)
