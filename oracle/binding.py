"""ctypes binding of the CPU checker (oracle/liboracle.so) and, when present, of the unmodified
reference serial build (oracle/_ref/libnsref.so).

TEST INFRASTRUCTURE ONLY -- imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product (navierstokes-parallel_b200) never imports it.

Grid convention here = the reference's: numpy arrays of shape (i_max+2, j_max+2), C order,
element [i, j]  (src/serial/memory.c:3-26).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_DIR = os.path.join(HERE, "_ref")
REF_LIB_PATH = os.path.join(REF_DIR, "libnsref.so")
REF_SERIAL = os.path.join(REF_DIR, "serial")
REF_STENCILS = os.path.join(REF_DIR, "stencils")

TOP, BOTTOM, LEFT, RIGHT = 0, 1, 2, 3
dbl_p = C.POINTER(C.c_double)
int_p = C.POINTER(C.c_int)


class Params(C.Structure):
    _fields_ = [("problem", C.c_int), ("f", C.c_double), ("i_max", C.c_int), ("j_max", C.c_int),
                ("a", C.c_double), ("b", C.c_double), ("T", C.c_double), ("Re", C.c_double),
                ("g_x", C.c_double), ("g_y", C.c_double), ("tau", C.c_double),
                ("omega", C.c_double), ("epsilon", C.c_double), ("max_it", C.c_int),
                ("n_print", C.c_int)]


class StepLog(C.Structure):
    _fields_ = [("t", C.c_double), ("dt", C.c_double), ("u_max", C.c_double),
                ("v_max", C.c_double), ("gamma", C.c_double), ("norm_p", C.c_double),
                ("res", C.c_double), ("sor_ret", C.c_int), ("sor_iters", C.c_int)]


def build(force: bool = False) -> None:
    """Compile liboracle.so (always possible) and oracle/_ref (only where /root/reference exists)."""
    if force or not os.path.exists(LIB_PATH) or \
            os.path.getmtime(LIB_PATH) < os.path.getmtime(os.path.join(HERE, "ns_oracle.c")):
        subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if os.path.isdir("/root/reference/src/serial") and (force or not os.path.exists(REF_LIB_PATH)):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def _ptr(a: np.ndarray):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(dbl_p)


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        L.orc_max_mat.restype = C.c_double
        L.orc_max_mat.argtypes = [C.c_int, C.c_int, dbl_p]
        L.orc_min3.restype = C.c_double
        L.orc_min3.argtypes = [C.c_double] * 3
        L.orc_dt_gamma.argtypes = [C.POINTER(Params), C.c_double, C.c_double, C.c_double,
                                   C.c_double, dbl_p, dbl_p]
        L.orc_set_inflow.argtypes = [C.c_int, C.c_int, dbl_p, dbl_p, C.c_int, C.c_double, C.c_double]
        L.orc_apply_bc.argtypes = [C.POINTER(Params), C.c_double, dbl_p, dbl_p]
        for name in ("orc_du2_dx", "orc_dv2_dy"):
            f = getattr(L, name)
            f.restype = C.c_double
            f.argtypes = [dbl_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double]
        for name in ("orc_duv_dy", "orc_duv_dx"):
            f = getattr(L, name)
            f.restype = C.c_double
            f.argtypes = [dbl_p, dbl_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double]
        L.orc_FG.argtypes = [dbl_p, dbl_p, dbl_p, dbl_p, C.c_int, C.c_int] + [C.c_double] * 7
        L.orc_RHS.argtypes = [dbl_p, dbl_p, dbl_p, C.c_int, C.c_int] + [C.c_double] * 3
        L.orc_L2.restype = C.c_double
        L.orc_L2.argtypes = [dbl_p, C.c_int, C.c_int]
        L.orc_SOR_lex.argtypes = [dbl_p, C.c_int, C.c_int, C.c_double, C.c_double, dbl_p, dbl_p,
                                  C.c_double, C.c_double, C.c_int, int_p, dbl_p, dbl_p]
        L.orc_SOR_rb.argtypes = [dbl_p, C.c_int, C.c_int, C.c_double, C.c_double, dbl_p,
                                 C.c_double, C.c_double, C.c_int, int_p, dbl_p, dbl_p]
        L.orc_SOR_rb_iter.restype = C.c_double
        L.orc_SOR_rb_iter.argtypes = [dbl_p, C.c_int, C.c_int, C.c_double, C.c_double, dbl_p,
                                      C.c_double]
        L.orc_SOR_rb_omp.argtypes = [dbl_p, C.c_int, C.c_int, C.c_double, C.c_double, dbl_p,
                                     C.c_double, C.c_double, C.c_int, int_p, dbl_p, C.c_int]
        L.orc_update_uv.argtypes = [dbl_p] * 5 + [C.c_int, C.c_int] + [C.c_double] * 3
        L.orc_run.argtypes = [C.POINTER(Params), C.c_int, C.c_int] + [dbl_p] * 6 + \
                             [C.POINTER(StepLog), C.c_int, dbl_p, dbl_p]
        L.orc_output.argtypes = [C.c_int, C.c_int, dbl_p, dbl_p, dbl_p, C.c_double, C.c_double,
                                 C.c_double, C.c_char_p]
        L.orc_read_params.argtypes = [C.c_char_p, C.POINTER(Params)]
        _lib = L
    return _lib


def make_params(**kw) -> Params:
    d = dict(problem=1, f=1.0, i_max=64, j_max=64, a=1.0, b=1.0, T=0.01, Re=1000.0, g_x=0.0,
             g_y=0.0, tau=1.0, omega=1.7, epsilon=1e-4, max_it=20000, n_print=1)
    d.update(kw)
    return Params(**d)


def read_params(path: str) -> Params:
    P = Params()
    if lib().orc_read_params(path.encode(), C.byref(P)) != 0:
        raise FileNotFoundError(path)
    return P


def zeros(P_or_imax, j_max=None) -> np.ndarray:
    if j_max is None:
        i_max, j_max = P_or_imax.i_max, P_or_imax.j_max
    else:
        i_max = P_or_imax
    return np.zeros((i_max + 2, j_max + 2), dtype=np.float64)


def run(P: Params, sor_order: int = 1, max_steps: int = 0, max_log: int = 100000):
    """Whole time loop.  Returns dict(u,v,p,F,G,RHS,log,t,steps,sor_seconds)."""
    L = lib()
    f = {k: zeros(P) for k in ("u", "v", "p", "F", "G", "RHS")}
    log = (StepLog * max_log)()
    t_end = C.c_double(0)
    sor_s = C.c_double(0)
    n = L.orc_run(C.byref(P), sor_order, max_steps, _ptr(f["u"]), _ptr(f["v"]), _ptr(f["p"]),
                  _ptr(f["F"]), _ptr(f["G"]), _ptr(f["RHS"]), log, max_log,
                  C.byref(t_end), C.byref(sor_s))
    if n < 0:
        raise ValueError("unknown problem type")
    f["log"] = [dict((k, getattr(log[i], k)) for k, _ in StepLog._fields_) for i in range(min(n, max_log))]
    f["t"] = t_end.value
    f["steps"] = n
    f["sor_seconds"] = sor_s.value
    return f


def sor_rb(p, RHS, dx, dy, omega, eps, max_it):
    L = lib()
    i_max, j_max = p.shape[0] - 2, p.shape[1] - 2
    it = C.c_int(0)
    r = C.c_double(0)
    n0 = C.c_double(0)
    ret = L.orc_SOR_rb(_ptr(p), i_max, j_max, dx, dy, _ptr(RHS), omega, eps, max_it,
                       C.byref(it), C.byref(r), C.byref(n0))
    return ret, it.value, r.value, n0.value


def sor_rb_omp(p, RHS, dx, dy, omega, eps, max_it, nthreads=0):
    L = lib()
    i_max, j_max = p.shape[0] - 2, p.shape[1] - 2
    it = C.c_int(0)
    r = C.c_double(0)
    ret = L.orc_SOR_rb_omp(_ptr(p), i_max, j_max, dx, dy, _ptr(RHS), omega, eps, max_it,
                           C.byref(it), C.byref(r), nthreads)
    return ret, it.value, r.value


def sor_lex(p, RHS, dx, dy, omega, eps, max_it):
    L = lib()
    i_max, j_max = p.shape[0] - 2, p.shape[1] - 2
    res = np.zeros_like(p)
    it = C.c_int(0)
    r = C.c_double(0)
    n0 = C.c_double(0)
    ret = L.orc_SOR_lex(_ptr(p), i_max, j_max, dx, dy, _ptr(res), _ptr(RHS), omega, eps, max_it,
                        C.byref(it), C.byref(r), C.byref(n0))
    return ret, it.value, r.value, n0.value


# ------------------------------------------------------------------ reference build (oracle/_ref)

def have_ref() -> bool:
    return os.path.exists(REF_LIB_PATH)


_ref = None


def ref():
    """The UNMODIFIED reference serial functions (double** grids)."""
    global _ref
    if _ref is None:
        build()
        R = C.CDLL(REF_LIB_PATH)
        pp = C.POINTER(dbl_p)
        R.max_mat.restype = C.c_double
        R.max_mat.argtypes = [C.c_int, C.c_int, pp]
        R.set_inflow.argtypes = [C.c_int, C.c_int, pp, pp, C.c_int, C.c_double, C.c_double]
        R.set_noslip.argtypes = [C.c_int, C.c_int, pp, pp, C.c_int]
        R.FG.argtypes = [pp, pp, pp, pp, C.c_int, C.c_int] + [C.c_double] * 7
        R.SOR.argtypes = [pp, C.c_int, C.c_int, C.c_double, C.c_double, pp, pp, C.c_double,
                          C.c_double, C.c_int]
        R.L2.restype = C.c_double
        R.L2.argtypes = [pp, C.c_int, C.c_int]
        R.dp_dx.restype = C.c_double
        R.dp_dx.argtypes = [pp, C.c_int, C.c_int, C.c_double]
        R.dp_dy.restype = C.c_double
        R.dp_dy.argtypes = [pp, C.c_int, C.c_int, C.c_double]
        for name in ("du2_dx", "duv_dy", "dv2_dy", "duv_dx"):
            f = getattr(R, name)
            f.restype = C.c_double
            f.argtypes = [pp, pp, C.c_int, C.c_int, C.c_double, C.c_double]
        R.output.argtypes = [C.c_int, C.c_int, pp, pp, pp, C.c_double, C.c_double, C.c_double,
                             C.c_char_p]
        R.init.argtypes = [int_p, dbl_p, int_p, int_p] + [dbl_p] * 9 + [int_p, int_p, C.c_char_p]
        _ref = R
    return _ref


def rows(a: np.ndarray):
    """double** view of a C-contiguous 2-D array (row pointers), as the reference's grids."""
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"] and a.ndim == 2
    n = a.shape[0]
    tab = (dbl_p * n)()
    base = a.ctypes.data
    stride = a.strides[0]
    for i in range(n):
        tab[i] = C.cast(base + i * stride, dbl_p)
    return tab


# ------------------------------------------------------------------ reference CUDA build (oracle/_ref/librefcuda.so)
REF_CUDA_LIB_PATH = os.path.join(REF_DIR, "librefcuda.so")
REF_CUDA_BIN = os.path.join(REF_DIR, "parallel")


def have_refcuda() -> bool:
    return os.path.exists(REF_CUDA_LIB_PATH)


_refcuda = None


def refcuda():
    """oracle/ref_cuda_harness.cu: the reference's unmodified src/parallel/main.cu behind two entry points.
    Needs a GPU to run (its CHECK_CUDA_ERROR exits the process otherwise: call refcuda_device_count first)."""
    global _refcuda
    if _refcuda is None:
        R = C.CDLL(REF_CUDA_LIB_PATH)
        R.refcuda_device_count.restype = C.c_int
        R.refcuda_sor.restype = C.c_int
        R.refcuda_sor.argtypes = [dbl_p, dbl_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double,
                                  C.c_int, C.c_int]
        R.refcuda_run.restype = C.c_int
        R.refcuda_run.argtypes = [C.c_int, C.c_double, C.c_int, C.c_int] + [C.c_double] * 9 + \
                                 [C.c_int, C.c_int, C.c_int, dbl_p, dbl_p, dbl_p, int_p, dbl_p, C.c_int]
        _refcuda = R
    return _refcuda


def refcuda_sor(p, RHS, dx, dy, omega, eps, max_it, block_size=16):
    """Reference CUDA SOR (main.cu:656-726) on [i][j] grids; p updated in place; returns it+1 / -1."""
    ni, nj = p.shape[0] - 2, p.shape[1] - 2
    return refcuda().refcuda_sor(_ptr(p), _ptr(RHS), ni, nj, dx, dy, omega, eps, max_it, block_size)


def refcuda_run(P: Params, block_size=16, max_steps=0, max_log=100000):
    """Whole run of the reference CUDA solver (main.cu:1046-1103) from zero fields."""
    u, v, p = zeros(P), zeros(P), zeros(P)
    ret = (C.c_int * max_log)()
    dts = (C.c_double * max_log)()
    n = refcuda().refcuda_run(P.problem, P.f, P.i_max, P.j_max, P.a, P.b, P.T, P.Re, P.g_x, P.g_y, P.tau, P.omega,
                              P.epsilon, P.max_it, block_size, max_steps, _ptr(u), _ptr(v), _ptr(p), ret, dts, max_log)
    k = min(n, max_log)
    return {"steps": n, "u": u, "v": v, "p": p, "sor_ret": list(ret[:k]), "dt": list(dts[:k])}
