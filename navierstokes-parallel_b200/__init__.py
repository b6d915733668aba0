"""navierstokes-parallel_b200 -- host-side mirror (ctypes) of the C ABI in include/nsb200.h.

The product is libnsb200.so (hand-written sm_100a CUDA behind `extern "C"`); this module only
binds it.  There is no CPU path: importing works anywhere (so CPU-only tests can check the ABI),
but creating a `Solver` without a CUDA device, or without the built library, raises.

The directory name contains a hyphen, so import it with
    importlib.import_module("navierstokes-parallel_b200")
(tests/, bench.py and __graft_entry__.py do exactly that).

Names follow the reference (src/serial/*.h): init -> read_params, max_mat, set_noslip/set_inflow,
FG, SOR, output.  Grids handed in or out in "ref" layout are numpy arrays of shape
(i_max+2, j_max+2) indexed [i, j] like the reference's `double**`.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_PATH = os.path.join(HERE, "libnsb200.so")
HEADER_PATH = os.path.join(ROOT, "include", "nsb200.h")

TOP, BOTTOM, LEFT, RIGHT = 0, 1, 2, 3
U, V, P, F, G, RHS = 0, 1, 2, 3, 4, 5
LAYOUT_REF, LAYOUT_XFAST = 0, 1
NOT_CONVERGED = -1
DIV_AUTO, DIV_EXACT, DIV_RECIP = 0, 1, 2
SOR_AUTO, SOR_STREAM, SOR_RESIDENT, SOR_SIMPLE, SOR_STREAM_LAUNCH, SOR_SERIAL_ORDER = 0, 1, 2, 3, 4, 5

dbl_p = C.POINTER(C.c_double)
dbl_pp = C.POINTER(dbl_p)
int_p = C.POINTER(C.c_int)


class Params(C.Structure):
    """io.h:14 init(): the 15 values of a parameter file."""
    _fields_ = [("problem", C.c_int), ("f", C.c_double), ("i_max", C.c_int), ("j_max", C.c_int),
                ("a", C.c_double), ("b", C.c_double), ("T", C.c_double), ("Re", C.c_double),
                ("g_x", C.c_double), ("g_y", C.c_double), ("tau", C.c_double),
                ("omega", C.c_double), ("epsilon", C.c_double), ("max_it", C.c_int),
                ("n_print", C.c_int)]

    def asdict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class Options(C.Structure):
    _fields_ = [("device", C.c_int), ("div_mode", C.c_int), ("sor_engine", C.c_int),
                ("sor_chunk_rows", C.c_int), ("sor_check_batch", C.c_int), ("stream", C.c_void_p),
                ("sor_pass_k", C.c_int), ("spin_timeout_ms", C.c_int), ("reserved", C.c_int * 6)]


class StepInfo(C.Structure):
    _fields_ = [("t", C.c_double), ("dt", C.c_double), ("gamma", C.c_double), ("u_max", C.c_double),
                ("v_max", C.c_double), ("norm_p", C.c_double), ("res", C.c_double),
                ("sor_ret", C.c_int), ("sor_iters", C.c_int)]

    def asdict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class NsbError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile libnsb200.so + build/parallel in-tree (nvcc cross-compiles without a GPU)."""
    out = subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "all"], capture_output=True, text=True)
    if out.returncode != 0:
        raise NsbError("building libnsb200.so failed:\n" + out.stdout[-4000:] + out.stderr[-4000:])
    if verbose:
        print(out.stdout[-2000:])
    return LIB_PATH


_lib = None

# name -> (restype, argtypes): every symbol include/nsb200.h declares
ABI = {
    "nsb_abi_version": (C.c_int, []),
    "nsb_last_error": (C.c_char_p, [C.c_void_p]),
    "nsb_default_options": (None, [C.POINTER(Options)]),
    "nsb_read_params": (C.c_int, [C.c_char_p, C.POINTER(Params)]),
    "nsb_create": (C.c_int, [C.POINTER(Params), C.POINTER(Options), C.POINTER(C.c_void_p)]),
    "nsb_destroy": (None, [C.c_void_p]),
    "nsb_get_params": (C.c_int, [C.c_void_p, C.POINTER(Params)]),
    "nsb_stream": (C.c_void_p, [C.c_void_p]),
    "nsb_max_uv": (C.c_int, [C.c_void_p, dbl_p, dbl_p]),
    "nsb_dt_gamma": (C.c_int, [C.c_void_p, C.c_double, C.c_double, dbl_p, dbl_p]),
    "nsb_set_inflow": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_double]),
    "nsb_set_noslip": (C.c_int, [C.c_void_p, C.c_int]),
    "nsb_apply_bc": (C.c_int, [C.c_void_p, C.c_double]),
    "nsb_fg": (C.c_int, [C.c_void_p, C.c_double, C.c_double]),
    "nsb_rhs": (C.c_int, [C.c_void_p, C.c_double]),
    "nsb_fg_rhs": (C.c_int, [C.c_void_p, C.c_double, C.c_double]),
    "nsb_sor": (C.c_int, [C.c_void_p, int_p, dbl_p, dbl_p]),
    "nsb_sor_iterations": (C.c_int, [C.c_void_p, C.c_int, dbl_p]),
    "nsb_update_uv": (C.c_int, [C.c_void_p, C.c_double]),
    "nsb_step": (C.c_int, [C.c_void_p, C.c_double, C.POINTER(StepInfo)]),
    "nsb_run": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(StepInfo), C.c_int, int_p, dbl_p]),
    "nsb_get_center": (C.c_int, [C.c_void_p, dbl_p, dbl_p]),
    "nsb_sor_seconds": (C.c_double, [C.c_void_p]),
    "nsb_launch_count": (C.c_longlong, [C.c_void_p]),
    "nsb_sor_plan_info": (C.c_int, [C.c_void_p, int_p, int_p, int_p, int_p]),
    "nsb_upload": (C.c_int, [C.c_void_p, C.c_int, dbl_p, C.c_int]),
    "nsb_download": (C.c_int, [C.c_void_p, C.c_int, dbl_p, C.c_int]),
    "nsb_output": (C.c_int, [C.c_void_p, C.c_double, C.c_char_p]),
    "nsb_output_async": (C.c_int, [C.c_void_p, C.c_double, C.c_char_p]),
    "nsb_output_wait": (C.c_int, [C.c_void_p]),
    "nsb_device_view": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_longlong), int_p, int_p]),
    "nsb_SOR": (C.c_int, [dbl_pp, C.c_int, C.c_int, C.c_double, C.c_double, dbl_pp, dbl_pp, C.c_double, C.c_double, C.c_int]),
    "nsb_SOR_iters": (C.c_int, [dbl_pp, C.c_int, C.c_int, C.c_double, C.c_double, dbl_pp, dbl_pp, C.c_double, C.c_double, C.c_int, int_p, dbl_p]),
    "nsb_FG": (C.c_int, [dbl_pp, dbl_pp, dbl_pp, dbl_pp, C.c_int, C.c_int] + [C.c_double] * 7),
    "nsb_max_mat": (C.c_double, [C.c_int, C.c_int, dbl_pp]),
    "nsb_timestep": (C.c_int, [C.POINTER(Params), C.c_double] + [dbl_pp] * 6 + [C.POINTER(StepInfo)]),
    "nsb_host_api_reset": (None, []),
    "nsb_max_mat_noalloc": (C.c_int, [C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    # y-slab decomposition
    "nsb_slab_partition": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, int_p, int_p, int_p, int_p]),
    "nsb_create_slab": (C.c_int, [C.POINTER(Params), C.POINTER(Options), C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "nsb_slab_desc_bytes": (C.c_int, []),
    "nsb_slab_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nsb_slab_connect": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nsb_slab_info": (C.c_int, [C.c_void_p, int_p, int_p, int_p, int_p, int_p]),
    "nsb_slab_error": (C.c_int, [C.c_void_p]),
    "nsb_group_create": (C.c_int, [C.POINTER(Params), C.POINTER(Options), C.c_int, int_p, C.POINTER(C.c_void_p)]),
    "nsb_group_destroy": (None, [C.c_void_p]),
    "nsb_group_last_error": (C.c_char_p, [C.c_void_p]),
    "nsb_group_size": (C.c_int, [C.c_void_p]),
    "nsb_group_ctx": (C.c_void_p, [C.c_void_p, C.c_int]),
    "nsb_group_step": (C.c_int, [C.c_void_p, C.c_double, C.POINTER(StepInfo)]),
    "nsb_group_run": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(StepInfo), C.c_int, int_p, dbl_p]),
    "nsb_group_sor": (C.c_int, [C.c_void_p, int_p, dbl_p, dbl_p]),
    "nsb_group_sor_iterations": (C.c_int, [C.c_void_p, C.c_int, dbl_p]),
    "nsb_group_sor_seconds": (C.c_double, [C.c_void_p]),
    "nsb_group_upload": (C.c_int, [C.c_void_p, C.c_int, dbl_p, C.c_int]),
    "nsb_group_download": (C.c_int, [C.c_void_p, C.c_int, dbl_p, C.c_int]),
    "nsb_group_get_center": (C.c_int, [C.c_void_p, dbl_p, dbl_p]),
    "nsb_group_output": (C.c_int, [C.c_void_p, C.c_double, C.c_char_p]),
    "nsb_group_output_async": (C.c_int, [C.c_void_p, C.c_double, C.c_char_p]),
    "nsb_group_output_wait": (C.c_int, [C.c_void_p]),
}


def lib():
    """Load libnsb200.so (never builds implicitly; never falls back to anything else)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NsbError(f"{LIB_PATH} is missing: run __graft_entry__.build() / make -C "
                           f"{os.path.join(HERE, 'csrc')}.  There is no CPU fallback.")
        # NSB200_LIB: developer hook to load a tuning variant of the same library (tools/build_variants.sh)
        L = C.CDLL(os.environ.get("NSB200_LIB") or LIB_PATH)
        for name, (res, args) in ABI.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def make_params(**kw) -> Params:
    d = dict(problem=1, f=1.0, i_max=64, j_max=64, a=1.0, b=1.0, T=0.01, Re=1000.0, g_x=0.0,
             g_y=0.0, tau=1.0, omega=1.7, epsilon=1e-4, max_it=20000, n_print=1)
    d.update(kw)
    return Params(**d)


def read_params(path: str) -> Params:
    """io.h:14 init()."""
    p = Params()
    if lib().nsb_read_params(os.fsencode(path), C.byref(p)) != 0:
        raise FileNotFoundError(path)
    return p


def rows(a: np.ndarray):
    """`double**` row-pointer table over a C-contiguous 2-D array: the reference's grid type."""
    assert a.dtype == np.float64 and a.ndim == 2 and a.flags["C_CONTIGUOUS"]
    tab = (dbl_p * a.shape[0])()
    base, stride = a.ctypes.data, a.strides[0]
    for i in range(a.shape[0]):
        tab[i] = C.cast(base + i * stride, dbl_p)
    return tab


class Solver:
    """One simulation resident in HBM (an `nsb_ctx`).  Method names follow the reference."""

    def __init__(self, params: Params, device: int = 0, div_mode: int = DIV_AUTO,
                 sor_engine: int = SOR_AUTO, chunk_rows: int = 0, check_batch: int = 0, stream: int = 0,
                 pass_k: int = 0, spin_timeout_ms: int = 0):
        self._L = lib()
        self.params = params
        o = Options()
        self._L.nsb_default_options(C.byref(o))
        o.device, o.div_mode, o.sor_engine = device, div_mode, sor_engine
        o.sor_chunk_rows, o.sor_check_batch = chunk_rows, check_batch
        o.sor_pass_k, o.spin_timeout_ms = pass_k, spin_timeout_ms
        o.stream = stream or None
        h = C.c_void_p()
        rc = self._L.nsb_create(C.byref(params), C.byref(o), C.byref(h))
        if rc != 0:
            raise NsbError(self._L.nsb_last_error(None).decode())
        self._h = h
        self.shape = (params.i_max + 2, params.j_max + 2)

    # -- plumbing
    def close(self):
        if getattr(self, "_h", None):
            self._L.nsb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc, allow=()):
        if rc != 0 and rc not in allow:
            raise NsbError(f"nsb error {rc}: {self._L.nsb_last_error(self._h).decode()}")
        return rc

    @property
    def stream(self) -> int:
        return self._L.nsb_stream(self._h) or 0

    # -- reference-named phases
    def max_mat(self):
        """io.h:42 for u and v: returns (u_max, v_max)."""
        a, b = C.c_double(), C.c_double()
        self._ck(self._L.nsb_max_uv(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def dt_gamma(self, u_max, v_max):
        a, b = C.c_double(), C.c_double()
        self._ck(self._L.nsb_dt_gamma(self._h, u_max, v_max, C.byref(a), C.byref(b)))
        return a.value, b.value

    def set_inflow(self, side, u_fix, v_fix):
        self._ck(self._L.nsb_set_inflow(self._h, side, u_fix, v_fix))

    def set_noslip(self, side):
        self._ck(self._L.nsb_set_noslip(self._h, side))

    def apply_bc(self, t):
        self._ck(self._L.nsb_apply_bc(self._h, t))

    def FG(self, dt, gamma):
        self._ck(self._L.nsb_fg(self._h, dt, gamma))

    def RHS(self, dt):
        self._ck(self._L.nsb_rhs(self._h, dt))

    def FG_RHS(self, dt, gamma):
        self._ck(self._L.nsb_fg_rhs(self._h, dt, gamma))

    def SOR(self):
        """Returns (ret, iters, res, norm_p); ret = it+1 when converged, -1 otherwise."""
        it, r, n0 = C.c_int(), C.c_double(), C.c_double()
        rc = self._L.nsb_sor(self._h, C.byref(it), C.byref(r), C.byref(n0))
        if rc < 0 and rc != NOT_CONVERGED:
            self._ck(rc)
        return rc, it.value, r.value, n0.value

    def sor_iterations(self, n):
        r = C.c_double()
        self._ck(self._L.nsb_sor_iterations(self._h, n, C.byref(r)))
        return r.value

    def update_uv(self, dt):
        self._ck(self._L.nsb_update_uv(self._h, dt))

    def step(self, t) -> dict:
        info = StepInfo()
        self._ck(self._L.nsb_step(self._h, t, C.byref(info)))
        return info.asdict()

    def run(self, max_steps=0, max_log=100000):
        log = (StepInfo * max_log)()
        n, t = C.c_int(), C.c_double()
        self._ck(self._L.nsb_run(self._h, max_steps, log, max_log, C.byref(n), C.byref(t)))
        return [log[i].asdict() for i in range(min(n.value, max_log))], t.value

    def center(self):
        a, b = C.c_double(), C.c_double()
        self._ck(self._L.nsb_get_center(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def sor_seconds(self):
        return self._L.nsb_sor_seconds(self._h)

    def launch_count(self):
        return self._L.nsb_launch_count(self._h)

    def sor_plan(self) -> dict:
        """Iterations per pass, rows per chunk, strips and chunks of the persistent SOR sweep as planned."""
        v = [C.c_int() for _ in range(4)]
        self._ck(self._L.nsb_sor_plan_info(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("pass_k", "chunk_rows", "n_strips", "n_chunks"), (x.value for x in v)))

    def output(self, t, prefix, wait=True):
        """io.h:31 output(); wait=False: returns once the fields are snapshotted, output_wait() joins the writer."""
        f = self._L.nsb_output if wait else self._L.nsb_output_async
        self._ck(f(self._h, t, os.fsencode(prefix)))

    def output_wait(self):
        self._ck(self._L.nsb_output_wait(self._h))

    # -- transfers
    def upload(self, field, a: np.ndarray, layout=LAYOUT_REF):
        a = np.ascontiguousarray(a, dtype=np.float64)
        want = self.shape if layout == LAYOUT_REF else self.shape[::-1]
        assert a.shape == want, (a.shape, want)
        self._ck(self._L.nsb_upload(self._h, field, a.ctypes.data_as(dbl_p), layout))

    def download(self, field, layout=LAYOUT_REF) -> np.ndarray:
        out = np.empty(self.shape if layout == LAYOUT_REF else self.shape[::-1], dtype=np.float64)
        self._ck(self._L.nsb_download(self._h, field, out.ctypes.data_as(dbl_p), layout))
        return out

    def device_view(self, field):
        p, pitch, c0, r0 = C.c_void_p(), C.c_longlong(), C.c_int(), C.c_int()
        self._ck(self._L.nsb_device_view(self._h, field, C.byref(p), C.byref(pitch), C.byref(c0), C.byref(r0)))
        return p.value, pitch.value, c0.value, r0.value


def slab_partition(j_max: int, chunk_rows: int, world: int, rank: int):
    """Rows of slab `rank`: (j_lo, nj, band0, nbands_global).  Pure host arithmetic (no GPU needed)."""
    a, b, c, d = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    rc = lib().nsb_slab_partition(j_max, chunk_rows, world, rank, C.byref(a), C.byref(b), C.byref(c), C.byref(d))
    if rc != 0:
        raise NsbError(f"cannot cut j_max={j_max} into {world} slabs of {chunk_rows}-row bands")
    return a.value, b.value, c.value, d.value


def _options(device, div_mode, sor_engine, chunk_rows, check_batch, stream=0, pass_k=0, spin_timeout_ms=0):
    o = Options()
    lib().nsb_default_options(C.byref(o))
    o.device, o.div_mode, o.sor_engine = device, div_mode, sor_engine
    o.sor_chunk_rows, o.sor_check_batch = chunk_rows, check_batch
    o.sor_pass_k, o.spin_timeout_ms = pass_k, spin_timeout_ms
    o.stream = stream or None
    return o


class Group:
    """y-slab decomposition driven by ONE process (`nsb_group`): slab k on devices[k].  Equal devices
    are allowed -- the slabs then run in lock step on one stream, which is how a 1-GPU box tests it."""

    def __init__(self, params: Params, devices, div_mode: int = DIV_AUTO, chunk_rows: int = 0, check_batch: int = 0,
                 pass_k: int = 0, spin_timeout_ms: int = 0):
        self._L = lib()
        self.params = params
        o = _options(0, div_mode, SOR_STREAM, chunk_rows, check_batch, pass_k=pass_k, spin_timeout_ms=spin_timeout_ms)
        dv = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        rc = self._L.nsb_group_create(C.byref(params), C.byref(o), len(devices), dv, C.byref(h))
        if rc != 0:
            raise NsbError(self._L.nsb_group_last_error(None).decode())
        self._h = h
        self.shape = (params.i_max + 2, params.j_max + 2)

    def close(self):
        if getattr(self, "_h", None):
            self._L.nsb_group_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc, allow=()):
        if rc != 0 and rc not in allow:
            raise NsbError(f"nsb error {rc}: {self._L.nsb_group_last_error(self._h).decode()}")
        return rc

    def slabs(self):
        out = []
        for k in range(self._L.nsb_group_size(self._h)):
            v = [C.c_int() for _ in range(5)]
            self._L.nsb_slab_info(self._L.nsb_group_ctx(self._h, k), *[C.byref(x) for x in v])
            out.append(dict(zip(("rank", "world", "j_lo", "nj", "chunk_rows"), (x.value for x in v))))
        return out

    def step(self, t) -> dict:
        info = StepInfo()
        self._ck(self._L.nsb_group_step(self._h, t, C.byref(info)))
        return info.asdict()

    def run(self, max_steps=0, max_log=100000):
        log = (StepInfo * max_log)()
        n, t = C.c_int(), C.c_double()
        self._ck(self._L.nsb_group_run(self._h, max_steps, log, max_log, C.byref(n), C.byref(t)))
        return [log[k].asdict() for k in range(min(n.value, max_log))], t.value

    def SOR(self):
        it, res, nrm = C.c_int(), C.c_double(), C.c_double()
        rc = self._L.nsb_group_sor(self._h, C.byref(it), C.byref(res), C.byref(nrm))
        if rc < 0 and rc != NOT_CONVERGED:
            self._ck(rc)
        return rc, it.value, res.value, nrm.value

    def sor_iterations(self, n):
        res = C.c_double()
        self._ck(self._L.nsb_group_sor_iterations(self._h, n, C.byref(res)))
        return res.value

    def sor_seconds(self):
        return self._L.nsb_group_sor_seconds(self._h)

    def center(self):
        a, b = C.c_double(), C.c_double()
        self._ck(self._L.nsb_group_get_center(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def output(self, t, prefix, wait=True):
        f = self._L.nsb_group_output if wait else self._L.nsb_group_output_async
        self._ck(f(self._h, t, os.fsencode(prefix)))

    def output_wait(self):
        self._ck(self._L.nsb_group_output_wait(self._h))

    def upload(self, field, a: np.ndarray, layout=LAYOUT_REF):
        a = np.ascontiguousarray(a, dtype=np.float64)
        assert a.size == self.shape[0] * self.shape[1]
        self._ck(self._L.nsb_group_upload(self._h, field, a.ctypes.data_as(dbl_p), layout))

    def download(self, field, layout=LAYOUT_REF) -> np.ndarray:
        shape = self.shape if layout == LAYOUT_REF else self.shape[::-1]
        out = np.empty(shape, dtype=np.float64)
        self._ck(self._L.nsb_group_download(self._h, field, out.ctypes.data_as(dbl_p), layout))
        return out


class SlabSolver(Solver):
    """One slab per PROCESS (one process per GPU, e.g. under torchrun): creates the slab of `rank`,
    exchanges the 256-byte descriptors through `all_gather` (a callable bytes -> list of bytes in rank
    order, e.g. built on torch.distributed) and wires the peers over CUDA IPC.  After that every rank
    calls step() / run() / SOR() as if it were alone; halos and reductions travel inside the kernels."""

    def __init__(self, params: Params, rank: int, world: int, all_gather, device: int = 0, div_mode: int = DIV_AUTO,
                 chunk_rows: int = 0, check_batch: int = 0, pass_k: int = 0, spin_timeout_ms: int = 0):
        self._L = lib()
        self.params = params
        o = _options(device, div_mode, SOR_STREAM, chunk_rows, check_batch, pass_k=pass_k, spin_timeout_ms=spin_timeout_ms)
        h = C.c_void_p()
        rc = self._L.nsb_create_slab(C.byref(params), C.byref(o), rank, world, C.byref(h))
        if rc != 0:
            raise NsbError(self._L.nsb_last_error(None).decode())
        self._h = h
        n = self._L.nsb_slab_desc_bytes()
        mine = C.create_string_buffer(n)
        self._ck(self._L.nsb_slab_export(self._h, mine))
        blobs = all_gather(mine.raw)
        assert len(blobs) == world and all(len(b) == n for b in blobs)
        self._ck(self._L.nsb_slab_connect(self._h, C.create_string_buffer(b"".join(blobs), n * world)))
        v = [C.c_int() for _ in range(5)]
        self._L.nsb_slab_info(self._h, *[C.byref(x) for x in v])
        self.rank, self.world, self.j_lo, self.nj, self.chunk_rows = (x.value for x in v)
        self.shape = (params.i_max + 2, self.nj + 2)          # local rows 0 .. nj+1


def torch_all_gather(group=None):
    """all_gather callable for SlabSolver on top of torch.distributed (any backend)."""
    import torch
    import torch.distributed as dist

    def gather(blob: bytes):
        world = dist.get_world_size(group)
        dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
        t = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev)
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t, group=group)
        return [bytes(x.cpu().numpy().tobytes()) for x in out]
    return gather
