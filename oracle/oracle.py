"""ctypes front-end of the CPU oracle — TEST INFRASTRUCTURE, not product code.

Loads either implementation of oracle/oracle_api.h:
  kind="port"       oracle/libgs_oracle.so            (C restatement, gs_oracle.c)
  kind="reference"  oracle/_ref/libgpusolve_ref.so    (the reference's own TUs)
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs import this module.  Fields come back as numpy arrays of shape (Z, Y, X), i.e.
indexed [z, y, x] by logical coordinates.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

V, F, R, E, RESTV, NEWTONV, NEWTONF = range(7)
LINEAR, NONLINEAR, NEWTON = 0, 1, 2

CANONICAL_VALUES = (6.0, -1.0, -1.0, -1.0, -1.0, -1.0, -1.0)
CANONICAL_OFFSETS = ((0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1))


class Params(C.Structure):
    """mirror of gs_params, include/gpusolve_cuda.h"""
    _fields_ = [
        ("maxiter", C.c_uint64), ("tol", C.c_double), ("omega", C.c_double), ("gamma", C.c_double),
        ("h", C.c_double), ("dim", C.c_uint64 * 3), ("pre", C.c_uint64), ("post", C.c_uint64),
        ("stencil_val", C.c_double * 7), ("stencil_off", (C.c_int32 * 3) * 7),
        ("mode", C.c_int32), ("print_progress", C.c_int32),
    ]


def make_params(n, mode=0, pre=2, post=2, maxiter=10, tol=1e-5, omega=0.8, gamma=1.0,
                values=CANONICAL_VALUES, offsets=CANONICAL_OFFSETS, print_progress=0):
    if isinstance(n, int):
        n = (n, n, n)
    p = Params()
    p.maxiter, p.tol, p.omega, p.gamma = maxiter, tol, omega, gamma
    p.dim[:] = n
    p.h = 1.0 / (n[1] + 1)
    p.pre, p.post = pre, post
    for i in range(7):
        p.stencil_val[i] = values[i]
        for d in range(3):
            p.stencil_off[i][d] = offsets[i][d]
    p.mode, p.print_progress = mode, print_progress
    return p


def lib_path(kind):
    return os.path.join(HERE, "libgs_oracle.so") if kind == "port" else os.path.join(HERE, "_ref", "libgpusolve_ref.so")


def available(kind):
    return os.path.exists(lib_path(kind))


def build(ref=True):
    """compile the checker(s); building the checker is not using it"""
    subprocess.run(["make", "-C", HERE, "port"], check=True, capture_output=True)
    if ref and os.path.isdir(os.path.join(os.environ.get("GPUSOLVE_REF", "/root/reference"), "src")):
        subprocess.run(["make", "-C", HERE, "ref"], check=True, capture_output=True)


_libs = {}


def _load(kind):
    if kind in _libs:
        return _libs[kind]
    lib = C.CDLL(lib_path(kind))
    vp, i, d, u64 = C.c_void_p, C.c_int, C.c_double, C.c_uint64
    dp = C.POINTER(C.c_double)
    sig = {
        "gso_kind": (C.c_char_p, []),
        "gso_create": (i, [C.POINTER(Params), C.POINTER(vp)]),
        "gso_destroy": (None, [vp]),
        "gso_num_levels": (i, [vp]),
        "gso_level_info": (i, [vp, i, C.POINTER(u64 * 3), dp]),
        "gso_get_field": (i, [vp, i, i, vp]),
        "gso_set_field": (i, [vp, i, i, vp]),
        "gso_residual": (d, [vp, i]),
        "gso_smooth": (None, [vp, i, u64]),
        "gso_restrict": (None, [vp, i, i, i]),
        "gso_apply_stencil": (None, [vp, i]),
        "gso_interpolate": (None, [vp, i]),
        "gso_axpy": (None, [vp, i, i, i, i]),
        "gso_vcycle": (d, [vp]),
        "gso_solve": (i, [vp, dp, i]),
        "gso_newton_begin": (None, [vp]),
        "gso_newton_compF": (d, [vp]),
        "gso_newton_find_error": (i, [vp]),
        "gso_newton_solve": (i, [vp, dp, i, C.POINTER(i)]),
        "gso_norm_exact": (d, [vp, i, i]),
        "gso_set_threads": (None, [i]),
    }
    if kind == "reference":
        sig["gso_ref_native_solve"] = (None, [vp])
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _libs[kind] = lib
    return lib


class Oracle:
    def __init__(self, params, kind="port"):
        self.lib = _load(kind)
        self.kind = kind
        self.params = params
        self.g = C.c_void_p()
        if self.lib.gso_create(C.byref(params), C.byref(self.g)) != 0:
            raise RuntimeError("gso_create failed")

    def close(self):
        if self.g:
            self.lib.gso_destroy(self.g)
            self.g = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def num_levels(self):
        return self.lib.gso_num_levels(self.g)

    def level_info(self, level):
        dim = (C.c_uint64 * 3)()
        h = C.c_double()
        if self.lib.gso_level_info(self.g, level, C.byref(dim), C.byref(h)) != 0:
            raise IndexError(level)
        return tuple(dim), h.value

    def shape(self, level):
        (nx, ny, nz), _ = self.level_info(level)
        return (nz + 2, ny + 2, nx + 2)

    def get(self, level, field):
        out = np.empty(self.shape(level), dtype=np.float64)
        if self.lib.gso_get_field(self.g, level, field, out.ctypes.data) != 0:
            raise KeyError((level, field))
        return out

    def set(self, level, field, arr):
        a = np.ascontiguousarray(arr, dtype=np.float64)
        assert a.shape == self.shape(level)
        if self.lib.gso_set_field(self.g, level, field, a.ctypes.data) != 0:
            raise KeyError((level, field))

    def residual(self, level):
        return self.lib.gso_residual(self.g, level)

    def smooth(self, level, iters):
        self.lib.gso_smooth(self.g, level, iters)

    def restrict(self, level, src, dst):
        self.lib.gso_restrict(self.g, level, src, dst)

    def apply_stencil(self, level):
        self.lib.gso_apply_stencil(self.g, level)

    def interpolate(self, level):
        self.lib.gso_interpolate(self.g, level)

    def axpy(self, level, dst, src, sign):
        self.lib.gso_axpy(self.g, level, dst, src, sign)

    def vcycle(self):
        return self.lib.gso_vcycle(self.g)

    def solve(self, cap=64):
        norms = (C.c_double * cap)()
        n = self.lib.gso_solve(self.g, norms, cap)
        return list(norms[: n + 1])

    def newton_begin(self):
        self.lib.gso_newton_begin(self.g)

    def newton_compF(self):
        return self.lib.gso_newton_compF(self.g)

    def newton_find_error(self):
        return self.lib.gso_newton_find_error(self.g)

    def newton_solve(self, cap=64):
        norms = (C.c_double * cap)()
        inner = C.c_int()
        n = self.lib.gso_newton_solve(self.g, norms, cap, C.byref(inner))
        return list(norms[: n + 1]), inner.value

    def native_solve(self):
        """the reference's own solve() entry point (kind='reference' only); prints its stdout"""
        self.lib.gso_ref_native_solve(self.g)

    def norm_exact(self, level, field):
        return self.lib.gso_norm_exact(self.g, level, field)

    def set_threads(self, n):
        self.lib.gso_set_threads(n)
