"""gpusolve_b200 — ctypes binding of the sm_100a CUDA backend (lib/libgpusolve_cuda.so).

The product is the C-ABI shared library declared in include/gpusolve_cuda.h plus the
`bin/GpuSolve-cuda <config>` driver; this module is the thin Python view of that ABI
used by tests/ and bench.py.  There is no CPU or PyTorch fallback: importing works
without the library (so tooling can call build()), but constructing a solver raises
if lib/libgpusolve_cuda.so is missing.

The directory is named `gpu-solve_b200` (not importable by that name); load it with
`__graft_entry__.load_package()` which registers it as `gpusolve_b200`.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "lib", "libgpusolve_cuda.so")
EXE = os.path.join(HERE, "bin", "GpuSolve-cuda")

LINEAR, NONLINEAR, NEWTON = 0, 1, 2
V, F, R, E, RESTV, NEWTONV, NEWTONF = range(7)
X_FASTEST, Z_FASTEST = 0, 1
OPS = ("smooth", "residual", "restrict", "prolong", "other")

CANONICAL_VALUES = (6.0, -1.0, -1.0, -1.0, -1.0, -1.0, -1.0)
CANONICAL_OFFSETS = ((0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1))


class Params(C.Structure):
    """gs_params, include/gpusolve_cuda.h (POD mirror of GridParams, src/gridParams.h:28-47)"""
    _fields_ = [
        ("maxiter", C.c_uint64), ("tol", C.c_double), ("omega", C.c_double), ("gamma", C.c_double),
        ("h", C.c_double), ("dim", C.c_uint64 * 3), ("pre", C.c_uint64), ("post", C.c_uint64),
        ("stencil_val", C.c_double * 7), ("stencil_off", (C.c_int32 * 3) * 7),
        ("mode", C.c_int32), ("print_progress", C.c_int32),
    ]


class SolveInfo(C.Structure):
    _fields_ = [
        ("initial_residual", C.c_double), ("final_residual", C.c_double), ("iterations", C.c_uint64),
        ("inner_cycles", C.c_uint64), ("converged", C.c_int32), ("reserved", C.c_int32),
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


def build(verbose=False):
    """compile the CUDA backend in-tree for sm_100a (nvcc cross-compiles without a GPU)"""
    jobs = str(max(1, min(8, os.cpu_count() or 1)))  # the five translation units compile independently
    r = subprocess.run(["make", "-C", HERE, "-j", jobs], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libgpusolve_cuda.so failed")
    return LIB


class GpuSolveError(RuntimeError):
    pass


_lib = None

# name -> (restype, argtypes); every symbol include/gpusolve_cuda.h declares
_vp, _i, _u64, _dp = C.c_void_p, C.c_int, C.c_uint64, C.POINTER(C.c_double)
SIGNATURES = {
    "gs_last_error": (C.c_char_p, []),
    "gs_version": (_i, []),
    "gs_device_count": (_i, []),
    "gs_parse_config": (_i, [C.c_char_p, C.POINTER(Params)]),
    "gs_create": (_i, [C.POINTER(Params), _i, C.POINTER(_vp)]),
    "gs_comm_unique_id": (_i, [_vp]),
    "gs_create_slab": (_i, [C.POINTER(Params), _i, _i, _i, _vp, C.POINTER(_vp)]),
    "gs_destroy": (_i, [_vp]),
    "gs_partition": (_i, [C.POINTER(Params), _i, _i, _i, C.POINTER(_u64), C.POINTER(_u64), C.POINTER(_i)]),
    "gs_plan_zchunk": (_i, [_i, _i, _i, _i, _i, _i, _i, _i]),
    "gs_time_halo_exchange": (_i, [C.c_void_p, _i, _i, _i, C.POINTER(C.c_double)]),
    "gs_num_levels": (_i, [_vp]),
    "gs_get_level_info": (_i, [_vp, _i, C.POINTER(_u64 * 3), _dp]),
    "gs_get_field": (_i, [_vp, _i, _i, _vp, _i]),
    "gs_set_field": (_i, [_vp, _i, _i, _vp, _i]),
    "gs_dump_field": (_i, [_vp, _i, _i, C.c_char_p]),
    "gs_smooth": (_i, [_vp, _i, _u64]),
    "gs_residual": (_i, [_vp, _i, _dp]),
    "gs_residual_static": (_i, [_vp, _i, _dp]),
    "gs_restrict": (_i, [_vp, _i, _i, _i]),
    "gs_restrict_residual": (_i, [_vp, _i]),
    "gs_fas_coarse_rhs": (_i, [_vp, _i]),
    "gs_prolong": (_i, [_vp, _i, _vp, _i]),
    "gs_prolong_correct": (_i, [_vp, _i]),
    "gs_vcycle": (_i, [_vp, _dp]),
    "gs_newton_compF": (_i, [_vp, _dp]),
    "gs_newton_step": (_i, [_vp, _dp, C.POINTER(_u64)]),
    "gs_solve": (_i, [_vp, C.POINTER(SolveInfo)]),
    "gs_get_history": (_i, [_vp, _dp, _i, C.POINTER(_i)]),
    "gs_reset_timing": (_i, [_vp]),
    "gs_get_timing": (_i, [_vp, C.POINTER(C.c_double * 5), C.POINTER(_u64 * 5)]),
    "gs_get_level_timing": (_i, [_vp, _i, C.POINTER(C.c_double * 5), C.POINTER(_u64 * 5)]),
    "gs_get_level_sweeps": (_i, [_vp, _i, C.POINTER(_u64)]),
    "gs_launch_count": (_u64, [_vp]),
    "gs_set_option": (_i, [_vp, C.c_char_p, C.c_int64]),
    "gs_stream": (_vp, [_vp]),
    "gs_synchronize": (_i, [_vp]),
}


def load():
    """dlopen the C-ABI library and bind every entry point; raises if it is not built"""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB):
        raise GpuSolveError(f"{LIB} is missing: run __graft_entry__.build() (there is no CPU fallback)")
    lib = C.CDLL(LIB)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library drift apart
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def _check(rc):
    if rc != 0:
        raise GpuSolveError(f"[{rc}] {load().gs_last_error().decode()}")


def parse_config(path):
    p = Params()
    _check(load().gs_parse_config(str(path).encode(), C.byref(p)))
    return p


def partition(params, world, rank, level):
    zb, ze, rep = C.c_uint64(), C.c_uint64(), C.c_int()
    _check(load().gs_partition(C.byref(params), world, rank, level, C.byref(zb), C.byref(ze), C.byref(rep)))
    return zb.value, ze.value, rep.value


def plan_zchunk(tiles, planes, sm_count=148, ctas_per_sm=2, overhead=6, min_chunk=16, max_chunk=128, even=False):
    rc = load().gs_plan_zchunk(tiles, planes, sm_count, ctas_per_sm, overhead, min_chunk, max_chunk, int(even))
    if rc < 0:
        _check(rc)
    return rc


def comm_unique_id():
    buf = C.create_string_buffer(128)
    _check(load().gs_comm_unique_id(buf))
    return buf.raw


class GpuSolve:
    """One grid hierarchy on one GPU (or one z-slab of it when world > 1)."""

    def __init__(self, params, device=0, rank=0, world=1, nccl_id=None):
        self.lib = load()
        self.params = params
        self.h = C.c_void_p()
        if world == 1:
            _check(self.lib.gs_create(C.byref(params), device, C.byref(self.h)))
        else:
            self._id = C.create_string_buffer(nccl_id, 128)
            _check(self.lib.gs_create_slab(C.byref(params), device, rank, world, self._id, C.byref(self.h)))

    def close(self):
        if self.h:
            self.lib.gs_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def num_levels(self):
        return self.lib.gs_num_levels(self.h)

    def level_info(self, level):
        dim, h = (C.c_uint64 * 3)(), C.c_double()
        _check(self.lib.gs_get_level_info(self.h, level, C.byref(dim), C.byref(h)))
        return tuple(dim), h.value

    def shape(self, level):
        (nx, ny, nz), _ = self.level_info(level)
        return (nz + 2, ny + 2, nx + 2)

    def get(self, level, field):
        """dense array indexed [z, y, x] by logical coordinates"""
        out = np.empty(self.shape(level), dtype=np.float64)
        _check(self.lib.gs_get_field(self.h, level, field, out.ctypes.data, X_FASTEST))
        return out

    def get_zfastest(self, level, field):
        """the reference's Vector3 order: indexed [x, y, z]"""
        z, y, x = self.shape(level)
        out = np.empty((x, y, z), dtype=np.float64)
        _check(self.lib.gs_get_field(self.h, level, field, out.ctypes.data, Z_FASTEST))
        return out

    def set(self, level, field, arr):
        a = np.ascontiguousarray(arr, dtype=np.float64)
        assert a.shape == self.shape(level), (a.shape, self.shape(level))
        _check(self.lib.gs_set_field(self.h, level, field, a.ctypes.data, X_FASTEST))

    def dump(self, level, field, path):
        _check(self.lib.gs_dump_field(self.h, level, field, str(path).encode()))

    def smooth(self, level, iters):
        _check(self.lib.gs_smooth(self.h, level, iters))

    def residual(self, level):
        n = C.c_double()
        _check(self.lib.gs_residual(self.h, level, C.byref(n)))
        return n.value

    def residual_static(self, level):
        """gs_residual with the reduction scratch in static device memory: nothing on the call path allocates"""
        n = C.c_double()
        _check(self.lib.gs_residual_static(self.h, level, C.byref(n)))
        return n.value

    def restrict(self, level, src, dst):
        _check(self.lib.gs_restrict(self.h, level, src, dst))

    def restrict_residual(self, level):
        _check(self.lib.gs_restrict_residual(self.h, level))

    def fas_coarse_rhs(self, level):
        _check(self.lib.gs_fas_coarse_rhs(self.h, level))

    def prolong(self, level):
        out = np.empty(self.shape(level), dtype=np.float64)
        _check(self.lib.gs_prolong(self.h, level, out.ctypes.data, X_FASTEST))
        return out

    def prolong_correct(self, level):
        _check(self.lib.gs_prolong_correct(self.h, level))

    def vcycle(self):
        n = C.c_double()
        _check(self.lib.gs_vcycle(self.h, C.byref(n)))
        return n.value

    def newton_compF(self):
        n = C.c_double()
        _check(self.lib.gs_newton_compF(self.h, C.byref(n)))
        return n.value

    def newton_step(self):
        n, inner = C.c_double(), C.c_uint64()
        _check(self.lib.gs_newton_step(self.h, C.byref(n), C.byref(inner)))
        return n.value, inner.value

    def solve(self):
        info = SolveInfo()
        _check(self.lib.gs_solve(self.h, C.byref(info)))
        return info

    def history(self):
        cnt = C.c_int()
        _check(self.lib.gs_get_history(self.h, None, 0, C.byref(cnt)))
        buf = (C.c_double * max(cnt.value, 1))()
        _check(self.lib.gs_get_history(self.h, buf, cnt.value, C.byref(cnt)))
        return list(buf[: cnt.value])

    def reset_timing(self):
        _check(self.lib.gs_reset_timing(self.h))

    def timing(self):
        ms, n = (C.c_double * 5)(), (C.c_uint64 * 5)()
        _check(self.lib.gs_get_timing(self.h, C.byref(ms), C.byref(n)))
        return {OPS[i]: (ms[i], n[i]) for i in range(5)}

    def level_timing(self, level):
        ms, n = (C.c_double * 5)(), (C.c_uint64 * 5)()
        _check(self.lib.gs_get_level_timing(self.h, level, C.byref(ms), C.byref(n)))
        return {OPS[i]: (ms[i], n[i]) for i in range(5)}

    def level_sweeps(self, level):
        n = C.c_uint64()
        _check(self.lib.gs_get_level_sweeps(self.h, level, C.byref(n)))
        return n.value

    @property
    def launches(self):
        return self.lib.gs_launch_count(self.h)

    def set_option(self, name, value):
        _check(self.lib.gs_set_option(self.h, name.encode(), int(value)))

    @property
    def stream(self):
        return self.lib.gs_stream(self.h)

    def time_halo_exchange(self, level, depth=2, reps=20):
        """slabs: ms per ghost-plane refresh of the level's v (collective)"""
        ms = C.c_double()
        _check(self.lib.gs_time_halo_exchange(self.h, level, depth, reps, C.byref(ms)))
        return ms.value

    def synchronize(self):
        _check(self.lib.gs_synchronize(self.h))
