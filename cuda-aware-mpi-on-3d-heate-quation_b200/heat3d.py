"""ctypes mirror of include/heat3d.h — the host-side interface tests and bench.py drive.

This is plumbing over the C ABI (the product is libheat3d.so + the C++ `heat3D` driver).  It never
falls back to a CPU implementation: if the shared library is missing or no CUDA device is usable the
calls raise.  Function names and argument meaning follow include/heat3d.h one to one; the reference
lines each entry point replaces are cited there.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libheat3d.so")
DRIVER_PATH = os.path.join(_HERE, "bin", "heat3D")
CSRC = os.path.join(_HERE, "csrc")

ARITH_STRICT, ARITH_FMA = 0, 1
KERNEL_AUTO, KERNEL_BASELINE, KERNEL_TMA = 0, 1, 2
FLAG_NO_OVERLAP, FLAG_TIME_HALO = 0x1, 0x2


class Heat3DError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"heat3d error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("nx", C.c_uint32), ("ny", C.c_uint32), ("nz", C.c_uint32),
                ("Dx", C.c_double), ("Dy", C.c_double), ("Dz", C.c_double),
                ("arith", C.c_int32), ("temporal_k", C.c_int32), ("kernel", C.c_int32),
                ("nslabs", C.c_int32), ("slab_first", C.c_int32), ("nlocal", C.c_int32),
                ("devices", C.POINTER(C.c_int32)), ("nccl_unique_id", C.c_void_p), ("stream", C.c_void_p),
                ("flags", C.c_uint32), ("tune", C.c_int32 * 6), ("nslabs_y", C.c_int32)]


class Timing(C.Structure):
    _fields_ = [("steps", C.c_uint64), ("launches", C.c_uint64), ("last_call_ms", C.c_double),
                ("total_ms", C.c_double), ("exposed_halo_ms", C.c_double),
                ("last_call_launches", C.c_uint64), ("last_call_stencil_launches", C.c_uint64)]


# every symbol include/heat3d.h declares: name -> (restype, argtypes)
_dp = C.POINTER(C.c_double)
_ctx = C.c_void_p
EXPORTS = {
    "heat3d_default_params": (None, [C.POINTER(Params), C.c_uint32, C.c_uint32, C.c_uint32]),
    "heat3d_create": (C.c_int, [C.POINTER(Params), C.POINTER(_ctx)]),
    "heat3d_destroy": (C.c_int, [_ctx]),
    "heat3d_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "heat3d_init_reference_field": (C.c_int, [_ctx, C.c_double]),
    "heat3d_upload": (C.c_int, [_ctx, _dp]),
    "heat3d_download": (C.c_int, [_ctx, _dp]),
    "heat3d_upload_planes": (C.c_int, [_ctx, _dp, C.c_uint32, C.c_uint32]),
    "heat3d_download_planes": (C.c_int, [_ctx, _dp, C.c_uint32, C.c_uint32]),
    "heat3d_upload_planes_async": (C.c_int, [_ctx, _dp, C.c_uint32, C.c_uint32]),
    "heat3d_upload_commit": (C.c_int, [_ctx]),
    "heat3d_local_range": (C.c_int, [_ctx, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "heat3d_step": (C.c_int, [_ctx, C.c_uint32]),
    "heat3d_step_residual": (C.c_int, [_ctx, C.c_uint32, _dp]),
    "heat3d_run": (C.c_int, [_ctx, C.c_uint32, C.c_double, C.c_uint32, C.POINTER(C.c_uint32),
                             C.POINTER(C.c_int32), C.POINTER(C.c_uint32)]),
    "heat3d_residual": (C.c_int, [_ctx, _dp]),
    "heat3d_mean_abs_error_vs_y": (C.c_int, [_ctx, C.c_double, _dp]),
    "heat3d_mean_abs_error_vs_y_noalloc": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_double,
                                                     C.c_void_p, C.c_void_p]),
    "heat3d_snapshot_async": (C.c_int, [_ctx, _dp]),
    "heat3d_snapshot_wait": (C.c_int, [_ctx]),
    "heat3d_snapshot_strided_async": (C.c_int, [_ctx, _dp, C.c_uint32, C.c_uint32, C.c_uint32]),
    "heat3d_device_count": (C.c_int, [C.POINTER(C.c_int32)]),
    "heat3d_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "heat3d_host_free": (C.c_int, [C.c_void_p]),
    "heat3d_sync": (C.c_int, [_ctx]),
    "heat3d_get_timing": (C.c_int, [_ctx, C.POINTER(Timing)]),
    "heat3d_reset_timing": (C.c_int, [_ctx]),
    "heat3d_compute_stream": (C.c_void_p, [_ctx, C.c_int32]),
    "heat3d_layout": (C.c_int, [_ctx, C.c_int32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "heat3d_last_error": (C.c_char_p, []),
    "heat3d_version": (C.c_int, []),
    "heat3d_kernel_name": (C.c_char_p, [_ctx]),
    "heat3d_exchange_name": (C.c_char_p, [_ctx]),
}

_lib = None


def build(verbose=False):
    """Compile libheat3d.so and the heat3D driver in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.run(["make", "-C", CSRC, "all"] + ([] if verbose else ["-s"]), check=True)


def lib():
    """Load libheat3d.so (raises if it has not been built — there is no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} not built; run __graft_entry__.build() (make -C {CSRC})")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in EXPORTS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise Heat3DError(rc, lib().heat3d_last_error().decode(errors="replace"))


def coefficients(nx, ny, nz):
    """spacing, dt, Dx/Dy/Dz with the reference's expressions (heat3D.cu:352-367); host arithmetic only."""
    alpha, cfl = 1.0, 0.4
    sp = [1.0 / (n - 1.0) for n in (nx, ny, nz)]
    dt = cfl * 1.0 / (3 * 2) * pow(min(sp), 2.0) / alpha
    D = [dt * alpha / pow(s, 2.0) for s in sp]
    return sp, dt, D


def nccl_unique_id():
    buf = (C.c_ubyte * 128)()
    _check(lib().heat3d_nccl_unique_id(buf))
    return bytes(buf)


def pinned_empty(shape):
    """float64 ndarray backed by cudaHostAlloc memory (freed when the array's base object dies)."""
    n = int(np.prod(shape))
    p = C.c_void_p()
    _check(lib().heat3d_host_alloc(C.byref(p), n * 8))

    class _Owner:
        def __init__(self, ptr):
            self.ptr = ptr

        def __del__(self):
            try:
                lib().heat3d_host_free(self.ptr)
            except Exception:
                pass
    buf = (C.c_double * n).from_address(p.value)
    arr = np.frombuffer(buf, dtype=np.float64).reshape(shape)
    arr_owner = _Owner(p)
    # keep the owner alive as long as any view of the buffer is
    buf._owner = arr_owner
    return arr


class Heat3D:
    """One heat3d_ctx.  Fields cross the boundary as float64 arrays of shape (nz, ny, nx), x fastest."""

    def __init__(self, nx, ny, nz, D=None, arith=ARITH_STRICT, kernel=KERNEL_AUTO, temporal_k=1, nslabs=1,
                 slab_first=0, nlocal=0, devices=None, nccl_id=None, stream=None, flags=0, tune=None, nslabs_y=0):
        L = lib()
        self.nx, self.ny, self.nz = nx, ny, nz
        self.sp, self.dt, Dref = coefficients(nx, ny, nz)
        D = Dref if D is None else D
        p = Params()
        L.heat3d_default_params(C.byref(p), nx, ny, nz)
        p.Dx, p.Dy, p.Dz = D
        p.arith, p.kernel, p.temporal_k = arith, kernel, temporal_k
        p.nslabs, p.slab_first, p.nlocal = nslabs, slab_first, nlocal
        self._keep = []
        if devices is not None:
            arr = (C.c_int32 * len(devices))(*devices)
            self._keep.append(arr)
            p.devices = C.cast(arr, C.POINTER(C.c_int32))
        if nccl_id is not None:
            idb = (C.c_ubyte * 128).from_buffer_copy(nccl_id)
            self._keep.append(idb)
            p.nccl_unique_id = C.cast(idb, C.c_void_p)
        if stream is not None:
            p.stream = stream
        p.flags = flags
        p.nslabs_y = nslabs_y
        if tune:
            for i, v in enumerate(tune):
                p.tune[i] = v
        self._ctx = C.c_void_p()
        _check(L.heat3d_create(C.byref(p), C.byref(self._ctx)))

    # -- lifecycle
    def close(self):
        if getattr(self, "_ctx", None) and self._ctx.value:
            lib().heat3d_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- field
    def init_reference_field(self):
        _check(lib().heat3d_init_reference_field(self._ctx, self.sp[1]))

    def _arr(self, a, planes=None):
        shape = (self.nz if planes is None else planes, self.ny, self.nx)
        assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"] and a.shape == shape, (a.shape, shape)
        return a.ctypes.data_as(_dp)

    def upload(self, T):
        _check(lib().heat3d_upload(self._ctx, self._arr(T)))

    def download(self, out=None):
        out = np.empty((self.nz, self.ny, self.nx)) if out is None else out
        _check(lib().heat3d_download(self._ctx, self._arr(out)))
        return out

    def upload_planes(self, T, z0):
        _check(lib().heat3d_upload_planes(self._ctx, self._arr(T, T.shape[0]), z0, T.shape[0]))

    def upload_planes_async(self, T_pinned, z0):
        """Start the host->device copy of pinned planes; install them with upload_commit()."""
        _check(lib().heat3d_upload_planes_async(self._ctx, self._arr(T_pinned, T_pinned.shape[0]), z0, T_pinned.shape[0]))

    def upload_commit(self):
        _check(lib().heat3d_upload_commit(self._ctx))

    def download_planes(self, z0, n, out=None):
        out = np.empty((n, self.ny, self.nx)) if out is None else out
        _check(lib().heat3d_download_planes(self._ctx, self._arr(out, n), z0, n))
        return out

    def local_range(self):
        a, b = C.c_uint32(), C.c_uint32()
        _check(lib().heat3d_local_range(self._ctx, C.byref(a), C.byref(b)))
        return a.value, b.value

    # -- stepping
    def step(self, n):
        _check(lib().heat3d_step(self._ctx, n))

    def step_residual(self, n):
        r = C.c_double()
        _check(lib().heat3d_step_residual(self._ctx, n, C.byref(r)))
        return r.value

    def run(self, iter_max, eps, check_every=1):
        fi, cv, ap = C.c_uint32(), C.c_int32(), C.c_uint32()
        _check(lib().heat3d_run(self._ctx, iter_max, eps, check_every, C.byref(fi), C.byref(cv), C.byref(ap)))
        return ap.value, fi.value, bool(cv.value)

    def residual(self):
        r = C.c_double()
        _check(lib().heat3d_residual(self._ctx, C.byref(r)))
        return r.value

    def mean_abs_error_vs_y(self):
        r = C.c_double()
        _check(lib().heat3d_mean_abs_error_vs_y(self._ctx, self.sp[1], C.byref(r)))
        return r.value

    # -- snapshots / sync / diagnostics
    def snapshot_async(self, pinned):
        _check(lib().heat3d_snapshot_async(self._ctx, self._arr(pinned)))

    def snapshot_async_local(self, pinned_local):
        """As snapshot_async, for a pinned array that holds only THIS context's planes (shape (z1-z0, ny, nx)): the C
        entry point addresses planes by their global index, so it is handed the address plane 0 would have."""
        z0, z1 = self.local_range()
        assert pinned_local.dtype == np.float64 and pinned_local.flags["C_CONTIGUOUS"] and pinned_local.shape == (z1 - z0, self.ny, self.nx)
        base = pinned_local.ctypes.data - z0 * self.ny * self.nx * 8
        _check(lib().heat3d_snapshot_async(self._ctx, C.cast(C.c_void_p(base), _dp)))

    def snapshot_strided_async(self, pinned, sx, sy, sz):
        """Decimated snapshot into a pinned array of shape (ceil(nz/sz), ceil(ny/sy), ceil(nx/sx))."""
        shape = (-(-self.nz // sz), -(-self.ny // sy), -(-self.nx // sx))
        assert pinned.dtype == np.float64 and pinned.flags["C_CONTIGUOUS"] and pinned.shape == shape, (pinned.shape, shape)
        _check(lib().heat3d_snapshot_strided_async(self._ctx, pinned.ctypes.data_as(_dp), sx, sy, sz))

    def strided_local_shape(self, sx, sy, sz):
        """Shape of the decimated array holding only THIS context's planes."""
        z0, z1 = self.local_range()
        first = -(-z0 // sz) * sz
        return (max(0, -(-(z1 - first) // sz)) if first < z1 else 0, -(-self.ny // sy), -(-self.nx // sx))

    def snapshot_strided_async_local(self, pinned_local, sx, sy, sz):
        """As snapshot_strided_async for a pinned array of strided_local_shape(): the C entry point addresses decimated
        planes by their global index, so it is handed the address decimated plane 0 would have."""
        shape = self.strided_local_shape(sx, sy, sz)
        assert pinned_local.dtype == np.float64 and pinned_local.flags["C_CONTIGUOUS"] and pinned_local.shape == shape, (pinned_local.shape, shape)
        z0, _ = self.local_range()
        first_k = -(-z0 // sz)
        base = pinned_local.ctypes.data - first_k * shape[1] * shape[2] * 8
        _check(lib().heat3d_snapshot_strided_async(self._ctx, C.cast(C.c_void_p(base), _dp), sx, sy, sz))

    def snapshot_wait(self):
        _check(lib().heat3d_snapshot_wait(self._ctx))

    def sync(self):
        _check(lib().heat3d_sync(self._ctx))

    def timing(self):
        t = Timing()
        _check(lib().heat3d_get_timing(self._ctx, C.byref(t)))
        return {k: getattr(t, k) for k, _ in Timing._fields_}

    def reset_timing(self):
        _check(lib().heat3d_reset_timing(self._ctx))

    def layout(self, i=0):
        a, b, g = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(lib().heat3d_layout(self._ctx, i, C.byref(a), C.byref(b), C.byref(g)))
        return a.value, b.value, g.value

    def kernel_name(self):
        return lib().heat3d_kernel_name(self._ctx).decode()

    def exchange_name(self):
        return lib().heat3d_exchange_name(self._ctx).decode()

    def compute_stream(self, i=0):
        return lib().heat3d_compute_stream(self._ctx, i)


NOALLOC_MAX_PLANES = 8192


def mean_abs_error_vs_y_noalloc(d_field_ptr, nx, ny, nz, pitch_doubles, sp_y, d_err_ptr, stream=None):
    """heat3d_mean_abs_error_vs_y_noalloc on the caller's own device buffers (raw addresses); queues work on
    `stream` and returns - synchronise before reading d_err.  Allocates nothing; see include/heat3d.h for the limits."""
    _check(lib().heat3d_mean_abs_error_vs_y_noalloc(d_field_ptr, nx, ny, nz, pitch_doubles, sp_y, d_err_ptr, stream))


def device_count():
    n = C.c_int32()
    _check(lib().heat3d_device_count(C.byref(n)))
    return n.value


def slab_bounds(nz, nslabs):
    """Global plane ranges of the z-slabs, as the library cuts them (include/heat3d.h, Decomposition)."""
    return [(s * nz // nslabs, (s + 1) * nz // nslabs) for s in range(nslabs)]
