"""ctypes loader for oracle/libheat3d_oracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this
module.  The product (libheat3d.so, the heat3D driver, the ctypes mirror in the package)
never does.  Function-by-function citations into the reference are in heat3d_oracle.h.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libheat3d_oracle.so")
REF_BIN = os.path.join(_HERE, "_ref", "heat3D_ref")
STRICT, FMA = 0, 1
_lib = None


def build(ref=True):
    """Compile the C restatement (and, when /root/reference is present, oracle/_ref)."""
    subprocess.run(["make", "-s", "-C", _HERE, "oracle"] + (["ref"] if ref else []), check=True,
                   stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build(ref=False)
        L = C.CDLL(_LIB_PATH)
        dp, u, d, i = C.POINTER(C.c_double), C.c_uint, C.c_double, C.c_int
        L.h3o_coefficients.argtypes = [u, u, u, dp, dp, dp]
        L.h3o_init_field.argtypes = [dp, u, u, u, d]
        L.h3o_step_planes.argtypes = [dp, dp, u, u, u, d, d, d, u, u, i, i]
        L.h3o_steps.argtypes = [dp, dp, u, u, u, d, d, d, u, i, i]
        L.h3o_residual.argtypes = [dp, dp, u, u, u]
        L.h3o_residual.restype = d
        L.h3o_run.argtypes = [dp, dp, u, u, u, d, d, d, u, d, C.POINTER(u), C.POINTER(i), i, i]
        L.h3o_run.restype = u
        L.h3o_mean_abs_error_vs_y.argtypes = [dp, u, u, u, d]
        L.h3o_mean_abs_error_vs_y.restype = d
        L.h3o_write_dump.argtypes = [C.c_char_p, dp, u, u, u, dp]
        L.h3o_write_dump.restype = i
        _lib = L
    return _lib


def _p(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_double))


def coefficients(nx, ny, nz):
    """-> (sp[3], dt, D[3]) per heat3D.cu:352-367."""
    sp = np.zeros(3)
    D = np.zeros(3)
    dt = C.c_double()
    lib().h3o_coefficients(nx, ny, nz, _p(sp), C.byref(dt), _p(D))
    return sp, dt.value, D


def init_field(nx, ny, nz):
    """Field of shape (nz, ny, nx) (x fastest) with the reference IC/BC, heat3D.cu:409-453."""
    sp, _, _ = coefficients(nx, ny, nz)
    T = np.empty((nz, ny, nx))
    lib().h3o_init_field(_p(T), nx, ny, nz, sp[1])
    return T


def step_planes(src, dst, D, z0, z1, arith=STRICT, threads=1):
    nz, ny, nx = src.shape
    lib().h3o_step_planes(_p(src), _p(dst), nx, ny, nz, D[0], D[1], D[2], z0, z1, arith, threads)


def steps(T, D, nsteps, arith=STRICT, threads=1):
    """Advance T (nz,ny,nx) in place by nsteps; returns T."""
    nz, ny, nx = T.shape
    scratch = np.empty_like(T)
    lib().h3o_steps(_p(T), _p(scratch), nx, ny, nz, D[0], D[1], D[2], nsteps, arith, threads)
    return T


def residual(T, T0):
    nz, ny, nx = T.shape
    return lib().h3o_residual(_p(T), _p(T0), nx, ny, nz)


def run(T, D, iter_max, eps, arith=STRICT, threads=1):
    """Reference time loop; -> (applied_updates, final_iters, converged)."""
    nz, ny, nx = T.shape
    scratch = np.empty_like(T)
    fi = C.c_uint()
    cv = C.c_int()
    n = lib().h3o_run(_p(T), _p(scratch), nx, ny, nz, D[0], D[1], D[2], iter_max, eps,
                      C.byref(fi), C.byref(cv), arith, threads)
    return n, fi.value, bool(cv.value)


def mean_abs_error_vs_y(T):
    nz, ny, nx = T.shape
    sp, _, _ = coefficients(nx, ny, nz)
    return lib().h3o_mean_abs_error_vs_y(_p(T), nx, ny, nz, sp[1])


def write_dump(path, T):
    nz, ny, nx = T.shape
    sp, _, _ = coefficients(nx, ny, nz)
    rc = lib().h3o_write_dump(os.fsencode(path), _p(T), nx, ny, nz, _p(sp))
    if rc != 0:
        raise OSError(f"cannot write {path}")


def have_reference_binary():
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK)


def run_reference_binary(nx, ny, nz, iter_max, eps, workdir, raw=True, dump=True):
    """Run oracle/_ref/heat3D_ref (the reference's own host program) in `workdir`.

    `eps` is passed as typed (a string) so the echoed value matches a command-line run.
    -> dict(stdout=str, loop_seconds=float, raw=ndarray|None, dump_path=str|None)
    """
    if not have_reference_binary():
        raise FileNotFoundError(REF_BIN + " (build it in the dev container: make -C oracle ref)")
    os.makedirs(os.path.join(workdir, "output"), exist_ok=True)
    env = dict(os.environ)
    raw_path = os.path.join(workdir, "raw.bin")
    if raw:
        env["HEAT3D_REF_RAW"] = raw_path
    if not dump:
        env["HEAT3D_REF_NODUMP"] = "1"
    out = subprocess.run([REF_BIN, str(nx), str(ny), str(nz), str(iter_max), str(eps)],
                         cwd=workdir, env=env, check=True, capture_output=True, text=True).stdout
    secs = None
    for line in out.splitlines():
        if line.startswith("Computational time (parallel):"):
            secs = float(line.split(":")[1])
    field = np.fromfile(raw_path, dtype=np.float64).reshape(nz, ny, nx) if raw else None
    return dict(stdout=out, loop_seconds=secs, raw=field,
                dump_path=os.path.join(workdir, "output", "out.dat") if dump else None)
