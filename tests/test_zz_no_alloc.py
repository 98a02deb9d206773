"""GPU: the no-allocation error-norm entry point against the context-based one.

heat3d_mean_abs_error_vs_y_noalloc exists for hosts that must not allocate device memory (everything is owned
by their pool or caching allocator, or a cudaMalloc stall is unacceptable where they call us).  Here the
"host" is torch: the field and the result live in torch tensors, and the library is given their raw addresses.
The value must agree with heat3d_mean_abs_error_vs_y on a context holding the same field to 1e-13 (the two
reduce in different orders; 1e-13 is the tolerance the other error-norm tests use), on two calls in a row and
then on a smaller, different field - the scratch array is shared by all calls and keeps old contents, and
nothing stale may leak into a later answer.  Runs last (file name) so a failure here cannot hide other tests.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def h3():
    import __graft_entry__ as ge
    return ge.load_package()


def _expected(h3, T, oracle=None):
    nz, ny, nx = T.shape
    with h3.Heat3D(nx, ny, nz) as sim:
        sim.upload(T)
        want = sim.mean_abs_error_vs_y()
        if oracle is not None:      # and the context entry point itself against the CPU oracle's serial sum
            assert abs(want - oracle.mean_abs_error_vs_y(T)) < 1e-13
        return want, sim.sp[1]


def test_noalloc_matches_context_entry_point(h3, oracle):
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    rng = np.random.default_rng(11)
    big = oracle.steps(oracle.init_field(70, 37, 45), oracle.coefficients(70, 37, 45)[2], 30) + 0.01 * rng.random((45, 37, 70))
    small = rng.random((9, 20, 33))
    for T in (big, small):            # the second field is smaller: slots used by the first call stay behind, stale
        nz, ny, nx = T.shape
        want, sp_y = _expected(h3, T, oracle)
        d_field = torch.from_numpy(T).to(dev)
        d_err = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        for call in range(2):         # two calls in a row on the same input
            d_err.fill_(-1.0)
            free_before = torch.cuda.mem_get_info()[0]
            h3.mean_abs_error_vs_y_noalloc(d_field.data_ptr(), nx, ny, nz, nx, sp_y, d_err.data_ptr())
            torch.cuda.synchronize()
            got = float(d_err.item())
            assert abs(got - want) < 1e-13, (call, got, want)
            assert abs(got - oracle.mean_abs_error_vs_y(T)) < 1e-13, (call, got)
            if call == 1:             # by the second call the module is resident: no device memory may have been taken
                assert torch.cuda.mem_get_info()[0] == free_before
    # pitched input: same field stored with padded rows
    nz, ny, nx = small.shape
    want, sp_y = _expected(h3, small)
    padded = torch.zeros((nz, ny, nx + 7), dtype=torch.float64, device=dev)
    padded[:, :, :nx] = torch.from_numpy(small).to(dev)
    d_err = torch.zeros(1, dtype=torch.float64, device=dev)
    h3.mean_abs_error_vs_y_noalloc(padded.data_ptr(), nx, ny, nz, nx + 7, sp_y, d_err.data_ptr())
    torch.cuda.synchronize()
    assert abs(float(d_err.item()) - want) < 1e-13


def test_noalloc_rejects_oversize_and_bad_arguments(h3):
    import torch
    torch.cuda.set_device(0)
    L = h3.lib()
    dummy = torch.zeros(8, dtype=torch.float64, device="cuda")
    p = ctypes.c_void_p(dummy.data_ptr())
    too_many = h3.NOALLOC_MAX_PLANES + 3          # nz - 2 interior planes = max + 1; refused before memory is touched
    assert L.heat3d_mean_abs_error_vs_y_noalloc(p, 4, 4, too_many, 4, 0.1, p, None) == -1
    assert b"HEAT3D_NOALLOC_MAX_PLANES" in L.heat3d_last_error()
    assert L.heat3d_mean_abs_error_vs_y_noalloc(None, 4, 4, 4, 4, 0.1, p, None) == -1
    assert L.heat3d_mean_abs_error_vs_y_noalloc(p, 4, 4, 4, 3, 0.1, p, None) == -1      # pitch < nx
    assert L.heat3d_mean_abs_error_vs_y_noalloc(p, 2, 4, 4, 4, 0.1, p, None) == -1
