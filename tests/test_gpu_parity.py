"""GPU: parity of libheat3d (through the C ABI) against the CPU oracle and the committed golden fields.

Tolerances: STRICT arithmetic is bit-exact (np.array_equal); FMA arithmetic is compared with the oracle's
FMA ordering bit-exactly and with the strict oracle to 1e-12 max-abs and max-rel (the north-star bound).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def h3():
    import __graft_entry__ as ge
    return ge.load_package()


KERNELS = [("baseline", dict(kernel=1)), ("tma", dict(kernel=2))]
TUNES = [(128, 8, 2, 4, 3), (128, 8, 4, 4, 4), (128, 16, 2, 4, 1), (128, 16, 4, 4, 2), (128, 32, 4, 4, 1), (64, 16, 4, 4, 4)]


def _oracle_field(oracle, nx, ny, nz, nsteps, arith=0):
    _, _, D = oracle.coefficients(nx, ny, nz)
    return oracle.steps(oracle.init_field(nx, ny, nz), D, nsteps, arith=arith, threads=8)


@pytest.mark.parametrize("kname,kw", KERNELS)
@pytest.mark.parametrize("shape,nsteps", [((27, 27, 27), 40), ((33, 17, 65), 25), ((64, 64, 64), 30), ((5, 4, 3), 7),
                                          ((3, 3, 3), 3), ((130, 35, 9), 12), ((257, 66, 20), 9), ((3, 200, 40), 5)])
def test_strict_bit_exact_vs_oracle(h3, oracle, kname, kw, shape, nsteps):
    nx, ny, nz = shape
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    with h3.Heat3D(nx, ny, nz, **kw) as sim:
        sim.init_reference_field()
        sim.step(nsteps)
        got = sim.download()
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("tune", TUNES)
def test_every_tma_variant_bit_exact(h3, oracle, tune):
    nx, ny, nz, nsteps = 150, 70, 23, 11
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    with h3.Heat3D(nx, ny, nz, kernel=2, tune=tune) as sim:
        assert sim.kernel_name() == "tma_%dx%d_r%d_s%d_c%d" % tune
        sim.init_reference_field()
        sim.step(nsteps - 1)
        res = sim.step_residual(1)
        got = sim.download()
    assert np.array_equal(got, ref)
    prev = _oracle_field(oracle, nx, ny, nz, nsteps - 1)
    assert res == oracle.residual(ref, prev)


@pytest.mark.parametrize("kname,kw", KERNELS)
def test_fma_mode(h3, oracle, kname, kw):
    nx, ny, nz, nsteps = 48, 40, 33, 200
    with h3.Heat3D(nx, ny, nz, arith=h3.ARITH_FMA, **kw) as sim:
        sim.init_reference_field()
        sim.step(nsteps)
        got = sim.download()
    assert np.array_equal(got, _oracle_field(oracle, nx, ny, nz, nsteps, arith=1))
    strict = _oracle_field(oracle, nx, ny, nz, nsteps)
    assert np.max(np.abs(got - strict)) <= 1e-12
    nzm = strict != 0
    assert np.max(np.abs(got[nzm] - strict[nzm]) / np.abs(strict[nzm])) <= 1e-12


@pytest.mark.parametrize("name", ["c27_converge", "aniso_33x17x65", "tiny_5x4x3"])
def test_golden_fields_from_reference_program(h3, manifest, golden_field, name):
    """Same run as the reference's own host program made (tests/golden/manifest.json): convergence loop
    with the residual checked every iteration, final field bit-identical, same iteration count."""
    nx, ny, nz, it, eps = manifest[name]["args"]
    G = golden_field(name)
    with h3.Heat3D(nx, ny, nz) as sim:
        sim.init_reference_field()
        applied, final_iters, conv = sim.run(it, float(eps), check_every=1)
        got = sim.download()
        err = sim.mean_abs_error_vs_y()
    assert np.array_equal(got, G)
    report = manifest[name]["stdout"]
    if conv:
        assert report[-2].startswith(f"Simulation has converged in {final_iters} iterations")
    else:
        assert report[-2] == f"Simulation did not converge within {it} iterations."
    assert report[-1] == "L2-norm error: %.4f %%" % (100 * err)


def test_init_field_matches_oracle(h3, oracle):
    for shape in [(9, 7, 5), (64, 64, 64), (130, 3, 17)]:
        with h3.Heat3D(*shape) as sim:
            sim.init_reference_field()
            assert np.array_equal(sim.download(), oracle.init_field(*shape))


def test_upload_download_roundtrip_and_step(h3, oracle):
    nx, ny, nz = 45, 31, 22
    rng = np.random.default_rng(7)
    T = rng.random((nz, ny, nx))
    _, _, D = oracle.coefficients(nx, ny, nz)
    with h3.Heat3D(nx, ny, nz) as sim:
        sim.upload(T)
        assert np.array_equal(sim.download(), T)
        sim.step(5)
        got = sim.download()
        part = sim.download_planes(3, 4)
    ref = oracle.steps(T.copy(), D, 5)
    assert np.array_equal(got, ref)
    assert np.array_equal(part, ref[3:7])


def test_residual_and_check_every(h3, oracle):
    nx = ny = nz = 27
    _, _, D = oracle.coefficients(nx, ny, nz)
    with h3.Heat3D(nx, ny, nz) as sim:
        sim.init_reference_field()
        r0 = sim.step_residual(1)
        assert sim.residual() == r0
        sim.step(1)
        with pytest.raises(h3.Heat3DError):
            sim.residual()
    a = oracle.init_field(nx, ny, nz)
    b = oracle.steps(a.copy(), D, 1)
    assert r0 == oracle.residual(b, a)
    # check_every > 1 can only overshoot the reference's iteration count, by less than check_every
    with h3.Heat3D(nx, ny, nz) as sim:
        sim.init_reference_field()
        applied, fin, conv = sim.run(100000, 1e-6, check_every=16)
    assert conv and 3302 <= fin < 3302 + 16 and applied == fin + 1


def test_steady_state_fixed_point_and_max_principle(h3):
    nx, ny, nz = 40, 33, 29
    with h3.Heat3D(nx, ny, nz) as sim:
        y = np.arange(ny) * sim.sp[1]
        T = np.broadcast_to(y[None, :, None], (nz, ny, nx)).copy()
        sim.upload(T)
        sim.step(50)
        assert np.max(np.abs(sim.download() - T)) < 1e-15
        sim.init_reference_field()
        sim.step(300)
        U = sim.download()
    assert U.min() >= 0.0 and U.max() <= 1.0


@pytest.mark.parametrize("kernel", [1, 2])
@pytest.mark.parametrize("shape,P,nsteps", [((40, 24, 31), 2, 20), ((21, 13, 30), 4, 25), ((70, 40, 17), 3, 13),
                                            ((33, 17, 9), 8, 6)])
def test_multi_slab_on_one_gpu_equals_single_domain_bitwise(h3, oracle, kernel, shape, P, nsteps):
    """z-slabs with ghost planes exchanged by peer copies (all slabs placed on device 0) must reproduce the
    single-domain field bit for bit, overlapped or not."""
    nx, ny, nz = shape
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    for flags in (0, h3.FLAG_NO_OVERLAP, h3.FLAG_TIME_HALO):
        with h3.Heat3D(nx, ny, nz, kernel=kernel, nslabs=P, devices=[0] * P, flags=flags) as sim:
            sim.init_reference_field()
            sim.step(nsteps - 1)
            res = sim.step_residual(1)
            got = sim.download()
            err = sim.mean_abs_error_vs_y()
        assert np.array_equal(got, ref), flags
        assert res == oracle.residual(ref, _oracle_field(oracle, nx, ny, nz, nsteps - 1))
        assert abs(err - oracle.mean_abs_error_vs_y(ref)) < 1e-13


def test_multi_slab_upload(h3, oracle):
    nx, ny, nz, P = 30, 20, 26, 3
    T = np.random.default_rng(3).random((nz, ny, nx))
    _, _, D = oracle.coefficients(nx, ny, nz)
    with h3.Heat3D(nx, ny, nz, nslabs=P, devices=[0] * P) as sim:
        sim.upload(T)
        sim.step(9)
        got = sim.download()
    assert np.array_equal(got, oracle.steps(T.copy(), D, 9))


def test_snapshot_async_overlaps_stepping(h3, oracle):
    nx, ny, nz = 64, 48, 40
    _, _, D = oracle.coefficients(nx, ny, nz)
    with h3.Heat3D(nx, ny, nz) as sim:
        sim.init_reference_field()
        sim.step(10)
        snap = h3.pinned_empty((nz, ny, nx))
        sim.snapshot_async(snap)
        sim.step(7)                     # keeps going while the copy is in flight
        sim.snapshot_wait()
        later = sim.download()
        assert np.array_equal(snap, _oracle_field(oracle, nx, ny, nz, 10))
        assert np.array_equal(later, _oracle_field(oracle, nx, ny, nz, 17))


@pytest.mark.parametrize("P", [1, 3])
def test_strided_snapshot_matches_decimated_oracle(h3, oracle, P):
    """heat3d_snapshot_strided_async packs every (sx,sy,sz)-th point on the device; the pinned array must be the
    oracle's field decimated the same way, while stepping continues underneath."""
    nx, ny, nz = 64, 48, 41
    with h3.Heat3D(nx, ny, nz, nslabs=P, devices=[0] * P) as sim:
        sim.init_reference_field()
        sim.step(10)
        for sx, sy, sz in ((2, 3, 4), (1, 1, 5), (7, 1, 1), (1, 1, 1)):
            snap = h3.pinned_empty((-(-nz // sz), -(-ny // sy), -(-nx // sx)))
            snap[:] = -1.0
            sim.snapshot_strided_async(snap, sx, sy, sz)
            sim.step(3)
            sim.snapshot_wait()
            want = _oracle_field(oracle, nx, ny, nz, 10)[::sz, ::sy, ::sx]
            assert np.array_equal(snap, want), (sx, sy, sz)
            sim.init_reference_field()
            sim.step(10)
    assert h3.device_count() >= 1


def test_large_grid_properties(h3):
    """BASELINE config-2 size (512^3): size-independent checks instead of an oracle run —
    x<->z transpose symmetry and x/z reflection symmetry to rounding, max principle, monotone residual."""
    n = 512
    with h3.Heat3D(n, n, n) as sim:
        sim.init_reference_field()
        r1 = sim.step_residual(1)
        sim.step(98)
        r2 = sim.step_residual(1)
        T = sim.download()
        t = sim.timing()
    assert 0 < r2 < r1
    assert T.min() >= 0.0 and T.max() <= 1.0
    assert np.max(np.abs(T - T.transpose(2, 1, 0))) < 1e-13
    assert np.max(np.abs(T - T[::-1, :, :])) < 1e-13
    assert np.max(np.abs(T - T[:, :, ::-1])) < 1e-13
    assert t["steps"] == 100 and t["launches"] >= 100


@pytest.mark.parametrize("k", [1, 2, 4])
def test_multi_device_one_process_bitwise(h3, oracle, k):
    """Slabs on distinct GPUs driven by one context (cudaMemcpyPeerAsync between devices)."""
    import torch
    nd = min(torch.cuda.device_count(), 4)
    if nd < 2:
        pytest.skip("needs >= 2 GPUs")
    nx, ny, nz, nsteps = 150, 70, 46, 15
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    with h3.Heat3D(nx, ny, nz, temporal_k=k, nslabs=nd, devices=list(range(nd))) as sim:
        assert sim.exchange_name() == "peer-copy"
        sim.init_reference_field()
        res = sim.step_residual(nsteps)
        got = sim.download()
    assert np.array_equal(got, ref)
    assert res == oracle.residual(ref, _oracle_field(oracle, nx, ny, nz, nsteps - 1))


@pytest.mark.parametrize("P", [1, 3])
def test_run_every_iteration_is_exact_and_state_stays_consistent(h3, oracle, P):
    """check_every=1 queues iteration i+1 before the host has seen residual i and drops it again on convergence:
    iteration count, field and later stepping must be exactly the reference's."""
    nx = ny = nz = 27
    _, _, D = oracle.coefficients(nx, ny, nz)
    with h3.Heat3D(nx, ny, nz, nslabs=P, devices=[0] * P) as sim:
        sim.init_reference_field()
        applied, fin, conv = sim.run(100000, 1e-6, check_every=1)
        assert (applied, fin, conv) == (3303, 3302, True)
        assert sim.timing()["steps"] == 3303
        T = sim.download()
        assert np.array_equal(T, _oracle_field(oracle, nx, ny, nz, 3303))
        assert sim.residual() == oracle.residual(T, _oracle_field(oracle, nx, ny, nz, 3302))
        sim.step(5)                                   # the dropped speculative iteration must leave no trace
        assert np.array_equal(sim.download(), _oracle_field(oracle, nx, ny, nz, 3308))
    with h3.Heat3D(nx, ny, nz, nslabs=P, devices=[0] * P) as sim:      # not converging: exactly iter_max updates
        sim.init_reference_field()
        assert sim.run(37, 1e-30, check_every=1) == (37, 0, False)
        assert np.array_equal(sim.download(), _oracle_field(oracle, nx, ny, nz, 37))


@pytest.mark.parametrize("k", [1, 2])
def test_long_runs_through_cuda_graphs(h3, oracle, k):
    """heat3d_step replays graphs of 32 passes on single-slab contexts; odd leftovers take plain launches."""
    nx, ny, nz, nsteps = 40, 33, 29, 2 * 32 * k + 32 * k + 7          # three graph launches and a remainder
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    with h3.Heat3D(nx, ny, nz, temporal_k=k) as sim:
        sim.init_reference_field()
        sim.step(nsteps)
        t = sim.timing()
        assert t["steps"] == nsteps
        assert np.array_equal(sim.download(), ref)
        sim.step(1)                                    # buffer parity flipped: the other graph gets built next
        sim.step(32 * k)
        assert np.array_equal(sim.download(), _oracle_field(oracle, nx, ny, nz, nsteps + 1 + 32 * k))


@pytest.mark.parametrize("shape,py,pz,nsteps", [((40, 24, 31), 2, 1, 20), ((45, 31, 22), 3, 2, 17), ((130, 35, 29), 2, 3, 12), ((33, 64, 16), 4, 2, 9),
                                                ((150, 70, 40), 5, 1, 11)])
def test_2d_decomposition_equals_single_domain_bitwise(h3, oracle, shape, py, pz, nsteps):
    """nslabs_y > 1: ranges of rows times z-slabs.  The faces between row ranges are strided in memory and travel through
    the pack / unpack kernels (the reference's host-side face packing, heat3D.cu:562-605); ownership is non-overlapping,
    so field, residual, error norm, uploads and the convergence loop must be the single domain's, bit for bit."""
    nx, ny, nz = shape
    _, _, D = oracle.coefficients(nx, ny, nz)
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    prev = _oracle_field(oracle, nx, ny, nz, nsteps - 1)
    with h3.Heat3D(nx, ny, nz, nslabs=pz, nslabs_y=py, devices=[0] * (py * pz)) as sim:
        assert sim.exchange_name().startswith("y: pack/unpack")
        sim.init_reference_field()
        assert np.array_equal(sim.download(), oracle.init_field(nx, ny, nz))
        sim.step(nsteps - 1)
        res = sim.step_residual(1)
        got = sim.download()
        assert np.array_equal(got, ref)
        assert res == oracle.residual(ref, prev)
        want_err = oracle.mean_abs_error_vs_y(ref)       # serial sum there, per-range tree sums here: relative 1e-12
        assert abs(sim.mean_abs_error_vs_y() - want_err) < 1e-12 * want_err
        t = sim.timing()
        assert t["steps"] == nsteps and t["last_call_ms"] > 0
        T = np.random.default_rng(12).random((nz, ny, nx))
        sim.upload(T)
        assert np.array_equal(sim.download(), T)
        sim.step(6)
        assert np.array_equal(sim.download(), oracle.steps(T.copy(), D, 6))
        assert np.array_equal(sim.download_planes(3, 5), oracle.steps(T.copy(), D, 6)[3:8])
    with pytest.raises(h3.Heat3DError):
        h3.Heat3D(nx, ny, nz, nslabs_y=2, temporal_k=2, devices=[0, 0])


def test_2d_decomposition_convergence_run_matches_reference_program(h3, manifest, golden_field):
    nx, ny, nz, it, eps = manifest["c27_converge"]["args"]
    with h3.Heat3D(nx, ny, nz, nslabs=2, nslabs_y=3, devices=[0] * 6) as sim:
        sim.init_reference_field()
        applied, fin, conv = sim.run(it, float(eps), check_every=1)
        assert (applied, fin, conv) == (3303, 3302, True)
        assert np.array_equal(sim.download(), golden_field("c27_converge"))
