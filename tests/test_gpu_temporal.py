"""GPU: temporal blocking (temporal_k=2: two FTCS updates per HBM pass) must be bit-identical to stepping twice.

Same bar as tests/test_gpu_parity.py: strict arithmetic bit-exact against the CPU oracle (np.array_equal), fma
arithmetic bit-exact against the oracle's fma ordering."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

# cx,cy,rows/thread,stages,CTAs/SM,chunk_len,xmode  (shared-memory access mode, kernels_t2.cuh: 0 "_L" plain loads,
# 2 "_P" parity-split mid plane, 3 "_X" that + swapped x-neighbour loads from the stage, 6 "_Q" = "_P" with
# neighbour-to-neighbour barriers)
T2_TUNES = ["64,32,4,4,2,7,2", "64,32,4,4,2,7,0", "64,32,4,4,2,7,3", "64,32,4,4,2,7,6", "64,16,2,4,3,7,2", "128,32,4,4,1,7,2", "64,64,4,4,1,7,2"]


@pytest.fixture(scope="module")
def h3():
    import __graft_entry__ as ge
    return ge.load_package()


@pytest.fixture(autouse=True)
def _clean_env():
    os.environ.pop("HEAT3D_T2_TUNE", None)
    yield
    os.environ.pop("HEAT3D_T2_TUNE", None)


def _oracle_field(oracle, nx, ny, nz, nsteps, arith=0):
    _, _, D = oracle.coefficients(nx, ny, nz)
    return oracle.steps(oracle.init_field(nx, ny, nz), D, nsteps, arith=arith, threads=8)


@pytest.mark.parametrize("shape,nsteps", [((27, 27, 27), 40), ((33, 17, 65), 25), ((64, 64, 64), 30), ((5, 4, 3), 7), ((3, 3, 3), 4),
                                          ((130, 35, 9), 12), ((257, 66, 20), 9), ((3, 200, 40), 6), ((61, 31, 4), 2), ((125, 64, 37), 1)])
def test_depth2_bit_exact_vs_oracle(h3, oracle, shape, nsteps):
    nx, ny, nz = shape
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    with h3.Heat3D(nx, ny, nz, temporal_k=2) as sim:
        assert sim.kernel_name().startswith("t2_")
        assert sim.layout()[2] == 2                   # ghost depth follows the temporal depth
        sim.init_reference_field()
        sim.step(nsteps)
        got = sim.download()
        assert sim.timing()["steps"] == nsteps
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("tune", T2_TUNES)
def test_every_depth2_variant(h3, oracle, tune):
    nx, ny, nz, nsteps = 150, 70, 41, 10
    f = tune.split(",")
    os.environ["HEAT3D_T2_TUNE"] = tune        # chunk_len 7: several chunk seams inside the grid
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    prev = _oracle_field(oracle, nx, ny, nz, nsteps - 1)
    with h3.Heat3D(nx, ny, nz, temporal_k=2) as sim:
        assert sim.kernel_name() == "t2_%sx%s_r%s_s%s_c%s" % tuple(f[:5]) + {"0": "_L", "2": "_P", "3": "_X", "6": "_Q"}[f[6]]
        sim.init_reference_field()
        res = sim.step_residual(nsteps)               # even count: the residual comes out of the depth-2 kernel
        got = sim.download()
    assert np.array_equal(got, ref)
    assert res == oracle.residual(ref, prev)


def test_depth2_residual_odd_and_even(h3, oracle):
    nx, ny, nz = 48, 40, 33
    for nsteps in (1, 2, 3, 8, 9):
        ref = _oracle_field(oracle, nx, ny, nz, nsteps)
        prev = _oracle_field(oracle, nx, ny, nz, nsteps - 1)
        with h3.Heat3D(nx, ny, nz, temporal_k=2) as sim:
            sim.init_reference_field()
            res = sim.step_residual(nsteps)
            got = sim.download()
        assert np.array_equal(got, ref), nsteps
        assert res == oracle.residual(ref, prev), nsteps


def test_depth2_fma_mode(h3, oracle):
    nx, ny, nz, nsteps = 48, 40, 33, 100
    with h3.Heat3D(nx, ny, nz, temporal_k=2, arith=h3.ARITH_FMA) as sim:
        sim.init_reference_field()
        sim.step(nsteps)
        got = sim.download()
    assert np.array_equal(got, _oracle_field(oracle, nx, ny, nz, nsteps, arith=1))
    assert np.max(np.abs(got - _oracle_field(oracle, nx, ny, nz, nsteps))) <= 1e-12


def test_depth2_random_field_and_convergence_run(h3, oracle, manifest, golden_field):
    nx, ny, nz = 45, 31, 22
    T = np.random.default_rng(7).random((nz, ny, nx))
    _, _, D = oracle.coefficients(nx, ny, nz)
    with h3.Heat3D(nx, ny, nz, temporal_k=2) as sim:
        sim.upload(T)
        sim.step(6)
        assert np.array_equal(sim.download(), oracle.steps(T.copy(), D, 6))
    # the reference's convergence run, residual examined every 2nd iteration so the depth-2 kernel does the work
    nx, ny, nz, it, eps = manifest["c27_converge"]["args"]
    with h3.Heat3D(nx, ny, nz, temporal_k=2) as sim:
        sim.init_reference_field()
        applied, fin, conv = sim.run(it, float(eps), check_every=2)
        got = sim.download()
    assert conv and fin in (3302, 3303) and applied == fin + 1
    assert np.array_equal(got, _oracle_field(oracle, nx, ny, nz, applied))


@pytest.mark.parametrize("shape,P,nsteps", [((40, 24, 31), 2, 20), ((21, 13, 30), 4, 25), ((70, 40, 17), 3, 13), ((33, 17, 16), 8, 6)])
def test_depth2_multi_slab_equals_single_domain_bitwise(h3, oracle, shape, P, nsteps):
    nx, ny, nz = shape
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    prev = _oracle_field(oracle, nx, ny, nz, nsteps - 1)
    for flags in (0, h3.FLAG_NO_OVERLAP):
        with h3.Heat3D(nx, ny, nz, temporal_k=2, nslabs=P, devices=[0] * P, flags=flags) as sim:
            sim.init_reference_field()
            res = sim.step_residual(nsteps)
            got = sim.download()
        assert np.array_equal(got, ref), flags
        assert res == oracle.residual(ref, prev)


@pytest.mark.parametrize("guided", ["1", "0"])
def test_depth2_guided_chunk_plan_and_partial_tiles(h3, oracle, guided):
    """The default launch plan (long chunks first, short ones last) against uniform chunks, on a grid with partial tiles
    at the high x and y edges (warps that lie wholly outside the domain sit out the compute loop) and enough planes for
    several chunk seams."""
    nx, ny, nz, nsteps = 129, 95, 300, 6
    os.environ["HEAT3D_T2_GUIDED"] = guided
    try:
        ref = _oracle_field(oracle, nx, ny, nz, nsteps)
        with h3.Heat3D(nx, ny, nz, temporal_k=2) as sim:
            sim.init_reference_field()
            sim.step(nsteps)
            assert np.array_equal(sim.download(), ref)
    finally:
        os.environ.pop("HEAT3D_T2_GUIDED", None)


@pytest.mark.parametrize("k", [1, 2])
@pytest.mark.parametrize("shape,P,nsteps", [((40, 24, 31), 2, 20), ((70, 40, 37), 3, 13), ((150, 70, 64), 4, 9), ((33, 17, 41), 8, 6)])
def test_peer_store_exchange_bitwise(h3, oracle, k, shape, P, nsteps):
    """HEAT3D_EXCHANGE=store: the stencil kernel writes its outer planes straight into the neighbouring slabs' ghost planes
    and publishes the arrival flag itself (no copies, no exchange stream).  Same bar: bit-identical to the single domain,
    also after an upload (copy-based ghost refresh) and across the speculative step of the convergence loop."""
    nx, ny, nz = shape
    os.environ["HEAT3D_EXCHANGE"] = "store"
    try:
        ref = _oracle_field(oracle, nx, ny, nz, nsteps)
        prev = _oracle_field(oracle, nx, ny, nz, nsteps - 1)
        with h3.Heat3D(nx, ny, nz, temporal_k=k, nslabs=P, devices=[0] * P, flags=h3.FLAG_TIME_HALO) as sim:
            assert sim.exchange_name() == "peer-store"
            sim.init_reference_field()
            res = sim.step_residual(nsteps)
            assert np.array_equal(sim.download(), ref)
            assert res == oracle.residual(ref, prev)
            T = np.random.default_rng(9).random((nz, ny, nx))
            _, _, D = oracle.coefficients(nx, ny, nz)
            sim.upload(T)
            sim.step(7)
            assert np.array_equal(sim.download(), oracle.steps(T.copy(), D, 7))
            sim.init_reference_field()
            applied, fin, conv = sim.run(11, 1e-30, check_every=1)
            assert (applied, conv) == (11, False)
            assert np.array_equal(sim.download(), _oracle_field(oracle, nx, ny, nz, 11))
    finally:
        os.environ.pop("HEAT3D_EXCHANGE", None)


def test_peer_store_falls_back_on_thin_slabs(h3, oracle):
    os.environ["HEAT3D_EXCHANGE"] = "store"
    try:
        with h3.Heat3D(21, 13, 12, temporal_k=2, nslabs=4, devices=[0] * 4) as sim:     # 3 planes per slab < 2 * ghost depth
            assert sim.exchange_name() == "peer-copy"
            sim.init_reference_field()
            sim.step(6)
            assert np.array_equal(sim.download(), _oracle_field(oracle, 21, 13, 12, 6))
    finally:
        os.environ.pop("HEAT3D_EXCHANGE", None)


@pytest.mark.parametrize("P", [1, 3])
def test_depth2_every_iteration_convergence_loop_is_exact(h3, oracle, P):
    """check_every=1 on a depth-2 context: each pass reports the residuals of BOTH its updates.  Iteration counts and
    fields must be the reference's whether the converging iteration is the first or the second of a pass."""
    nx = ny = nz = 27
    _, _, D = oracle.coefficients(nx, ny, nz)
    parities = set()
    for eps, iter_max in ((1e-6, 100000), (2e-6, 100000), (3e-6, 100000), (5e-6, 100000), (1e-30, 37), (1e-30, 40)):
        T = oracle.init_field(nx, ny, nz)
        applied_o, fin_o, conv_o = oracle.run(T, D, iter_max, eps)
        if conv_o:
            parities.add(fin_o % 2)
        with h3.Heat3D(nx, ny, nz, temporal_k=2, nslabs=P, devices=[0] * P) as sim:
            sim.init_reference_field()
            assert sim.run(iter_max, eps, check_every=1) == (applied_o, fin_o, conv_o)
            assert sim.timing()["steps"] == applied_o
            assert np.array_equal(sim.download(), T)
            sim.step(3)
            assert np.array_equal(sim.download(), oracle.steps(T.copy(), D, 3))
    assert parities == {0, 1}, parities          # both kinds of ending were exercised


def test_depth2_first_update_residual_matches_oracle(h3, oracle):
    nx, ny, nz = 70, 37, 45
    _, _, D = oracle.coefficients(nx, ny, nz)
    a = oracle.steps(oracle.init_field(nx, ny, nz), D, 8)
    b = oracle.steps(a.copy(), D, 1)
    c2 = oracle.steps(b.copy(), D, 1)
    with h3.Heat3D(nx, ny, nz, temporal_k=2, nslabs=2, devices=[0, 0]) as sim:
        sim.init_reference_field()
        sim.step(8)
        assert sim.step_residual(2) == oracle.residual(c2, b)       # the second update's residual is what this call returns
        assert np.array_equal(sim.download(), c2)


@pytest.mark.parametrize("shape,P,nsteps", [((40, 24, 31), 2, 20), ((21, 13, 34), 4, 27), ((70, 40, 29), 3, 13), ((33, 17, 33), 8, 9),
                                            ((150, 70, 64), 4, 16)])
def test_depth4_blocks_multi_slab_equal_single_domain_bitwise(h3, oracle, shape, P, nsteps):
    """temporal_k=4: ghost depth 4, blocks of four updates = two depth-2 passes and ONE exchange (the first pass also
    advances two ghost planes per side).  Same bar as every other decomposition: bit-identical to the single domain -
    with step counts that leave depth-2 and depth-1 tails, residuals out of the last block, uploads, and the
    every-iteration convergence loop (which exchanges after every pass)."""
    nx, ny, nz = shape
    _, _, D = oracle.coefficients(nx, ny, nz)
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    prev = _oracle_field(oracle, nx, ny, nz, nsteps - 1)
    with h3.Heat3D(nx, ny, nz, temporal_k=4, nslabs=P, devices=[0] * P, flags=h3.FLAG_TIME_HALO) as sim:
        assert sim.layout()[2] == 4
        sim.init_reference_field()
        res = sim.step_residual(nsteps)
        assert np.array_equal(sim.download(), ref)
        assert res == oracle.residual(ref, prev)
        sim.step(8)                                    # two whole blocks on top
        assert np.array_equal(sim.download(), _oracle_field(oracle, nx, ny, nz, nsteps + 8))
        res8 = sim.step_residual(8)                    # residual out of a depth-4 block
        want = _oracle_field(oracle, nx, ny, nz, nsteps + 16)
        assert np.array_equal(sim.download(), want)
        assert res8 == oracle.residual(want, _oracle_field(oracle, nx, ny, nz, nsteps + 15))
        T = np.random.default_rng(4).random((nz, ny, nx))
        sim.upload(T)
        sim.step(11)
        assert np.array_equal(sim.download(), oracle.steps(T.copy(), D, 11))
        sim.init_reference_field()
        assert sim.run(13, 1e-30, check_every=1) == (13, 0, False)
        assert np.array_equal(sim.download(), _oracle_field(oracle, nx, ny, nz, 13))
    with h3.Heat3D(nx, ny, nz, temporal_k=4) as one:   # a single slab has nobody to talk to: plain depth-2 passes
        one.init_reference_field()
        one.step(nsteps)
        assert np.array_equal(one.download(), ref)
