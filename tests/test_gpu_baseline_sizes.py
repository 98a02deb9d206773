"""GPU: the BASELINE.json sizes against the CPU oracle itself (not against size-independent properties, and not against
another path of this library).

 - config 1 (64^3 x 1000 updates): sha256 of the raw fp64 field equals the digest of the field the reference's own host
   program produced (tests/golden/manifest.json, made by tests/golden/make_golden.py from oracle/_ref).
 - config 2's grid (512^3): 21 updates (odd: the depth-2 path ends with a depth-1 launch), both temporal depths,
   np.array_equal against oracle.steps on all host cores.  This is the 9 x 17 (depth 2) / 4 x 64 (depth 1) tile grid and
   the multi-wave chunk schedule the benchmark runs.
 - one slab-shaped case with 1024-wide planes (configs 3/4: 1024 x 1024 x 24, 18 x 35 depth-2 tiles, several chunks).
Tolerance: none - strict arithmetic is bit-exact (heat3D.cu:128-131 evaluated in the host compiler's order)."""
import hashlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def h3():
    import __graft_entry__ as ge
    return ge.load_package()


def _oracle_field(oracle, nx, ny, nz, nsteps):
    _, _, D = oracle.coefficients(nx, ny, nz)
    return oracle.steps(oracle.init_field(nx, ny, nz), D, nsteps, threads=os.cpu_count() or 8)


@pytest.mark.parametrize("k", [1, 2])
def test_config1_raw_field_digest(h3, manifest, k):
    nx, ny, nz, it, eps = manifest["c64_1000"]["args"]
    with h3.Heat3D(nx, ny, nz, temporal_k=k) as sim:
        sim.init_reference_field()
        applied, fin, conv = sim.run(it, float(eps), check_every=1)
        got = sim.download()
    assert (applied, conv) == (it, False)
    assert hashlib.sha256(np.ascontiguousarray(got).tobytes()).hexdigest() == manifest["c64_1000"]["raw_sha256"]


@pytest.fixture(scope="module")
def oracle_512(oracle):
    _, _, D = oracle.coefficients(512, 512, 512)
    prev = _oracle_field(oracle, 512, 512, 512, 20)
    ref = oracle.steps(prev.copy(), D, 1, threads=os.cpu_count() or 8)
    return ref, oracle.residual(ref, prev)


@pytest.mark.parametrize("k", [1, 2])
def test_config2_grid_bit_exact_vs_oracle(h3, oracle, oracle_512, k):
    n, nsteps = 512, 21
    with h3.Heat3D(n, n, n, temporal_k=k) as sim:
        sim.init_reference_field()
        sim.step(nsteps - 1)
        res = sim.step_residual(1)
        got = sim.download()
    ref, want_res = oracle_512
    assert np.array_equal(got, ref)
    assert res == want_res


@pytest.mark.parametrize("k", [1, 2])
def test_1024_wide_slab_bit_exact_vs_oracle(h3, oracle, k):
    nx, ny, nz, nsteps = 1024, 1024, 24, 9
    ref = _oracle_field(oracle, nx, ny, nz, nsteps)
    os.environ["HEAT3D_T2_TUNE"] = "64,32,4,4,2,5,2"       # chunks of 5 planes: several seams inside 22 interior planes
    try:
        with h3.Heat3D(nx, ny, nz, temporal_k=k) as sim:
            sim.init_reference_field()
            sim.step(nsteps)
            assert np.array_equal(sim.download(), ref)
    finally:
        os.environ.pop("HEAT3D_T2_TUNE", None)
    with h3.Heat3D(nx, ny, nz, temporal_k=k) as sim:     # and with the default launch plan
        sim.init_reference_field()
        sim.step(nsteps)
        assert np.array_equal(sim.download(), ref)
