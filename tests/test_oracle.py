"""CPU: pins the C restatement (oracle/heat3d_oracle.c) against the reference's own host program.

Golden fixtures under tests/golden/ were produced by oracle/_ref/heat3D_ref = the reference's
src/heat3D.cpp compiled here (tests/golden/make_golden.py).  Where that binary is present the
comparison is also re-run live.  Tolerance: bit-exact (strict arithmetic both sides).
"""
import hashlib
import os

import numpy as np
import pytest


def _case_field(oracle, args):
    nx, ny, nz, it, eps = args
    _, _, D = oracle.coefficients(nx, ny, nz)
    T = oracle.init_field(nx, ny, nz)
    applied, final_iters, conv = oracle.run(T, D, it, float(eps))
    return T, applied, final_iters, conv


@pytest.mark.parametrize("name", ["c27_converge", "aniso_33x17x65", "tiny_5x4x3", "c64_1000"])
def test_oracle_bitwise_vs_reference_golden(oracle, manifest, golden_field, name, tmp_path):
    m = manifest[name]
    T, applied, final_iters, conv = _case_field(oracle, m["args"])
    assert hashlib.sha256(T.tobytes()).hexdigest() == m["raw_sha256"]
    if m["raw_file"]:
        G = golden_field(name)
        assert np.array_equal(T, G)
    for key, hx in m["probes_hex"].items():
        k, j, i = map(int, key.split(","))
        assert float(T[k, j, i]).hex() == hx
    # report lines
    report = m["stdout"]
    if conv:
        assert report[-2].startswith(f"Simulation has converged in {final_iters} iterations")
    else:
        assert report[-2] == f"Simulation did not converge within {m['args'][3]} iterations."
    err = oracle.mean_abs_error_vs_y(T)
    assert report[-1] == "L2-norm error: %.4f %%" % (100 * err)
    # dump: byte-identical to the reference's iostream output
    p = tmp_path / "out.dat"
    oracle.write_dump(str(p), T)
    data = p.read_bytes()
    assert len(data) == m["dump_bytes"]
    assert hashlib.md5(data).hexdigest() == m["dump_md5"]


def test_known_answers_from_survey(manifest):
    # SURVEY.md Appendix C values, obtained independently during the survey.
    assert manifest["c27_converge"]["dump_md5"] == "5a4d37977671d00f9a6456d9a3174438"
    assert manifest["c27_converge"]["dump_bytes"] == 1200746     # same size as the shipped output/out.dat
    assert "converged in 3302 iterations" in manifest["c27_converge"]["stdout"][-2]
    assert manifest["c64_1000"]["dump_md5"] == "09d0e9b189b42d81359e2ad1d3848fc0"
    assert manifest["c64_1000"]["stdout"][-1] == "L2-norm error: 18.5525 %"


def test_live_reference_binary_matches_oracle(oracle, tmp_path):
    if not oracle.have_reference_binary():
        pytest.skip("oracle/_ref/heat3D_ref not built (needs /root/reference)")
    nx, ny, nz, it = 19, 23, 11, 57
    r = oracle.run_reference_binary(nx, ny, nz, it, "1e-30", str(tmp_path))
    _, _, D = oracle.coefficients(nx, ny, nz)
    T = oracle.steps(oracle.init_field(nx, ny, nz), D, it)
    assert np.array_equal(T, r["raw"])


def test_coefficients(oracle):
    for n in (27, 64, 512, 1024, 2048):
        sp, dt, D = oracle.coefficients(n, n, n)
        assert D[0] == D[1] == D[2] == 0.4 / 6
        assert (n - 1) * sp[1] == 1.0
    sp, dt, D = oracle.coefficients(1024, 1024, 4096)
    assert D[2] == 0.4 / 6 and D[0] == D[1] < D[2]


def test_init_field_shell(oracle):
    nx, ny, nz = 9, 7, 5
    sp, _, _ = oracle.coefficients(nx, ny, nz)
    T = oracle.init_field(nx, ny, nz)
    assert np.all(T[1:-1, 1:-1, 1:-1] == 0)
    assert np.all(T[:, 0, :] == 0)
    assert np.all(T[1:-1, -1, 1:-1] == 1.0)
    y = np.arange(ny) * sp[1]
    for face in (T[:, :, 0], T[:, :, -1]):
        assert np.array_equal(face, np.broadcast_to(y, face.shape))
    for face in (T[0], T[-1]):
        assert np.array_equal(face, np.broadcast_to(y[:, None], face.shape))


def test_steady_state_is_fixed_point_and_max_principle(oracle):
    nx, ny, nz = 17, 17, 17
    sp, _, D = oracle.coefficients(nx, ny, nz)
    T = oracle.init_field(nx, ny, nz)
    T[:, :, :] = (np.arange(ny) * sp[1])[None, :, None]
    T[:, -1, :] = 1.0
    T0 = T.copy()
    oracle.steps(T, D, 5)
    assert np.max(np.abs(T - T0)) < 1e-15
    U = oracle.steps(oracle.init_field(nx, ny, nz), D, 200)
    assert U.min() >= 0.0 and U.max() <= 1.0


def test_fma_mode_within_tolerance(oracle):
    nx = ny = nz = 33
    _, _, D = oracle.coefficients(nx, ny, nz)
    A = oracle.steps(oracle.init_field(nx, ny, nz), D, 500, arith=oracle.STRICT)
    B = oracle.steps(oracle.init_field(nx, ny, nz), D, 500, arith=oracle.FMA)
    assert 0 < np.max(np.abs(A - B)) < 1e-14          # differs in the last bits, far inside 1e-12


def test_threads_do_not_change_bits(oracle):
    nx, ny, nz = 40, 24, 31
    _, _, D = oracle.coefficients(nx, ny, nz)
    A = oracle.steps(oracle.init_field(nx, ny, nz), D, 50, threads=1)
    B = oracle.steps(oracle.init_field(nx, ny, nz), D, 50, threads=4)
    assert np.array_equal(A, B)


def test_slab_split_equals_single_domain(oracle):
    """Non-overlapping z-slabs with a 1-deep ghost exchange reproduce the single domain bitwise."""
    nx, ny, nz, P, nsteps = 21, 13, 30, 4, 25
    _, _, D = oracle.coefficients(nx, ny, nz)
    ref = oracle.steps(oracle.init_field(nx, ny, nz), D, nsteps)
    cur = oracle.init_field(nx, ny, nz)
    nxt = cur.copy()
    bounds = [(p * nz // P, (p + 1) * nz // P) for p in range(P)]
    for _ in range(nsteps):
        for z0, z1 in bounds:
            oracle.step_planes(cur, nxt, D, z0, z1)
        cur, nxt = nxt, cur
    assert np.array_equal(cur, ref)
