"""GPU: black-box test of the heat3D executable against what the reference's own host program printed and
dumped (tests/golden/manifest.json): same report lines, byte-identical output/out.dat."""
import hashlib
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def h3():
    import __graft_entry__ as ge
    return ge.load_package()


def _run(h3, args, cwd, env=None):
    os.makedirs(os.path.join(cwd, "output"), exist_ok=True)
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([h3.DRIVER_PATH] + [str(a) for a in args], cwd=cwd, env=e, capture_output=True, text=True)
    return r


def _report(stdout):
    return [ln for ln in stdout.splitlines() if not ln.startswith("executable:") and not ln.startswith("Computational time")]


@pytest.mark.parametrize("name", ["c27_converge", "aniso_33x17x65", "tiny_5x4x3", "c64_1000"])
@pytest.mark.parametrize("ndev_env", [{}, {"HEAT3D_KERNEL": "baseline"}])
def test_driver_matches_reference_program(h3, manifest, tmp_path, name, ndev_env):
    m = manifest[name]
    r = _run(h3, m["args"], str(tmp_path), ndev_env)
    assert r.returncode == 0, r.stderr
    assert _report(r.stdout) == m["stdout"]
    assert any(ln.startswith("Computational time (parallel): ") for ln in r.stdout.splitlines())
    data = open(tmp_path / "output" / "out.dat", "rb").read()
    assert len(data) == m["dump_bytes"]
    assert hashlib.md5(data).hexdigest() == m["dump_md5"]


def test_driver_usage_aborts_like_reference(h3, tmp_path):
    r = _run(h3, [27, 27, 27], str(tmp_path))
    assert r.returncode != 0          # std::abort(), as heat3D.cu:291
    assert "Incorrect number of command line arguments specified" in r.stdout
    assert "bin/HeatEquation3D NUM_CELLS_X NUM_CELLS_Y NUM_CELLS_Z ITER_MAX EPS" in r.stdout


def test_driver_periodic_dumps_and_perf_line(h3, manifest, tmp_path):
    m = manifest["aniso_33x17x65"]
    r = _run(h3, m["args"], str(tmp_path), {"HEAT3D_OUT_EVERY": "100", "HEAT3D_PERF": "1", "HEAT3D_CHECK_EVERY": "50"})
    assert r.returncode == 0, r.stderr
    assert "[heat3D perf]" in r.stderr and "GLUPS" in r.stderr
    files = sorted(os.listdir(tmp_path / "output"))
    assert files == ["out.dat", "out_100.dat", "out_200.dat"]
    assert hashlib.md5(open(tmp_path / "output" / "out.dat", "rb").read()).hexdigest() == m["dump_md5"]
    assert os.path.getsize(tmp_path / "output" / "out_100.dat") == m["dump_bytes"]


def test_driver_temporal_blocking_same_output(h3, manifest, tmp_path):
    """HEAT3D_TEMPORAL_K=2 + residual checked every 4th iteration: different kernels, byte-identical dump."""
    m = manifest["aniso_33x17x65"]
    r = _run(h3, m["args"], str(tmp_path), {"HEAT3D_TEMPORAL_K": "2", "HEAT3D_CHECK_EVERY": "4"})
    assert r.returncode == 0, r.stderr
    assert _report(r.stdout) == m["stdout"]
    assert hashlib.md5(open(tmp_path / "output" / "out.dat", "rb").read()).hexdigest() == m["dump_md5"]


@pytest.mark.parametrize("ndev", [2, 4])
@pytest.mark.parametrize("k", [1, 2])
def test_driver_multi_gpu_in_one_process(h3, manifest, tmp_path, ndev, k):
    """HEAT3D_NDEV=P: one process drives P GPUs (peer copies between slabs) - the reference's `mpirun -n P`;
    the dump is the single-zone form and must equal the single-domain result byte for byte."""
    import torch
    if torch.cuda.device_count() < ndev:
        pytest.skip(f"needs {ndev} GPUs")
    for name in ("c27_converge", "c64_1000"):
        m = manifest[name]
        wd = tmp_path / f"{name}_{ndev}_{k}"
        wd.mkdir()
        r = _run(h3, m["args"], str(wd), {"HEAT3D_NDEV": str(ndev), "HEAT3D_TEMPORAL_K": str(k)})
        assert r.returncode == 0, r.stderr
        assert _report(r.stdout) == m["stdout"]
        assert hashlib.md5(open(wd / "output" / "out.dat", "rb").read()).hexdigest() == m["dump_md5"]
