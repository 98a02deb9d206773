"""GPU: black-box test of the heat3D executable against what the reference's own host program printed and
dumped (tests/golden/manifest.json): same report lines, byte-identical output/out.dat."""
import hashlib
import os
import socket
import subprocess

import numpy as np
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


def test_driver_periodic_dumps_and_perf_line(h3, oracle, manifest, tmp_path):
    """HEAT3D_OUT_EVERY: snapshots leave the GPU while stepping goes on and their files are written an interval later;
    every file must be the oracle's dump of the field at that iteration, byte for byte."""
    m = manifest["aniso_33x17x65"]
    r = _run(h3, m["args"], str(tmp_path), {"HEAT3D_OUT_EVERY": "100", "HEAT3D_PERF": "1", "HEAT3D_CHECK_EVERY": "50"})
    assert r.returncode == 0, r.stderr
    assert "[heat3D perf]" in r.stderr and "GLUPS" in r.stderr
    files = sorted(os.listdir(tmp_path / "output"))
    assert files == ["out.dat", "out_100.dat", "out_200.dat"]
    assert hashlib.md5(open(tmp_path / "output" / "out.dat", "rb").read()).hexdigest() == m["dump_md5"]
    nx, ny, nz = m["args"][:3]
    _, _, D = oracle.coefficients(nx, ny, nz)
    T = oracle.init_field(nx, ny, nz)
    for it in (100, 200):
        T = oracle.steps(T, D, 100)
        want = tmp_path / f"oracle_{it}.dat"
        oracle.write_dump(str(want), T)
        assert open(tmp_path / "output" / f"out_{it}.dat", "rb").read() == open(want, "rb").read(), it


def _read_raw(path):
    with open(path, "rb") as f:
        head = f.readline().decode().split()
        i = head.index("points")
        nx, ny, nz = (int(v) for v in head[i + 1:i + 4])
        j = head.index("stride")
        stride = tuple(int(v) for v in head[j + 1:j + 4])
        it = int(head[head.index("iteration") + 1])
        data = np.frombuffer(f.read(), dtype=np.float64)
    return data.reshape(nz, ny, nx), stride, it


@pytest.mark.parametrize("ndev", [1, 3])
def test_driver_raw_format_and_stride(h3, oracle, tmp_path, ndev):
    """HEAT3D_OUT_FORMAT=raw + HEAT3D_OUT_STRIDE: the decimated fp64 field, bit for bit, for the periodic snapshots
    (decimated on the GPU) and for the final dump; and the decimated Tecplot text when the format stays ascii."""
    nx, ny, nz, iters = 45, 31, 38, 70
    _, _, D = oracle.coefficients(nx, ny, nz)
    sx, sy, sz = 2, 3, 4
    env = {"HEAT3D_OUT_FORMAT": "raw", "HEAT3D_OUT_STRIDE": f"{sx},{sy},{sz}", "HEAT3D_OUT_EVERY": "25", "HEAT3D_DEVICES": ",".join(["0"] * ndev),
           "HEAT3D_CHECK_EVERY": "7"}
    r = _run(h3, [nx, ny, nz, iters, "1e-30"], str(tmp_path), env)
    assert r.returncode == 0, r.stderr
    assert sorted(os.listdir(tmp_path / "output")) == ["out.raw", "out_25.raw", "out_50.raw"]
    for name, it in (("out_25.raw", 25), ("out_50.raw", 50), ("out.raw", iters)):
        got, stride, hit = _read_raw(tmp_path / "output" / name)
        want = oracle.steps(oracle.init_field(nx, ny, nz), D, it)[::sz, ::sy, ::sx]
        assert stride == (sx, sy, sz) and hit == it
        assert np.array_equal(got, want), name
    # ascii + stride: same decimated values in the reference's line format, coordinates of the kept points
    wd = tmp_path / "ascii"
    wd.mkdir()
    r = _run(h3, [nx, ny, nz, iters, "1e-30"], str(wd), {"HEAT3D_OUT_STRIDE": "2"})
    assert r.returncode == 0, r.stderr
    lines = open(wd / "output" / "out.dat").read().splitlines()
    want = oracle.steps(oracle.init_field(nx, ny, nz), D, iters)[::2, ::2, ::2]
    assert lines[2] == 'ZONE T = "0", I=%d, J=%d, K=%d, F=POINT' % (want.shape[2], want.shape[1], want.shape[0])
    vals = np.array([float(ln.split()[3]) for ln in lines[3:]]).reshape(want.shape)
    assert len(lines) == 3 + want.size and np.allclose(vals, want, rtol=1e-5, atol=1e-300)
    sp = oracle.coefficients(nx, ny, nz)[0]
    assert lines[3 + 1].split()[0] == "%.5e" % (2 * sp[0])          # second point along x is grid point 2


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


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 4])
def test_driver_one_process_per_slab(h3, manifest, tmp_path, world):
    """The reference's launch model, `mpirun -n P heat3D ...` (heat3D.cu:203-263), without MPI: P processes of the
    executable, rank / world from the environment, the NCCL id through a file; rank 0 reports and joins the per-rank
    parts into one single-zone out.dat that must equal the single-domain dump byte for byte."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    for name, extra in (("c27_converge", {}), ("c64_1000", {"HEAT3D_TEMPORAL_K": "2", "HEAT3D_CHECK_EVERY": "10"})):
        m = manifest[name]
        wd = tmp_path / f"{name}_{world}"
        (wd / "output").mkdir(parents=True)
        idfile = str(wd / "nccl_id")
        procs = []
        for rank in range(world):
            e = dict(os.environ, HEAT3D_RANK=str(rank), HEAT3D_WORLD=str(world), HEAT3D_LOCAL_RANK=str(rank), HEAT3D_ID_FILE=idfile, **extra)
            for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
                e.pop(k, None)
            procs.append(subprocess.Popen([h3.DRIVER_PATH] + [str(a) for a in m["args"]], cwd=str(wd), env=e, stdout=subprocess.PIPE,
                                          stderr=subprocess.PIPE, text=True))
        outs = [p.communicate(timeout=600) for p in procs]
        assert all(p.returncode == 0 for p in procs), [o[1][-500:] for o in outs]
        assert _report(outs[0][0]) == m["stdout"]
        assert all(o[0] == "" for o in outs[1:])               # only rank 0 talks
        assert sorted(os.listdir(wd / "output")) == ["out.dat"]
        assert hashlib.md5(open(wd / "output" / "out.dat", "rb").read()).hexdigest() == m["dump_md5"]


def test_driver_2d_decomposition_same_output(h3, manifest, tmp_path):
    """HEAT3D_NSLABS_Y: rows cut into ranges on top of the z-slabs; report and dump stay the single-domain ones."""
    m = manifest["aniso_33x17x65"]
    r = _run(h3, m["args"], str(tmp_path), {"HEAT3D_NSLABS_Y": "2", "HEAT3D_DEVICES": "0,0,0,0"})
    assert r.returncode == 0, r.stderr
    assert _report(r.stdout) == m["stdout"]
    assert hashlib.md5(open(tmp_path / "output" / "out.dat", "rb").read()).hexdigest() == m["dump_md5"]
