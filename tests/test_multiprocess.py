"""The N>1 path, one process per slab under torch.distributed.run: gloo on CPU (host-side logic), NCCL on GPUs."""
import os
import socket
import subprocess
import sys

import pytest

from conftest import ROOT


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _launch(world, backend, shape_steps, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "mp_worker.py"), backend] + [str(v) for v in shape_steps]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    for k in range(world):
        assert f"rank {k}/{world} {backend} ok" in r.stdout


@pytest.mark.parametrize("world,shape_steps", [(2, (21, 13, 30, 12)), (3, (16, 9, 11, 7)), (3, (16, 9, 17, 11))])
def test_gloo_slabs_equal_single_domain(world, shape_steps):
    _launch(world, "gloo", shape_steps)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4])
def test_nccl_slabs_equal_single_domain_bitwise(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    _launch(world, "nccl", (150, 70, 46, 15))
