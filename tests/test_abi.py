"""CPU: the C-ABI library builds, loads, and exports every symbol include/heat3d.h declares; no compute."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def h3():
    import __graft_entry__ as ge
    ge.build()
    return ge.load_package()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "heat3d.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(heat3d_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported(h3):
    names = _declared_symbols()
    assert len(names) >= 25
    L = ctypes.CDLL(h3.LIB_PATH)
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    # and the Python mirror binds exactly the header's set
    assert sorted(h3.EXPORTS) == names


def test_version_and_param_struct(h3):
    assert h3.lib().heat3d_version() == 100
    p = h3.Params()
    h3.lib().heat3d_default_params(ctypes.byref(p), 5, 6, 7)
    assert (p.nx, p.ny, p.nz) == (5, 6, 7)
    assert p.struct_size == ctypes.sizeof(h3.Params)      # header and mirror agree on the layout
    assert p.nslabs == 1 and p.nlocal == 1 and p.temporal_k == 1 and p.arith == h3.ARITH_STRICT


def test_invalid_arguments_are_rejected_before_any_cuda_call(h3):
    L = h3.lib()
    ctx = ctypes.c_void_p()
    p = h3.Params()
    L.heat3d_default_params(ctypes.byref(p), 2, 9, 9)
    assert L.heat3d_create(ctypes.byref(p), ctypes.byref(ctx)) == -1
    assert b">= 3 points" in L.heat3d_last_error()
    L.heat3d_default_params(ctypes.byref(p), 9, 9, 9)
    p.struct_size = 4
    assert L.heat3d_create(ctypes.byref(p), ctypes.byref(ctx)) == -1
    assert L.heat3d_destroy(None) == 0
    assert L.heat3d_step(None, 1) == -1


def test_no_cpu_fallback(h3):
    """Without a GPU the product must fail loudly, never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(h3.Heat3DError) as e:
        h3.Heat3D(16, 16, 16)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_product_does_not_reference_oracle():
    pkg = os.path.join(ROOT, "cuda-aware-mpi-on-3d-heate-quation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in text.lower(), os.path.join(dirpath, f)


def test_coefficients_match_reference_expressions(h3, oracle):
    for shape in [(27, 27, 27), (64, 64, 64), (33, 17, 65), (1024, 1024, 4096)]:
        sp, dt, D = h3.coefficients(*shape)
        osp, odt, oD = oracle.coefficients(*shape)
        assert list(osp) == sp and odt == dt and list(oD) == D


def test_slab_bounds(h3):
    for nz, P in [(1024, 8), (30, 4), (7, 3), (65, 2)]:
        b = h3.slab_bounds(nz, P)
        assert b[0][0] == 0 and b[-1][1] == nz
        assert all(b[i][1] == b[i + 1][0] for i in range(P - 1))
        sizes = [z1 - z0 for z0, z1 in b]
        assert max(sizes) - min(sizes) <= 1
