"""tests/adapter_mock.cpp -- the Feel++-side adapter of INTEGRATION.md compiled with g++ against mock DofTable / GraphCSR /
Mesh objects: it builds and links against the C ABI on a machine without a GPU (and fails loudly there: no CPU fallback);
on a GPU its P1 Laplacian and load vector, flattened and assembled through the ABI from C++, equal the oracle's."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "feelpp_b200", "lib")


def build(tmp):
    exe = os.path.join(tmp, "adapter_mock")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "adapter_mock.cpp"),
           "-o", exe, "-L" + LIBDIR, "-lfeelgpu", "-Wl,-rpath," + LIBDIR, "-Wl,--allow-shlib-undefined"]
    out = subprocess.run(cmd, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    return exe


def test_adapter_mock_builds_links_and_fails_loudly_without_a_device(tmp_path):
    assert shutil.which("g++")
    exe = build(str(tmp_path))
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present: covered by the gpu test")
    out = subprocess.run([exe, "3", str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert out.returncode == 3 and "feelgpu_context_create" in out.stderr and "CUDA" in out.stderr


@pytest.mark.gpu
def test_adapter_mock_values_match_the_oracle(tmp_path):
    from oracle import oracle as orc
    exe = build(str(tmp_path))
    path = str(tmp_path / "o.bin")
    out = subprocess.run([exe, "5", path], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    raw = open(path, "rb").read()
    npts, nelt, nrows, nnz = np.frombuffer(raw, dtype=np.int64, count=4)
    o = 32
    def take(dt, cnt):
        nonlocal o
        a = np.frombuffer(raw, dtype=dt, count=cnt, offset=o)
        o += a.nbytes
        return a
    xyz, conn, l2g = take(np.float64, 3 * npts).reshape(-1, 3), take(np.int32, 4 * nelt).reshape(-1, 4), take(np.int32, 4 * nelt).reshape(-1, 4)
    ia, ja, a, b = take(np.int64, nrows + 1), take(np.int64, nnz), take(np.float64, nnz), take(np.float64, nrows)
    # the mock's DofTable and graph are the reference's first-touch numbering and COUPLED pattern
    l2g_o, nd = orc.doftable(npts, conn, 1)
    assert nd == nrows and np.array_equal(l2g_o, l2g)
    ia_o, ja_o = orc.graph(l2g, nd)
    assert np.array_equal(ia_o, ia) and np.array_equal(ja_o, ja)
    ref = orc.form2(orc.STIFFNESS, 1, 1, 1, xyz, conn, l2g, ia, ja)
    np.testing.assert_allclose(a, ref, rtol=1e-12, atol=1e-14 * np.abs(ref).max())
    refb = orc.form1(1, 1, 1, xyz, conn, l2g, nd, orc.Coeff(value=2.0))
    np.testing.assert_allclose(b, refb, rtol=1e-12, atol=1e-16)
