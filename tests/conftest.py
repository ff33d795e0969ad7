import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def engine():
    import feelpp_b200 as fpp
    eng = fpp.Engine(0)
    yield eng
    eng.close()


def scrambled(mesh, seed=3):
    """the same triangulation as gmsh would hand it over: vertex ids in random order and every tet's four vertices in
    random local order (both orientations occur; the reference takes J = |det K|, feelpoly/geomap.hpp:1971-1976)"""
    import numpy as np
    import feelpp_b200 as fpp
    rng = np.random.default_rng(seed)
    perm = rng.permutation(mesh.numPoints())          # new id of old vertex p
    xyz = np.empty_like(mesh.xyz)
    xyz[perm] = mesh.xyz
    conn = perm[mesh.conn].astype(np.int32)
    for e in range(conn.shape[0]):
        conn[e] = conn[e][rng.permutation(4)]
    return fpp.Mesh(xyz, conn)
