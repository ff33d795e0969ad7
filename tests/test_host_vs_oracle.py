"""Host-side mirror (feelpp_b200/host, C++) against the oracle's restatement of DofTable::build and
Stencil::computeGraph: local-to-global tables and CSR pattern must be bit-exact (BASELINE.json north_star)."""
import numpy as np
import pytest

import feelpp_b200 as fpp
from oracle import oracle as orc


@pytest.mark.parametrize("m,jit,shuf", [(1, 0, False), (2, 0, False), (4, 0.2, True), (6, 0.0, True)])
@pytest.mark.parametrize("order,ncomp", [(1, 1), (2, 1), (3, 1), (1, 3), (2, 3)])
def test_doftable_and_graph_bit_exact(m, jit, shuf, order, ncomp):
    mesh = fpp.unitCube(m, jit, shuf)
    Xh = fpp.FunctionSpace(mesh, order, ncomp)
    l2g, ndof = orc.doftable(mesh.numPoints(), mesh.conn, order, ncomp)
    assert ndof == Xh.nDof()
    assert np.array_equal(l2g, Xh.l2g)
    ia, ja = Xh.graph()
    oia, oja = orc.graph(l2g, ndof)
    assert np.array_equal(ia, oia) and np.array_equal(ja, oja)
    # rows ascending, diagonal present
    for r in (0, ndof // 2, ndof - 1):
        row = ja[ia[r]:ia[r + 1]]
        assert np.all(np.diff(row) > 0) and r in row


@pytest.mark.parametrize("order,ncomp", [(1, 1), (2, 1), (3, 1), (2, 3)])
def test_doftable_and_graph_bit_exact_scrambled_mesh(order, ncomp):
    """random vertex ids and random local vertex order: edge / face orientation by point ids (feelmesh/mesh3d.hpp:755-782)"""
    from conftest import scrambled
    mesh = scrambled(fpp.unitCube(3, 0.1, True))
    Xh = fpp.FunctionSpace(mesh, order, ncomp)
    l2g, ndof = orc.doftable(mesh.numPoints(), mesh.conn, order, ncomp)
    assert ndof == Xh.nDof() and np.array_equal(l2g, Xh.l2g)
    ia, ja = Xh.graph()
    oia, oja = orc.graph(l2g, ndof)
    assert np.array_equal(ia, oia) and np.array_equal(ja, oja)


def test_dof_counts_unit_cube():
    m = 4
    mesh = fpp.unitCube(m)
    nv = (m + 1) ** 3
    ne = 3 * m * (m + 1) ** 2 + 3 * m * m * (m + 1) + m ** 3      # axis + face-diagonal + body-diagonal edges
    nf = 2 * 6 * m ** 3 + 2 * 3 * m * m                            # (4*nElt + boundary faces)/2
    assert fpp.Pch(mesh, 1).nDof() == nv
    assert fpp.Pch(mesh, 2).nDof() == nv + ne
    assert fpp.Pch(mesh, 3).nDof() == nv + 2 * ne + nf


def test_mesh_is_positively_oriented_and_fills_cube():
    mesh = fpp.unitCube(3, 0.2, True)
    P = mesh.xyz[mesh.conn]
    vol = np.linalg.det(P[:, 1:] - P[:, :1]) / 6
    assert (vol > 0).all() and abs(vol.sum() - 1) < 1e-12


@pytest.mark.parametrize("order", [1, 2, 3])
def test_tabulation_matches_oracle(order):
    xi, w = fpp.im(6)
    phi, dphi = fpp.tabulate(order, xi)
    ophi, odphi = orc.tabulate(order, xi)
    assert np.abs(phi - ophi).max() < 1e-15 and np.abs(dphi - odphi).max() < 1e-14


def test_expr_parser():
    g = fpp.GinacProgram("1+x*y+z^2")
    assert g.polyDegree == 2
    assert fpp.GinacProgram("sin(pi*x)*cos(pi*y)*z").polyDegree == -1
    assert fpp.GinacProgram("2*(x+1)^3 - y/4").polyDegree == 3
    assert fpp.GinacProgram("3.5").polyDegree == 0
    with pytest.raises(RuntimeError):
        fpp.GinacProgram("1+*x")
    with pytest.raises(RuntimeError):
        fpp.GinacProgram("foo(x)")


def test_unsupported_expression_is_refused():
    mesh = fpp.unitCube(1)
    Xh = fpp.Pch(mesh, 1)
    u = v = fpp.Element(Xh)
    with pytest.raises(NotImplementedError):
        fpp.classify_form2(fpp.gradt(u) * fpp.id(v), 1)
