"""on(): Dirichlet elimination (feelvf/on.hpp:395-449 -> MatrixPetsc::zeroRows, feelalg/matrixpetsc.cpp:1762-1860).
CPU: the oracle restatement against its defining properties; GPU: feelgpu_form2_zero_rows against the oracle."""
import numpy as np
import pytest
import scipy.sparse as sps
import scipy.sparse.linalg as spla

import feelpp_b200 as fpp
from feelpp_b200 import cst, elements, boundaryfaces, expr, grad, gradt, id, integrate, on, trans
from oracle import oracle as orc


def problem(order, m=4):
    mesh = fpp.unitCube(m, 0.15, True)
    Xh = fpp.Pch(mesh, order)
    ia, ja = Xh.graph()
    a = orc.form2(orc.STIFFNESS, order, 1, orc.quad_order_form2(orc.STIFFNESS, order), mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    b = orc.form1(order, 1, orc.quad_order_form1(order, 0), mesh.xyz, mesh.conn, Xh.l2g, Xh.nDof())
    return mesh, Xh, ia, ja, a, b


@pytest.mark.parametrize("order", [1, 2, 3])
def test_boundary_dofs_are_the_nodes_on_the_cube_surface(order):
    mesh = fpp.unitCube(3)
    Xh = fpp.Pch(mesh, order)
    p = Xh.dofPoints()
    onb = (np.abs(p) < 1e-12).any(1) | (np.abs(p - 1) < 1e-12).any(1)
    assert np.array_equal(Xh.boundaryDofs(), np.nonzero(onb)[0])


@pytest.mark.parametrize("strategy", [orc.ON_ELIMINATION, orc.ON_ELIMINATION | orc.ON_SYMMETRIC,
                                      orc.ON_ELIMINATION | orc.ON_SYMMETRIC | orc.ON_KEEP_DIAGONAL])
def test_oracle_elimination_solves_the_dirichlet_problem(strategy):
    """-Laplace u = 0 with u = g on the boundary, g in P1: the discrete solution is g at every dof, for every strategy"""
    mesh, Xh, ia, ja, a, b = problem(2)
    g = lambda x, y, z: 1 + 2 * x - y + 0.5 * z
    rows = Xh.boundaryDofs()
    p = Xh.dofPoints()
    x = g(p[rows, 0], p[rows, 1], p[rows, 2])
    a2, b2 = orc.zero_rows(ia, ja, a, np.zeros_like(b), rows, x, strategy, 1.0)
    A = sps.csr_matrix((a2, ja, ia))
    if strategy & orc.ON_SYMMETRIC:
        assert abs(A - A.T).max() < 1e-12
    u = spla.spsolve(A.tocsc(), b2)
    assert np.allclose(u, g(p[:, 0], p[:, 1], p[:, 2]), atol=1e-10)
    d = A.diagonal()[rows]
    if strategy & orc.ON_KEEP_DIAGONAL:
        assert np.allclose(d, sps.csr_matrix((a, ja, ia)).diagonal()[rows])
    else:
        assert np.allclose(d, 1.0)


@pytest.mark.gpu
@pytest.mark.parametrize("order", [1, 2, 3])
@pytest.mark.parametrize("typ,strategy", [("elimination", 1), ("elimination_symmetric", 5),
                                          ("elimination_keep_diagonal", 3), ("elimination_symmetric_keep_diagonal", 7)])
def test_gpu_on_matches_oracle(engine, order, typ, strategy):
    mesh, Xh, ia, ja, aref, bref = problem(order)
    u = v = fpp.Element(Xh)
    a = fpp.form2(_test=Xh, _trial=Xh, engine=engine)
    l = fpp.form1(_test=Xh, engine=engine)
    a.assign(integrate(_range=elements(mesh), _expr=gradt(u) * trans(grad(v))))
    l.assign(integrate(_range=elements(mesh), _expr=cst(1.0) * id(v)))
    a0, b0 = a.values(), l.values()
    g = expr("1+2*x-y*z")
    a += on(_range=boundaryfaces(mesh), _rhs=l, _element=u, _expr=g, _type=typ, _value_on_diagonal=2.0)
    rows = Xh.boundaryDofs()
    x = g.prog.evaluate(Xh.dofPoints()[rows])
    a2, b2 = orc.zero_rows(ia, ja, a0, b0, rows, x, strategy, 2.0)
    # the elimination only copies, zeroes and (symmetric) subtracts one product per entry: bit-exact up to the order of
    # the atomic subtractions into b
    assert np.array_equal(a.values(), a2)
    np.testing.assert_allclose(l.values(), b2, rtol=1e-13, atol=1e-14 * np.abs(b2).max())


def test_context_on_values_are_the_reference_enum():
    """ContextOn (feelalg/enums.hpp:56-66) and ContextOnMap (feelalg/enums.cpp:28-35): the adapter forwards the strategy of
    form.zeroRows(...) unchanged, so header, mirror and oracle must use the reference's numeric values"""
    import re, os
    from feelpp_b200 import vf
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(ROOT, "include", "feelgpu.h")).read()
    got = {k: int(v) for k, v in re.findall(r"FEELGPU_ON_(\w+) = (\d+)", hdr)}
    assert got == {"PENALISATION": 0, "ELIMINATION": 1, "KEEP_DIAGONAL": 2, "SYMMETRIC": 4}
    assert (vf.ON_ELIMINATION, vf.ON_KEEP_DIAGONAL, vf.ON_SYMMETRIC) == (1, 2, 4)
    assert (orc.ON_ELIMINATION, orc.ON_KEEP_DIAGONAL, orc.ON_SYMMETRIC) == (1, 2, 4)
    assert vf._ON_TYPES == {"elimination": 1, "elimination_keep_diagonal": 3, "elimination_symmetric": 5,
                            "elimination_symmetric_keep_diagonal": 7}


def test_oracle_zero_rows_needs_the_elimination_bit():
    """MatrixPetsc::zeroRows does nothing unless ContextOn::ELIMINATION is set (matrixpetsc.cpp:1784)"""
    mesh, Xh, ia, ja, a, b = problem(1)
    rows = Xh.boundaryDofs()
    a2, b2 = orc.zero_rows(ia, ja, a, b, rows, np.ones(len(rows)), orc.ON_SYMMETRIC | orc.ON_KEEP_DIAGONAL, 1.0)
    assert np.array_equal(a2, a) and np.array_equal(b2, b)


@pytest.mark.gpu
def test_gpu_on_duplicate_rows_and_penalisation(engine):
    """a row listed twice keeps its first value (on.hpp de-duplicates while collecting); a strategy without ELIMINATION is
    refused (penalisation stays on the host)"""
    import ctypes as C
    from feelpp_b200._capi import check, c_dblp, c_i32p, ptr, FeelGpuError
    mesh, Xh, ia, ja, aref, bref = problem(2)
    u = v = fpp.Element(Xh)
    a = fpp.form2(_test=Xh, _trial=Xh, engine=engine)
    l = fpp.form1(_test=Xh, engine=engine)
    a.assign(integrate(_range=elements(mesh), _expr=gradt(u) * trans(grad(v))))
    l.assign(integrate(_range=elements(mesh), _expr=cst(1.0) * id(v)))
    a0, b0 = a.values(), l.values()
    rows = Xh.boundaryDofs().astype(np.int32)
    x = np.linspace(1.0, 2.0, len(rows))
    rows2 = np.concatenate([rows, rows[::3]]).astype(np.int32)
    x2 = np.concatenate([x, 7.0 + x[::3]])
    check(engine.lib.feelgpu_form2_zero_rows(a.pat, len(rows2), ptr(rows2, c_i32p), ptr(x2, c_dblp), 5, 1.5, a.d_values,
                                             l.d_values, None), engine.ctx)
    engine.sync()
    a2, b2 = orc.zero_rows(ia, ja, a0, b0, rows, x, 5, 1.5)
    assert np.array_equal(a.values(), a2)
    np.testing.assert_allclose(l.values(), b2, rtol=1e-13, atol=1e-14 * np.abs(b2).max())
    for bad in (0, 4, 6, 8 | 1):
        with pytest.raises(FeelGpuError):
            check(engine.lib.feelgpu_form2_zero_rows(a.pat, len(rows), ptr(rows, c_i32p), ptr(x, c_dblp), bad, 1.0, a.d_values,
                                                     l.d_values, None), engine.ctx)
