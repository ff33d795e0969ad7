"""Boundary-face integrals: integrate(_range=boundaryfaces(mesh), ...) for Neumann terms g*id(v)
and Robin terms k*idt(u)*id(v) (feelvf/integrator.hpp:3242-3704).  CPU: the oracle's face rule against the committed
triangle literals and against closed forms; the host mirror's tables against the oracle's.  GPU: parity with the oracle."""
import json
import os

import numpy as np
import pytest

import feelpp_b200 as fpp
from feelpp_b200 import hostmirror as hm
from feelpp_b200 import boundaryfaces, cst, expr, id, idt, integrate
from oracle import oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_triangle_rules_match_fixture_and_are_symmetric():
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "tri_quadrature.json")))
    for order in range(0, 9):
        q = np.array(g[str(order)]["q"]).reshape(-1, 3)
        assert abs(q[:, 2].sum() - 2.0) < 1e-13                     # area of the reference triangle
        # barycentric point set invariant under vertex permutations: the face permutation the reference applies
        # (pointsetquadrature.hpp:712-715) only reorders the points
        lam = np.stack([-(q[:, 0] + q[:, 1]) / 2, (1 + q[:, 0]) / 2, (1 + q[:, 1]) / 2], axis=1)
        key = lambda L, w: sorted((round(a, 12), round(b, 12), round(c, 12), round(x, 12)) for (a, b, c), x in zip(L, w))
        for perm in ((1, 0, 2), (2, 1, 0), (0, 2, 1), (1, 2, 0)):
            assert key(lam, q[:, 2]) == key(lam[:, perm], q[:, 2]), order
        # exactness: monomials up to the order
        for a in range(order + 1):
            for b in range(order + 1 - a):
                num = (q[:, 2] * lam[:, 1] ** a * lam[:, 2] ** b).sum()
                from math import factorial as F
                assert abs(num - 2.0 * 2 * F(a) * F(b) / F(a + b + 2)) < 1e-13, (order, a, b)


def test_face_rule_on_reference_tet():
    for order in (1, 2, 4, 6):
        xi, w = orc.face_quadrature(order)
        assert np.allclose(w.sum(1), [2 * np.sqrt(3.0), 2, 2, 2], rtol=1e-14)
        assert np.allclose(xi[0].sum(1), -1)                        # x + y + z = -1 on face 0
        for f in (1, 2, 3):
            assert np.allclose(xi[f][:, f - 1], -1)                  # xi_{f-1} = -1 on face f
        hx, hw = hm.im_faces(order)
        assert np.array_equal(hx.reshape(xi.shape), xi) and np.array_equal(hw.reshape(w.shape), w)


@pytest.mark.parametrize("order", [1, 2, 3])
def test_oracle_face_forms_closed_forms(order):
    mesh = fpp.unitCube(4, 0.15, True)                              # boundary vertices stay on the cube: area 6
    Xh = fpp.Pch(mesh, order)
    ia, ja = Xh.graph()
    fe, fl = orc.boundary_faces(mesh.conn, mesh.numPoints())
    assert len(fe) == 6 * 2 * 4 * 4
    he, hf = boundaryfaces(mesh).faces()
    assert np.array_equal(he, fe) and np.array_equal(hf, fl)
    b = orc.form1_faces(order, 1, order, mesh.xyz, mesh.conn, Xh.l2g, Xh.nDof(), fe, fl)
    assert abs(b.sum() - 6.0) < 1e-12
    a = orc.form2_faces(order, 1, 2 * order, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, fe, fl)
    assert abs(a.sum() - 6.0) < 1e-11
    # u = x + 2y - z interpolated: u^T R u = int_boundary u^2 = closed form on the cube
    import scipy.sparse as sps
    R = sps.csr_matrix((a, ja, ia), shape=(Xh.nDof(), Xh.nDof()))
    X = Xh.dofPoints()
    u = X[:, 0] + 2 * X[:, 1] - X[:, 2]
    f2 = lambda x, y, z: (x + 2 * y - z) ** 2
    from scipy.integrate import dblquad
    exact = 0.0
    for ax in range(3):
        for v in (0.0, 1.0):
            def g(s, t, ax=ax, v=v):
                p = [s, t]
                p.insert(ax, v)
                return f2(*p)
            exact += dblquad(g, 0, 1, 0, 1)[0]
    assert abs(u @ (R @ u) - exact) < 1e-10 * exact
    assert abs((R - R.T)).max() < 1e-14


@pytest.mark.gpu
@pytest.mark.parametrize("order,ncomp", [(1, 1), (2, 1), (3, 1), (2, 3)])
def test_gpu_face_forms_match_oracle(order, ncomp):
    eng = fpp.Engine(0)
    mesh = fpp.unitCube(5, 0.15, True)
    Xh = fpp.FunctionSpace(mesh, order, ncomp)
    ia, ja = Xh.graph()
    u = v = fpp.Element(Xh)
    fe, fl = orc.boundary_faces(mesh.conn, mesh.numPoints())
    rng = boundaryfaces(mesh)
    if ncomp == 1:
        # Neumann: constant and closed-form flux (non polynomial: order k + 2, feelvf/detail/ginacexvf.hpp:836-845)
        l = fpp.form1(Xh, eng)
        l.assign(integrate(_range=rng, _expr=cst(2.5) * id(v)))
        ref = orc.form1_faces(order, 1, order, mesh.xyz, mesh.conn, Xh.l2g, Xh.nDof(), fe, fl, orc.Coeff(value=2.5))
        assert np.abs(l.values() - ref).max() <= 1e-12 * np.abs(ref).max()
        g = expr("sin(pi*x)*cos(y)+z")
        l += integrate(_range=rng, _expr=g * id(v))
        ref = orc.form1_faces(order, 1, order + 2, mesh.xyz, mesh.conn, Xh.l2g, Xh.nDof(), fe, fl,
                              orc.Coeff(kind=orc.COEFF_EXPR, ops=g.prog.ops, consts=g.prog.consts), b=ref)
        assert np.abs(l.values() - ref).max() <= 1e-12 * np.abs(ref).max()
        # a volume term and a face term in one vector
        l.assign(integrate(_range=fpp.elements(mesh), _expr=cst(1.0) * id(v)))
        l += integrate(_range=rng, _expr=cst(1.0) * id(v))
        assert abs(l.values().sum() - 7.0) < 1e-11
    # Robin: k idt(u) id(v) on the boundary, on top of a stiffness matrix
    a = fpp.form2(Xh, Xh, eng)
    from feelpp_b200 import grad, gradt, inner, trans
    vol = integrate(_range=fpp.elements(mesh), _expr=inner(gradt(u), grad(v)) if ncomp > 1 else gradt(u) * trans(grad(v)))
    rob = integrate(_range=rng, _expr=cst(3.0) * (inner(idt(u), id(v)) if ncomp > 1 else idt(u) * id(v)))
    a.assign(vol)
    a += rob
    K = orc.form2(orc.STIFFNESS, order, ncomp, orc.quad_order_form2(orc.STIFFNESS, order), mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    ref = orc.form2_faces(order, ncomp, 2 * order, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, fe, fl, orc.Coeff(value=3.0), a=K)
    assert np.abs(a.values() - ref).max() <= 1e-12 * np.abs(ref).max()
    a.assign(rob)                                                    # '=' zeroes first
    ref = orc.form2_faces(order, ncomp, 2 * order, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, fe, fl, orc.Coeff(value=3.0))
    assert np.abs(a.values() - ref).max() <= 1e-12 * np.abs(ref).max()
    assert abs(a.values().sum() - 3.0 * 6.0 * ncomp) < 1e-10
    eng.close()
