"""Pins the oracle with the analytic identities the reference's own tests use (no golden matrices exist upstream):
testsuite/feelvf/test_forms.cpp:135-139 (a(1,1) == measure), testsuite/feeldiscr/test_wire_basket.cpp:141-142
(mass energy), testsuite/feelvf/test_laplaciant.cpp:108-126 (energy of an interpolant), testsuite/feelpoly/
testlagrange.hpp:100-137 (phi_i(x_j) = delta_ij within 3000 eps), testsuite/feelintegration/
test_auto_order_integrate.cpp:169-211 (quadrature-order rule), and the extracted imsimplex.cpp literals."""
import json
import os

import numpy as np
import pytest
import scipy.sparse as sps

import feelpp_b200 as fpp
from oracle import oracle as orc

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_quadrature_tables_match_reference_literals():
    js = json.load(open(os.path.join(GOLD, "tet_quadrature.json")))
    for order, rule in js.items():
        xi, w = orc.quadrature(int(order))
        q = np.array(rule["q"]).reshape(-1, 4)
        assert len(w) == rule["npts"]
        assert np.array_equal(xi, q[:, :3]) and np.array_equal(w, q[:, 3])
        hxi, hw = fpp.im(int(order))          # the product's copy of the same tables
        assert np.array_equal(hxi, xi) and np.array_equal(hw, w)
    # point counts of SURVEY.md section 8: orders 1,2,4,5,6 -> 1,4,11,14,24
    assert [js[str(o)]["npts"] for o in (1, 2, 4, 5, 6)] == [1, 4, 11, 14, 24]


@pytest.mark.parametrize("order", range(1, 10))
def test_quadrature_exactness(order):
    """integrates monomials of total degree <= order exactly on the reference tet (test_quad_order.cpp:132-385)"""
    from math import factorial
    xi, w = orc.quadrature(order)
    lam = np.stack([(1 + xi[:, 0]) / 2, (1 + xi[:, 1]) / 2, (1 + xi[:, 2]) / 2], 1)
    for a in range(order + 1):
        for b in range(order + 1 - a):
            for c in range(order + 1 - a - b):
                exact = 8 * factorial(a) * factorial(b) * factorial(c) / factorial(a + b + c + 3)
                got = np.sum(w * lam[:, 0] ** a * lam[:, 1] ** b * lam[:, 2] ** c)
                assert abs(got - exact) < 2e-13, (order, a, b, c)


@pytest.mark.parametrize("order", [1, 2, 3])
def test_lagrange_kronecker_and_partition_of_unity(order):
    pts = orc.lagrange_points(order)
    phi, dphi = orc.tabulate(order, pts)
    assert np.abs(phi - np.eye(len(pts))).max() < 3000 * np.finfo(float).eps
    xi, _ = orc.quadrature(6)
    phi, dphi = orc.tabulate(order, xi)
    assert np.abs(phi.sum(0) - 1).max() < 1e-14 and np.abs(dphi.sum(0)).max() < 1e-13
    # derivative consistency: central finite differences
    h = 1e-6
    for c in range(3):
        e = np.zeros(3); e[c] = h
        fd = (orc.tabulate(order, xi + e)[0] - orc.tabulate(order, xi - e)[0]) / (2 * h)
        assert np.abs(fd - dphi[:, :, c]).max() < 1e-8


def test_p3_edge_nodes_are_gauss_lobatto():
    pts = orc.lagrange_points(3)
    # edge 2 = (v0,v1) runs along xi_0 from -1 to 1: interior nodes at -+1/sqrt(5) (warpblend.hpp:96-121)
    assert np.allclose(pts[8], [-1 / np.sqrt(5), -1, -1]) and np.allclose(pts[9], [1 / np.sqrt(5), -1, -1])
    assert np.allclose(pts[16], [1 / 3 - 2 / 3 + 0, 1 / 3 - 2 / 3, 1 / 3 - 2 / 3])  # face 0 centroid (v1,v2,v3)


def test_quadrature_order_rule():
    # (kind, order, coeff degree) -> order ; SURVEY.md section 8 table
    assert orc.quad_order_form2(orc.STIFFNESS, 1) == 1
    assert orc.quad_order_form2(orc.STIFFNESS, 2) == 2
    assert orc.quad_order_form2(orc.MASS, 2) == 4
    assert orc.quad_order_form2(orc.STIFFNESS, 3) == 4
    assert orc.quad_order_form2(orc.STIFFNESS, 3, 1) == 5
    assert orc.quad_order_form2(orc.STIFFNESS, 3, 2) == 6
    assert orc.quad_order_form2(orc.STIFFNESS, 3, -1) == 6
    assert orc.quad_order_form2(orc.ELASTICITY, 2) == 2
    assert orc.quad_order_form1(2, 0) == 2 and orc.quad_order_form1(2, -1) == 4
    # the harness' DSEL picks the same orders
    mesh = fpp.unitCube(1)
    for k in (1, 2, 3):
        Xh = fpp.Pch(mesh, k)
        u = v = fpp.Element(Xh)
        assert fpp.classify_form2(fpp.gradt(u) * fpp.trans(fpp.grad(v)), k).quad_order == orc.quad_order_form2(orc.STIFFNESS, k)
        assert fpp.classify_form2(fpp.idt(u) * fpp.id(v), k).quad_order == orc.quad_order_form2(orc.MASS, k)
        e = fpp.expr("1+x*y+z^2") * fpp.inner(fpp.gradt(u), fpp.grad(v))
        assert fpp.classify_form2(e, k).quad_order == orc.quad_order_form2(orc.STIFFNESS, k, 2)
        e = fpp.expr("sin(pi*x)*cos(pi*y)*z") * fpp.id(v)
        assert fpp.classify_form1(e, k).quad_order == orc.quad_order_form1(k, -1)


@pytest.fixture(scope="module")
def mesh():
    return fpp.unitCube(3, 0.15, True)


@pytest.mark.parametrize("order", [1, 2, 3])
def test_mass_and_stiffness_identities(mesh, order):
    Xh = fpp.Pch(mesh, order)
    ia, ja = Xh.graph()
    M = orc.form2(orc.MASS, order, 1, orc.quad_order_form2(orc.MASS, order), mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    K = orc.form2(orc.STIFFNESS, order, 1, orc.quad_order_form2(orc.STIFFNESS, order), mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    M, K = sps.csr_matrix((M, ja, ia)), sps.csr_matrix((K, ja, ia))
    one = np.ones(Xh.nDof())
    assert abs(one @ M @ one - 1.0) < 1e-10                 # a(1,1) == measure (test_forms.cpp:135-139)
    assert np.abs(K @ one).max() < 1e-12 * np.abs(K).max()    # constants are in the kernel
    assert abs(K - K.T).max() == 0 or abs(K - K.T).max() < 1e-15 * np.abs(K).max()
    f = {1: lambda x, y, z: x + 2 * y - z, 2: lambda x, y, z: x * x + 2 * y * z, 3: lambda x, y, z: x ** 3 + 2 * y * z}[order]
    exact = {1: 6.0, 2: 4.0, 3: 9 / 5 + 8 / 3}[order]
    u = Xh.interpolate(f)
    assert abs(u @ K @ u - exact) < 1e-9                      # energy of an exactly represented polynomial


def test_elasticity_rigid_body_modes(mesh):
    Xh = fpp.Pchv(mesh, 2)
    ia, ja = Xh.graph()
    lam = mu = 4e7
    E = orc.form2(orc.ELASTICITY, 2, 3, 2, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, orc.Coeff(lam=lam, mu=mu))
    Es = sps.csr_matrix((E, ja, ia))
    p = Xh.dofPoints()
    rot = np.zeros(Xh.nDof()); rot[0::3] = -p[:, 1]; rot[1::3] = p[:, 0]
    assert np.abs(Es @ rot).max() < 1e-12 * np.abs(E).max()
    stretch = np.zeros(Xh.nDof()); stretch[0::3] = p[:, 0]
    assert abs(stretch @ Es @ stretch - (lam + 2 * mu)) < 1e-9 * (lam + 2 * mu)


def test_local_matrix_p1_closed_form():
    """P1 stiffness of the reference tet scaled to unit edge: A_ij = V grad(l_i).grad(l_j)"""
    G = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1.0]])
    A = orc.local_matrix(orc.STIFFNESS, 1, 1, 1, G)
    g = np.array([[-1, -1, -1], [1, 0, 0], [0, 1, 0], [0, 0, 1.0]])
    assert np.abs(A - (g @ g.T) / 6).max() < 1e-15


def test_threaded_baseline_matches_serial(mesh):
    Xh = fpp.Pch(mesh, 2)
    ia, ja = Xh.graph()
    ref = orc.form2(orc.STIFFNESS, 2, 1, 2, mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    par, secs = orc.form2(orc.STIFFNESS, 2, 1, 2, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, threads=3)
    assert secs >= 0 and np.abs(par - ref).max() < 1e-13 * np.abs(ref).max()


def test_threaded_loop_matches_serial_loop():
    """the threaded element loop (CPU baseline, full-size parity) gives the serial loop's matrix up to summation order,
    and the same cancellation scale in ABS mode"""
    import feelpp_b200 as fpp
    m = fpp.unitCube(4, 0.15, False)
    Xh = fpp.Pch(m, 2)
    ia, ja = Xh.graph()
    for kind in (orc.STIFFNESS, orc.MASS):
        q = orc.quad_order_form2(kind, 2)
        s = orc.form2(kind, 2, 1, q, m.xyz, m.conn, Xh.l2g, ia, ja)
        t, _ = orc.form2(kind, 2, 1, q, m.xyz, m.conn, Xh.l2g, ia, ja, threads=3)
        sa = orc.form2(kind | orc.ABS_FLAG, 2, 1, q, m.xyz, m.conn, Xh.l2g, ia, ja)
        ta, _ = orc.form2(kind | orc.ABS_FLAG, 2, 1, q, m.xyz, m.conn, Xh.l2g, ia, ja, threads=3)
        assert np.abs(s - t).max() <= 1e-15 * np.abs(sa).max()
        np.testing.assert_allclose(ta, sa, rtol=1e-14)
        assert (ta >= np.abs(t) - 1e-18).all()


@pytest.mark.parametrize("order", [1, 2, 3])
def test_local_matrices_match_exact_barycentric_integration(order):
    """Known-answer check that shares nothing with the oracle's evaluation path but the nodal points: the nodal basis is
    expanded in barycentric monomials (Vandermonde solve) and mass / stiffness are integrated exactly with
    int_T lambda^alpha = 6|T| alpha! / (|alpha|+3)!  -- no quadrature table, no Dubiner basis, no geomap."""
    from itertools import product
    from math import factorial
    pts = orc.lagrange_points(order)
    n = len(pts)
    lam = np.empty((n, 4))
    lam[:, 1:] = (pts + 1) / 2
    lam[:, 0] = 1 - lam[:, 1:].sum(1)
    alphas = [a for a in product(range(order + 1), repeat=4) if sum(a) == order]
    assert len(alphas) == n
    V = np.array([[np.prod(lam[i] ** np.array(a)) for a in alphas] for i in range(n)])
    Cf = np.linalg.inv(V)                         # phi_i = sum_a Cf[a, i] lambda^a
    rng = np.random.default_rng(5)
    G = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1.0]]) + 0.2 * rng.standard_normal((4, 3))
    T = np.hstack([np.ones((4, 1)), G])
    vol = abs(np.linalg.det(T)) / 6
    dlam = np.linalg.inv(T)[1:, :].T              # dlam[m] = grad lambda_m

    def integral(a):
        if min(a) < 0:
            return 0.0
        return 6 * vol * np.prod([factorial(x) for x in a]) / factorial(sum(a) + 3)

    na = len(alphas)
    Ma = np.array([[integral(tuple(np.add(a, b))) for b in alphas] for a in alphas])
    Ka = np.zeros((na, na))
    gg = dlam @ dlam.T
    for ia_, a in enumerate(alphas):
        for ib, b in enumerate(alphas):
            s = 0.0
            for m in range(4):
                if not a[m]:
                    continue
                for l in range(4):
                    if not b[l]:
                        continue
                    e = list(np.add(a, b)); e[m] -= 1; e[l] -= 1
                    s += a[m] * b[l] * gg[m, l] * integral(tuple(e))
            Ka[ia_, ib] = s
    M_exact, K_exact = Cf.T @ Ma @ Cf, Cf.T @ Ka @ Cf
    M = orc.local_matrix(orc.MASS, order, 1, orc.quad_order_form2(orc.MASS, order), G).reshape(n, n)
    K = orc.local_matrix(orc.STIFFNESS, order, 1, orc.quad_order_form2(orc.STIFFNESS, order), G).reshape(n, n)
    assert np.abs(M - M_exact).max() < 1e-12 * np.abs(M_exact).max()
    assert np.abs(K - K_exact).max() < 1e-12 * np.abs(K_exact).max()


def test_local_elasticity_matches_exact_barycentric_integration():
    """P2 vector elasticity, local index I = n*c + i (doftable.hpp:1289-1292):
    E[(i,c),(j,d)] = lam D[c,d,i,j] + mu (delta_cd tr D[.,.,i,j] + D[d,c,i,j]), D[c,d,i,j] = int d_c phi_i d_d phi_j, integrated
    exactly in barycentric monomials."""
    from itertools import product
    from math import factorial
    order, n = 2, 10
    pts = orc.lagrange_points(order)
    lam4 = np.empty((n, 4))
    lam4[:, 1:] = (pts + 1) / 2
    lam4[:, 0] = 1 - lam4[:, 1:].sum(1)
    alphas = [a for a in product(range(order + 1), repeat=4) if sum(a) == order]
    Cf = np.linalg.inv(np.array([[np.prod(lam4[i] ** np.array(a)) for a in alphas] for i in range(n)]))
    rng = np.random.default_rng(11)
    G = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1.0]]) + 0.2 * rng.standard_normal((4, 3))
    T = np.hstack([np.ones((4, 1)), G])
    vol = abs(np.linalg.det(T)) / 6
    dlam = np.linalg.inv(T)[1:, :].T              # [m][c]

    def integral(a):
        return 0.0 if min(a) < 0 else 6 * vol * np.prod([factorial(x) for x in a]) / factorial(sum(a) + 3)

    Da = np.zeros((3, 3, n, n))                  # in the monomial basis
    for ia_, a in enumerate(alphas):
        for ib, b in enumerate(alphas):
            for m in range(4):
                for l in range(4):
                    if a[m] and b[l]:
                        e = list(np.add(a, b)); e[m] -= 1; e[l] -= 1
                        Da[:, :, ia_, ib] += a[m] * b[l] * integral(tuple(e)) * np.outer(dlam[m], dlam[l])
    D = np.einsum("ai,cdab,bj->cdij", Cf, Da, Cf)
    lam, mu = 4e7, 3e7
    E_exact = np.zeros((3 * n, 3 * n))
    for c in range(3):
        for d in range(3):
            blk = lam * D[c, d] + mu * D[d, c]
            if c == d:
                blk = blk + mu * (D[0, 0] + D[1, 1] + D[2, 2])
            E_exact[n * c:n * c + n, n * d:n * d + n] = blk
    E = orc.local_matrix(orc.ELASTICITY, 2, 3, 2, G, orc.Coeff(lam=lam, mu=mu)).reshape(3 * n, 3 * n)
    assert np.abs(E - E_exact).max() < 1e-12 * np.abs(E_exact).max()


@pytest.mark.parametrize("order", [1, 2, 3])
def test_load_vector_and_field_weighted_stiffness_match_exact_integration(order):
    """form1 f*id(v) with f = 1 (b_i = int phi_i) and form2 idv(k)*inner(gradt,grad) with k a P1 field
    (K_ij = sum_p k_p int lambda_p grad phi_i . grad phi_j), both exactly in barycentric monomials; the oracle uses the
    quadrature orders the reference selects (expr.hpp:1075-1076)."""
    from itertools import product
    from math import factorial
    pts = orc.lagrange_points(order)
    n = len(pts)
    lam4 = np.empty((n, 4))
    lam4[:, 1:] = (pts + 1) / 2
    lam4[:, 0] = 1 - lam4[:, 1:].sum(1)
    alphas = [a for a in product(range(order + 1), repeat=4) if sum(a) == order]
    Cf = np.linalg.inv(np.array([[np.prod(lam4[i] ** np.array(a)) for a in alphas] for i in range(n)]))
    rng = np.random.default_rng(order)
    G = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1.0]]) + 0.2 * rng.standard_normal((4, 3))
    T = np.hstack([np.ones((4, 1)), G])
    vol = abs(np.linalg.det(T)) / 6
    gg = np.linalg.inv(T)[1:, :].T
    gg = gg @ gg.T

    def integral(a):
        return 0.0 if min(a) < 0 else 6 * vol * np.prod([factorial(x) for x in a]) / factorial(sum(a) + 3)

    b_exact = Cf.T @ np.array([integral(a) for a in alphas])
    conn = np.array([[0, 1, 2, 3]], dtype=np.int32)
    l2g = np.arange(n, dtype=np.int32).reshape(1, n)
    b = orc.form1(order, 1, orc.quad_order_form1(order, 0), G, conn, l2g, n)
    assert np.abs(b - b_exact).max() < 1e-12 * np.abs(b_exact).max()

    kv = 1 + rng.random(4)
    Ka = np.zeros((n, n))
    for ia_, a in enumerate(alphas):
        for ib, bb in enumerate(alphas):
            for m in range(4):
                for l in range(4):
                    if a[m] and bb[l]:
                        for p in range(4):
                            e = list(np.add(a, bb)); e[m] -= 1; e[l] -= 1; e[p] += 1
                            Ka[ia_, ib] += kv[p] * a[m] * bb[l] * gg[m, l] * integral(tuple(e))
    K_exact = Cf.T @ Ka @ Cf
    co = orc.Coeff(kind=orc.COEFF_FIELD, field_order=1, field_l2g=conn, field_values=kv)
    K = orc.local_matrix(orc.STIFFNESS, order, 1, orc.quad_order_form2(orc.STIFFNESS, order, 1), G, co).reshape(n, n)
    assert np.abs(K - K_exact).max() < 1e-12 * np.abs(K_exact).max()


@pytest.mark.parametrize("order", [2, 3])
def test_polynomial_coefficient_stiffness_matches_exact_integration(order):
    """config 3's form: k*inner(gradt(u),grad(v)) with k = 1+x*y+z^2 given as expr("...") (degree 2 -> quadrature order
    2(k-1)+2, Q = 24 for P3).  k is rewritten in barycentric coordinates (x = sum_m lambda_m X_m) and the integral taken
    exactly."""
    from itertools import product
    from math import factorial
    pts = orc.lagrange_points(order)
    n = len(pts)
    lam4 = np.empty((n, 4))
    lam4[:, 1:] = (pts + 1) / 2
    lam4[:, 0] = 1 - lam4[:, 1:].sum(1)
    alphas = [a for a in product(range(order + 1), repeat=4) if sum(a) == order]
    Cf = np.linalg.inv(np.array([[np.prod(lam4[i] ** np.array(a)) for a in alphas] for i in range(n)]))
    rng = np.random.default_rng(40 + order)
    G = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1.0]]) + 0.2 * rng.standard_normal((4, 3))
    T = np.hstack([np.ones((4, 1)), G])
    vol = abs(np.linalg.det(T)) / 6
    gg = np.linalg.inv(T)[1:, :].T
    gg = gg @ gg.T

    def integral(a):
        return 0.0 if min(a) < 0 else 6 * vol * np.prod([factorial(x) for x in a]) / factorial(sum(a) + 3)

    kterms = [((0, 0, 0, 0), 1.0)]                                  # k = 1 + x*y + z^2 as a polynomial in lambda
    for m in range(4):
        for l in range(4):
            e = [0, 0, 0, 0]; e[m] += 1; e[l] += 1
            kterms.append((tuple(e), G[m, 0] * G[l, 1] + G[m, 2] * G[l, 2]))
    Ka = np.zeros((n, n))
    for ia_, a in enumerate(alphas):
        for ib, b in enumerate(alphas):
            for m in range(4):
                for l in range(4):
                    if a[m] and b[l]:
                        e = np.add(a, b); e[m] -= 1; e[l] -= 1
                        Ka[ia_, ib] += a[m] * b[l] * gg[m, l] * sum(c * integral(tuple(e + np.array(kb))) for kb, c in kterms)
    K_exact = Cf.T @ Ka @ Cf
    g = fpp.expr("1+x*y+z^2")
    assert orc.quad_order_form2(orc.STIFFNESS, order, 2) == 2 * (order - 1) + 2
    co = orc.Coeff(kind=orc.COEFF_EXPR, ops=g.prog.ops, consts=g.prog.consts)
    K = orc.local_matrix(orc.STIFFNESS, order, 1, orc.quad_order_form2(orc.STIFFNESS, order, 2), G, co).reshape(n, n)
    assert np.abs(K - K_exact).max() < 1e-12 * np.abs(K_exact).max()
