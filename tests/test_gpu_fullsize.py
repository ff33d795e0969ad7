"""Full-size parity (BASELINE.json configs[0] and configs[1]): the CUDA path on the meshes the bench runs, against the
oracle's threaded element loop entry by entry, and through the size-independent properties the reference's own tests use
(testsuite/feelvf/test_laplaciant.cpp:108-126, test_forms.cpp:135-139): K 1 = 0, 1^T M 1 = |Omega|, u^T K u and u^T M u
for a quadratic u, symmetry, `=` then `+=`, bit-identical repeats."""
import os

import numpy as np
import pytest

import feelpp_b200 as fpp
from feelpp_b200 import cst, elements, expr, grad, gradt, id, idt, integrate, trans
from oracle import oracle as orc
from oracle import parity

pytestmark = pytest.mark.gpu

THREADS = max(1, len(os.sched_getaffinity(0)))
ROWS_PER_CHUNK = 1 << 20


def check_chunked(vals, ref, scale, ia, ja, what):
    """oracle/parity.py's bound, evaluated in row chunks (308 M nonzeros do not go through numpy temporaries at once)"""
    n = len(ia) - 1
    diag = np.zeros(n)
    for r0 in range(0, n, ROWS_PER_CHUNK):
        r1 = min(n, r0 + ROWS_PER_CHUNK)
        k0, k1 = int(ia[r0]), int(ia[r1])
        rows = np.repeat(np.arange(r0, r1), np.diff(ia[r0:r1 + 1]))
        on = rows == ja[k0:k1]
        diag[rows[on]] = scale[k0:k1][on]
    assert (diag > 0).all()
    worst = 0.0
    for r0 in range(0, n, ROWS_PER_CHUNK):
        r1 = min(n, r0 + ROWS_PER_CHUNK)
        k0, k1 = int(ia[r0]), int(ia[r1])
        rows = np.repeat(np.arange(r0, r1), np.diff(ia[r0:r1 + 1]))
        tol = parity.RTOL * np.abs(ref[k0:k1]) + parity.FLOOR * np.sqrt(diag[rows] * diag[ja[k0:k1]])
        err = np.abs(vals[k0:k1] - ref[k0:k1])
        bad = err > tol
        assert not bad.any(), "%s: %d nonzeros out of tolerance in rows [%d,%d), worst err %.3e" % (
            what, int(bad.sum()), r0, r1, float(err[bad].max()))
        worst = max(worst, float((err / tol).max()))
    return worst


def oracle_pair(kind, Xh, ia, ja):
    m = Xh.mesh
    q = orc.quad_order_form2(kind, Xh.order)
    ref, _ = orc.form2(kind, Xh.order, 1, q, m.xyz, m.conn, Xh.l2g, ia, ja, threads=THREADS)
    scale, _ = orc.form2(kind | orc.ABS_FLAG, Xh.order, 1, q, m.xyz, m.conn, Xh.l2g, ia, ja, threads=THREADS)
    return ref, scale


def row_abs_sums(vals, ia):
    return np.add.reduceat(np.abs(vals), ia[:-1])


@pytest.mark.parametrize("order,m", [(1, 55), (2, 110)])
def test_full_size_against_oracle_and_invariants(engine, order, m):
    mesh = fpp.unitCube(m, 0.15, False)
    Xh = fpp.Pch(mesh, order)
    u = v = fpp.Element(Xh)
    a = fpp.form2(_test=Xh, _trial=Xh, engine=engine)
    ia, ja = a.ia, a.ja
    n = Xh.nDof()
    one = np.ones(n)
    f2 = Xh.interpolate(lambda x, y, z: x * x + y * z) if order >= 2 else Xh.interpolate(lambda x, y, z: x + 2 * y - z)

    # ---- stiffness: a = integrate(elements(mesh), gradt(u)*trans(grad(v))) ----
    stiff = integrate(_range=elements(mesh), _expr=gradt(u) * trans(grad(v)))
    a.assign(stiff)
    K = a.values()
    refK, scaleK = oracle_pair(orc.STIFFNESS, Xh, ia, ja)
    check_chunked(K, refK, scaleK, ia, ja, "stiffness P%d m=%d" % (order, m))
    rabs = row_abs_sums(K, ia)
    y = a.multVector(one)
    assert (np.abs(y) <= 1e-12 * rabs).all(), "K 1 != 0"
    # energy of an exactly represented function: |grad(x^2 + y z)|^2 = 4x^2 + z^2 + y^2 -> 2 ; |grad(x+2y-z)|^2 = 6
    exact = 2.0 if order >= 2 else 6.0
    assert abs(a.energy(f2, f2) - exact) <= 1e-11 * exact
    rng = np.random.default_rng(11)
    p, q = rng.standard_normal(n), rng.standard_normal(n)
    Kq = a.multVector(q)
    s1, s2 = float(p @ Kq), float(q @ a.multVector(p))
    assert abs(s1 - s2) <= 1e-11 * float(np.abs(p) @ np.abs(Kq)), "K is not symmetric"
    a.assign(stiff)
    assert np.array_equal(a.values(), K), "two assemblies of the same form differ bitwise"
    del K, y, Kq

    # ---- config 2's sequence: a = mass ; a += stiffness ----
    a.assign(integrate(_range=elements(mesh), _expr=idt(u) * id(v)))
    M = a.values()
    refM, scaleM = oracle_pair(orc.MASS, Xh, ia, ja)
    check_chunked(M, refM, scaleM, ia, ja, "mass P%d m=%d" % (order, m))
    M1 = a.multVector(one)
    assert abs(M1.sum() - 1.0) <= 1e-12, "1^T M 1 != |Omega|"
    if order >= 2:   # int (x^2 + y z)^2 = 1/5 + 1/6 + 1/9
        assert abs(a.energy(f2, f2) - (1 / 5 + 1 / 6 + 1 / 9)) <= 1e-11
    del M
    a += stiff
    refM += refK
    scaleM += scaleK
    check_chunked(a.values(), refM, scaleM, ia, ja, "mass + stiffness P%d m=%d" % (order, m))
    y = a.multVector(one)
    assert (np.abs(y - M1) <= 1e-12 * (rabs + np.abs(M1))).all(), "(M + K) 1 != M 1"
    del refM, scaleM, refK, scaleK

    # ---- form1: f = 1 and the non-polynomial f of config 2 ----
    l = fpp.form1(_test=Xh, engine=engine)
    l.assign(integrate(_range=elements(mesh), _expr=cst(1.0) * id(v)))
    b = l.values()
    assert abs(b.sum() - 1.0) <= 1e-12
    np.testing.assert_allclose(b, M1, rtol=1e-11, atol=1e-18)   # int phi_i = (M 1)_i
    f = expr("sin(pi*x)*cos(pi*y)*z")
    l.assign(integrate(_range=elements(mesh), _expr=f * id(v)))
    co = orc.Coeff(kind=orc.COEFF_EXPR, ops=f.prog.ops, consts=f.prog.consts)
    ref = orc.form1(order, 1, orc.quad_order_form1(order, -1), mesh.xyz, mesh.conn, Xh.l2g, n, co)
    np.testing.assert_allclose(l.values(), ref, rtol=1e-12, atol=1e-12 * float(np.abs(ref).max()))


# ---- configs[2] and configs[3] at full size (one GPU's share of config 4: m = 139 / 2) -------------------------------
def _fullsize_form(engine, order, ncomp, m, make_expr, kind, coeff, quad, what, invariant):
    """assemble `make_expr(Xh, mesh)` over the whole mesh with the default algorithm and compare every nonzero with the
    oracle's threaded element loop; then the form's size-independent property"""
    mesh = fpp.unitCube(m, 0.15, False)
    Xh = fpp.Pchv(mesh, order) if ncomp == 3 else fpp.Pch(mesh, order)
    a = fpp.form2(_test=Xh, _trial=Xh, engine=engine)
    ia, ja = a.ia, a.ja
    co = coeff(mesh)
    a.assign(integrate(_range=elements(mesh), _expr=make_expr(Xh, mesh)))
    vals = a.values()
    ref, _ = orc.form2(kind, order, ncomp, quad, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, co, threads=THREADS)
    scale, _ = orc.form2(kind | orc.ABS_FLAG, order, ncomp, quad, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, co, threads=THREADS)
    worst = check_chunked(vals, ref, scale, ia, ja, what)
    del ref, scale
    invariant(a, Xh, vals, ia)
    return worst


K_EXPR = "1+x*y+z^2"


def _k_field(mesh):
    Kh = fpp.Pch(mesh, 1)
    return Kh, Kh.interpolate(lambda x, y, z: 1.0 + 0.5 * np.sin(2 * np.pi * x))


def _rows_sum_to_zero(a, Xh, vals, ia):
    y = a.multVector(np.ones(Xh.nDof()))
    assert (np.abs(y) <= 1e-12 * row_abs_sums(vals, ia)).all(), "K 1 != 0"


@pytest.mark.parametrize("coeff", ["expr", "idv"])
def test_full_size_config3_p3_variable_conductivity(engine, coeff):
    """configs[2]: P3, k(x) inner(gradt(u), grad(v)) on m=87 (3 951 018 tets): k = expr("1+x*y+z^2") (order 6, 24 points)
    and k = idv(P1 field) (order 5, 14 points) -- heat-toolbox style (toolboxes/feel/feelmodels/heat/heatassembly.hpp:69-86)"""
    from feelpp_b200 import idv, inner
    if coeff == "expr":
        prog = fpp.GinacProgram(K_EXPR)
        mk = lambda Xh, mesh: expr(K_EXPR) * inner(gradt(fpp.Element(Xh)), grad(fpp.Element(Xh)))
        co = lambda mesh: orc.Coeff(kind=orc.COEFF_EXPR, ops=prog.ops, consts=prog.consts)
        quad = orc.quad_order_form2(orc.STIFFNESS, 3, 2)
    else:
        def mk(Xh, mesh):
            Kh, kv = _k_field(mesh)
            return idv(fpp.Element(Kh, kv)) * inner(gradt(fpp.Element(Xh)), grad(fpp.Element(Xh)))

        def co(mesh):
            Kh, kv = _k_field(mesh)
            return orc.Coeff(kind=orc.COEFF_FIELD, field_order=1, field_l2g=Kh.l2g, field_values=kv)
        quad = orc.quad_order_form2(orc.STIFFNESS, 3, 1)
    _fullsize_form(engine, 3, 1, 87, mk, orc.STIFFNESS, co, quad, "P3 k*stiffness (%s) m=87" % coeff, _rows_sum_to_zero)


def test_full_size_config4_p2_vector_elasticity(engine):
    """configs[3] per-GPU share: P2 vector elasticity, E=1e8, nu=0.25 (benchmarks/ATTIC/perf/assembly.cpp:86-89), m=70
    (2 058 000 tets, 8 x that on 8 GPUs = m=139); rigid translations are in the kernel of the operator"""
    from feelpp_b200 import Id, inner, sym, trace
    E, nu = 1e8, 0.25
    lam, mu = E * nu / ((1 + nu) * (1 - 2 * nu)), E / (2 * (1 + nu))

    def mk(Xh, mesh):
        u = v = fpp.Element(Xh)
        return inner(lam * trace(sym(gradt(u))) * Id() + 2 * mu * sym(gradt(u)), grad(v))

    def inv(a, Xh, vals, ia):
        rabs = row_abs_sums(vals, ia)
        for c in range(3):
            t = np.zeros(Xh.nDof())
            t[c::3] = 1.0
            assert (np.abs(a.multVector(t)) <= 1e-12 * rabs).all(), "a rigid translation is not in the kernel"

    _fullsize_form(engine, 2, 3, 70, mk, orc.ELASTICITY, lambda mesh: orc.Coeff(lam=lam, mu=mu),
                   orc.quad_order_form2(orc.ELASTICITY, 2), "P2v elasticity m=70", inv)
