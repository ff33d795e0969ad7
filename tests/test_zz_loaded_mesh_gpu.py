"""GPU parity on a mesh that came through the .msh reader, with marked face ranges (see test_gmsh.py for the reader's own
tests).  Kept in a file that sorts last: it was written after round 1's GPU minutes were spent, so under `-x` an
unexpected failure here cannot hide the parity tests that have already run green on a B200."""
import numpy as np
import pytest

import feelpp_b200 as fpp
from oracle import oracle as orc
from test_gmsh import cube, cube_with_sides, write_v2  # noqa: F401  (cube is a fixture)


@pytest.mark.gpu
@pytest.mark.parametrize("order", [1, 2])
def test_gpu_assembly_on_a_loaded_mesh_with_marked_faces(tmp_path, cube, order):
    """the drop-in path on a mesh that came through the .msh reader: volume term, Robin term on one marked side,
    Neumann term on two, Dirichlet elimination on another -- against the oracle on the same ranges"""
    from feelpp_b200 import cst, grad, gradt, id, idt, integrate, on, trans, expr, elements, markedfaces
    mesh0, tri, marker, ids, perm = cube
    path = str(tmp_path / "cube.msh")
    write_v2(path, mesh0, tri, marker, ids, perm, True)
    mesh = fpp.loadMesh(path)
    eng = fpp.Engine(0)
    Xh = fpp.Pch(mesh, order)
    ia, ja = Xh.graph()
    u = v = fpp.Element(Xh)
    a = fpp.form2(_test=Xh, _trial=Xh, engine=eng)
    l = fpp.form1(_test=Xh, engine=eng)
    a.assign(integrate(_range=elements(mesh), _expr=gradt(u) * trans(grad(v))))
    rob = markedfaces(mesh, "x0")
    a += integrate(_range=rob, _expr=cst(3.0) * (idt(u) * id(v)))
    K = orc.form2(orc.STIFFNESS, order, 1, orc.quad_order_form2(orc.STIFFNESS, order), mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    e, f = rob.faces()
    ref = orc.form2_faces(order, 1, 2 * order, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, e, f, orc.Coeff(value=3.0), a=K)
    assert np.abs(a.values() - ref).max() <= 1e-12 * np.abs(ref).max()
    neu = markedfaces(mesh, ["x1", "z0"])
    l.assign(integrate(_range=neu, _expr=cst(2.5) * id(v)))
    e, f = neu.faces()
    bref = orc.form1_faces(order, 1, order, mesh.xyz, mesh.conn, Xh.l2g, Xh.nDof(), e, f, orc.Coeff(value=2.5))
    assert np.abs(l.values() - bref).max() <= 1e-12 * np.abs(bref).max()
    assert abs(l.values().sum() - 5.0) < 1e-11
    dir_ = markedfaces(mesh, "y1")
    g = expr("1+2*x-y*z")
    a0, b0 = a.values(), l.values()
    a += on(_range=dir_, _rhs=l, _element=u, _expr=g, _type="elimination", _value_on_diagonal=1.0)
    rows = Xh.boundaryDofs(dir_.faces())
    assert (np.abs(Xh.dofPoints()[rows][:, 1] - 1) < 1e-14).all()
    a2, b2 = orc.zero_rows(ia, ja, a0, b0, rows, g.prog.evaluate(Xh.dofPoints()[rows]), 1, 1.0)
    assert np.array_equal(a.values(), a2)
    np.testing.assert_allclose(l.values(), b2, rtol=1e-13, atol=1e-14 * np.abs(b2).max())
    eng.close()
