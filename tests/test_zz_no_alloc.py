"""feelgpu_form2_assemble_noalloc (include/feelgpu.h): the allocation-free variant of the row-thread assembly, whose
coefficient scratch, parameter block and tables live in a static device array of the library.  Called through the C ABI and
compared with the regular entry point (feelgpu_form2_assemble, FEELGPU_ALGO_RPLAN) and with the oracle, on calls in a row
that share the static array.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import feelpp_b200 as fpp
from feelpp_b200 import elements, expr, grad, gradt, id, idt, idv, inner, integrate
from feelpp_b200.engine import _Keep
from oracle import oracle as orc
from oracle import parity

pytestmark = pytest.mark.gpu

NOALLOC_MAX_KQ = 2097152      # FEELGPU_NOALLOC_MAX_KQ


def _free_device_memory():
    """cudaMemGetInfo of the runtime the library is linked with; None when that runtime cannot be reached by name"""
    for name in ("libcudart.so.12", "libcudart.so"):
        try:
            rt = C.CDLL(name)
        except OSError:
            continue
        free, total = C.c_size_t(), C.c_size_t()
        if rt.cudaMemGetInfo(C.byref(free), C.byref(total)) == 0:
            return int(free.value)
    return None


class _Pair:
    """one pattern, two device value arrays: `a` filled by the regular entry point, `d_na` by the allocation-free one"""

    def __init__(self, engine, Xh):
        self.eng, self.Xh = engine, Xh
        self.a = fpp.form2(_test=Xh, _trial=Xh, engine=engine, algorithm=fpp.ALGO_RPLAN, pattern_flags=9)
        self.d_na = engine.malloc(8 * max(self.a.nnz, 1))
        self.keeps = []

    def desc(self, integ):
        keep = _Keep()
        self.keeps.append(keep)          # tables and coefficient data stay alive until close()
        return self.a._desc(integ, keep)

    def regular(self, integ, accumulate):
        if accumulate:
            self.a += integ
        else:
            self.a.assign(integ)
        return self.a.values()

    def noalloc(self, integ, accumulate):
        d = self.desc(integ)
        rc = self.eng.lib.feelgpu_form2_assemble_noalloc(self.a.pat, C.byref(d), accumulate, self.d_na, None)
        fpp._capi.check(rc, self.eng.ctx)
        self.eng.sync()
        return self.eng.to_host(self.d_na, max(self.a.nnz, 1))[:self.a.nnz]

    def close(self):
        self.eng.sync()
        for keep in self.keeps:
            for r in keep.refs:
                if isinstance(r, tuple):
                    self.eng.free(r[1])
        self.eng.free(self.d_na)
        self.a.close()


def _oracle(kind, Xh, quad, coeff):
    m = Xh.mesh
    ia, ja = Xh.graph()
    ref = orc.form2(kind, Xh.order, Xh.ncomp, quad, m.xyz, m.conn, Xh.l2g, ia, ja, coeff)
    scale = orc.form2(kind | orc.ABS_FLAG, Xh.order, Xh.ncomp, quad, m.xyz, m.conn, Xh.l2g, ia, ja, coeff)
    return ref, scale, ia, ja


@pytest.mark.parametrize("order", [1, 2, 3])
def test_noalloc_matches_regular_entry_point_expr(engine, order):
    """k(x)*stiffness (=), then k2(x)*mass (+=) with another quadrature size, then the first form again (=): three calls in a
    row through the shared static array, each against the regular entry point and the oracle"""
    mesh = fpp.unitCube(6, 0.15, True)
    Xh = fpp.Pch(mesh, order)
    u = v = fpp.Element(Xh)
    pr = _Pair(engine, Xh)
    try:
        k = expr("2+sin(x)*y-z^2")
        k2 = expr("1+0.5*x")
        f1 = integrate(_range=elements(mesh), _expr=k * inner(gradt(u), grad(v)))
        f2 = integrate(_range=elements(mesh), _expr=k2 * idt(u) * id(v))
        co = orc.Coeff(kind=orc.COEFF_EXPR, ops=k.prog.ops, consts=k.prog.consts)
        co2 = orc.Coeff(kind=orc.COEFF_EXPR, ops=k2.prog.ops, consts=k2.prog.consts)
        ref1, sc1, ia, ja = _oracle(orc.STIFFNESS, Xh, orc.quad_order_form2(orc.STIFFNESS, order, -1), co)
        ref2, sc2, _, _ = _oracle(orc.MASS, Xh, orc.quad_order_form2(orc.MASS, order, 1), co2)

        r = pr.regular(f1, 0)
        n = pr.noalloc(f1, 0)
        parity.check(n, r, sc1, ia, ja, "noalloc vs regular, k(x)*stiffness P%d" % order)
        parity.check(n, ref1, sc1, ia, ja, "noalloc vs oracle, k(x)*stiffness P%d" % order)

        r = pr.regular(f2, 1)
        free1 = _free_device_memory()
        n = pr.noalloc(f2, 1)
        free2 = _free_device_memory()
        parity.check(n, r, sc1 + sc2, ia, ja, "noalloc vs regular, += k2(x)*mass P%d" % order)
        parity.check(n, ref1 + ref2, sc1 + sc2, ia, ja, "noalloc vs oracle, += k2(x)*mass P%d" % order)
        if free1 is not None and free2 is not None:
            assert free2 >= free1, "device memory shrank across the allocation-free call (%d -> %d)" % (free1, free2)

        n = pr.noalloc(f1, 0)
        parity.check(n, ref1, sc1, ia, ja, "noalloc vs oracle, third call P%d" % order)
    finally:
        pr.close()


def test_noalloc_matches_regular_entry_point_field(engine):
    """idv(k) with k a P1 field on the device: P3 stiffness then P3 mass, two calls in a row"""
    mesh = fpp.unitCube(5, 0.15, True)
    Xh, Kh = fpp.Pch(mesh, 3), fpp.Pch(mesh, 1)
    kv = Kh.interpolate(lambda x, y, z: 1 + 0.5 * np.sin(2 * np.pi * x))
    k = fpp.Element(Kh, kv)
    u = v = fpp.Element(Xh)
    pr = _Pair(engine, Xh)
    try:
        co = orc.Coeff(kind=orc.COEFF_FIELD, field_order=1, field_l2g=Kh.l2g, field_values=kv)
        f1 = integrate(_range=elements(mesh), _expr=idv(k) * inner(gradt(u), grad(v)))
        f2 = integrate(_range=elements(mesh), _expr=idv(k) * idt(u) * id(v))
        ref1, sc1, ia, ja = _oracle(orc.STIFFNESS, Xh, orc.quad_order_form2(orc.STIFFNESS, 3, 1), co)
        ref2, sc2, _, _ = _oracle(orc.MASS, Xh, orc.quad_order_form2(orc.MASS, 3, 1), co)
        r = pr.regular(f1, 0)
        n = pr.noalloc(f1, 0)
        parity.check(n, r, sc1, ia, ja, "noalloc vs regular, idv(k)*stiffness P3")
        parity.check(n, ref1, sc1, ia, ja, "noalloc vs oracle, idv(k)*stiffness P3")
        r = pr.regular(f2, 0)
        n = pr.noalloc(f2, 0)
        parity.check(n, r, sc2, ia, ja, "noalloc vs regular, idv(k)*mass P3")
        parity.check(n, ref2, sc2, ia, ja, "noalloc vs oracle, idv(k)*mass P3")
    finally:
        pr.close()


def test_noalloc_refuses_what_it_cannot_do(engine):
    """sizes beyond the static scratch and forms without scratch are refused with a message, nothing is launched"""
    mesh = fpp.unitCube(48)                          # 663 552 elements x 4 points > FEELGPU_NOALLOC_MAX_KQ
    Xh = fpp.Pch(mesh, 1)
    u = v = fpp.Element(Xh)
    pr = _Pair(engine, Xh)
    try:
        k = expr("2+sin(x)*y-z^2")
        d = pr.desc(integrate(_range=elements(mesh), _expr=k * inner(gradt(u), grad(v))))
        assert mesh.numElements() * d.tables.nq > NOALLOC_MAX_KQ
        before = engine.launch_count()
        rc = engine.lib.feelgpu_form2_assemble_noalloc(pr.a.pat, C.byref(d), 0, pr.d_na, None)
        assert rc == 5
        msg = engine.lib.feelgpu_last_error(engine.ctx).decode()
        assert "FEELGPU_NOALLOC_MAX_KQ" in msg and str(mesh.numElements() * d.tables.nq) in msg
        # a constant coefficient uses no scratch: the regular entry point is the one to call
        d = pr.desc(integrate(_range=elements(mesh), _expr=inner(gradt(u), grad(v))))
        rc = engine.lib.feelgpu_form2_assemble_noalloc(pr.a.pat, C.byref(d), 0, pr.d_na, None)
        assert rc == 5 and "noalloc" in engine.lib.feelgpu_last_error(engine.ctx).decode()
        assert engine.launch_count() == before
        # the regular entry point still takes the large form
        pr.regular(integrate(_range=elements(mesh), _expr=k * inner(gradt(u), grad(v))), 0)
    finally:
        pr.close()
