"""Engine-side objects: context, device mirrors of mesh/space/pattern, and the form2/form1 objects.

Everything here goes through the C ABI of include/feelgpu.h (libfeelgpu.so).  No computation happens in Python.
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from . import hostmirror as hm
from . import vf
from ._capi import Coeff, Form1Desc, Form2Desc, Tables, c_dblp, c_i32p, c_i64p, check, ptr

ALGO_AUTO, ALGO_SCATTER, ALGO_ROWPLAN, ALGO_EPLAN, ALGO_RPLAN = 0, 1, 2, 3, 4
PATTERN_SLOTMAP, PATTERN_ROWPLAN, PATTERN_EPLAN, PATTERN_RPLAN, PATTERN_ASYNC = 1, 2, 4, 8, 16


class Engine:
    """one context per GPU (feelgpu_context_create)."""

    def __init__(self, device=0):
        self.lib = capi.gpu()
        self.ctx = C.c_void_p()
        check(self.lib.feelgpu_context_create(device, C.byref(self.ctx)), None)
        self._meshes, self._spaces, self._faces = {}, {}, {}

    def close(self):
        if self.ctx:
            self.lib.feelgpu_context_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def sync(self):
        check(self.lib.feelgpu_sync(self.ctx), self.ctx)

    def launch_count(self):
        n = C.c_int64()
        check(self.lib.feelgpu_launch_count(self.ctx, C.byref(n)), self.ctx)
        return n.value

    # ---- device mirrors (cached per host object) ----
    def mesh(self, mesh):
        key = builtins_id(mesh)
        if key not in self._meshes:
            h = C.c_void_p()
            check(self.lib.feelgpu_mesh_create(self.ctx, mesh.numPoints(), ptr(mesh.xyz, c_dblp), mesh.numElements(),
                                               ptr(mesh.conn, c_i32p), C.byref(h)), self.ctx)
            self._meshes[key] = (h, mesh)
        return self._meshes[key][0]

    def space(self, Xh):
        key = builtins_id(Xh)
        if key not in self._spaces:
            h = C.c_void_p()
            p2c = getattr(Xh, "p2c", None)     # distributed spaces: process -> global-cluster ids, owned dofs first
            nown = getattr(Xh, "nDofOwned", Xh.nDof())
            check(self.lib.feelgpu_space_create(self.ctx, self.mesh(Xh.mesh), Xh.order, Xh.ncomp,
                                                ptr(Xh.l2g, c_i32p), Xh.nDof(), nown,
                                                ptr(p2c, c_i64p) if p2c is not None else None, C.byref(h)), self.ctx)
            self._spaces[key] = (h, Xh)
        return self._spaces[key][0]

    def graph(self, Xh):
        """(ia, ja) of the COUPLED graph of (Xh, Xh), built on the device (feelgpu_graph_create): the device-side
        Stencil::computeGraph; bit-exact with Xh.graph()"""
        g, nnz = C.c_void_p(), C.c_int64()
        check(self.lib.feelgpu_graph_create(self.space(Xh), C.byref(g), C.byref(nnz)), self.ctx)
        ia = np.empty(Xh.nDof() + 1, dtype=np.int64)
        ja = np.empty(nnz.value, dtype=np.int64)
        check(self.lib.feelgpu_graph_get(g, ptr(ia, c_i64p), ptr(ja, c_i64p)), self.ctx)
        self.lib.feelgpu_graph_destroy(g)
        return ia, ja

    def device_graph(self, Xh):
        """the same graph kept on the device next to its host copies: form2(..., device_graph=g) builds the pattern from it
        without sending the columns back (feelgpu_pattern_create_from_graph); g.destroy() when the forms are built"""
        return DeviceGraph(self, Xh)

    def faces(self, rng):
        """device list of the (element, local face) pairs of a face range (feelgpu_faces_create)"""
        key = builtins_id(rng)
        if key not in self._faces:
            h = C.c_void_p()
            e, f = rng.faces()
            check(self.lib.feelgpu_faces_create(self.mesh(rng.mesh), len(e), ptr(e, c_i32p), ptr(f, c_i32p), C.byref(h)),
                  self.ctx)
            self._faces[key] = (h, rng)
        return self._faces[key][0]

    # ---- raw device memory ----
    def malloc(self, nbytes):
        p = C.c_void_p()
        check(self.lib.feelgpu_malloc(self.ctx, nbytes, C.byref(p)), self.ctx)
        return p

    def free(self, p):
        check(self.lib.feelgpu_free(self.ctx, p), self.ctx)

    def to_device(self, arr):
        arr = np.ascontiguousarray(arr)
        p = self.malloc(arr.nbytes)
        check(self.lib.feelgpu_memcpy_h2d(self.ctx, p, arr.ctypes.data_as(C.c_void_p), arr.nbytes), self.ctx)
        return p

    def to_host(self, p, shape, dtype=np.float64):
        out = np.empty(shape, dtype=dtype)
        check(self.lib.feelgpu_memcpy_d2h(self.ctx, out.ctypes.data_as(C.c_void_p), p, out.nbytes), self.ctx)
        return out


def builtins_id(o):
    import builtins
    return builtins.id(o)


class _Keep:
    """keeps numpy buffers referenced by a ctypes descriptor alive"""

    def __init__(self):
        self.refs = []

    def arr(self, a, typ):
        a = np.ascontiguousarray(a)
        self.refs.append(a)
        return a.ctypes.data_as(typ)


def _fill_tables(keep, order, quad_order, field_order=None, faces=False):
    xi, w = hm.im_faces(quad_order) if faces else hm.im(quad_order)
    phi, dphi = hm.tabulate(order, xi)
    t = Tables()
    t.nq, t.n = len(w), phi.shape[0]
    t.w, t.xi = keep.arr(w, c_dblp), keep.arr(xi, c_dblp)
    t.phi, t.dphi = keep.arr(phi, c_dblp), keep.arr(dphi, c_dblp)
    t.nf, t.phif = 0, None
    if field_order is not None:
        phif, _ = hm.tabulate(field_order, xi)
        t.nf, t.phif = phif.shape[0], keep.arr(phif, c_dblp)
    return t


def _fill_coeff(eng, keep, cl, c):
    co = cl.coeff
    c.kind = co.kind()
    c.value[0] = co.scale
    c.value[1] = c.value[2] = 0.0
    if co.vector is not None:
        for k in range(3):
            c.value[k] = co.scale * co.vector[k]
    c.lambda_, c.mu = cl.lam, cl.mu
    if co.field is not None:
        fe = co.field
        c.field_space = eng.space(fe.space)
        dv = eng.to_device(np.asarray(fe.values, dtype=np.float64) * co.scale)
        keep.refs.append(("dev", dv))
        c.field_values = dv
    if co.prog is not None:
        ops, consts = co.prog.ops, co.prog.consts
        if co.scale != 1.0:  # fold the constant factor into the program: prog * scale
            ops = np.concatenate([ops, np.array([3, 6], dtype=np.int32)])
            consts = np.concatenate([consts, np.array([co.scale, 0.0])])
        c.expr_len = len(ops)
        c.expr_ops = keep.arr(ops.astype(np.int32), c_i32p)
        c.expr_consts = keep.arr(consts.astype(np.float64), c_dblp)


class DeviceGraph:
    def __init__(self, eng, Xh):
        self.eng, self.Xh = eng, Xh
        self.g, nnz = C.c_void_p(), C.c_int64()
        check(eng.lib.feelgpu_graph_create(eng.space(Xh), C.byref(self.g), C.byref(nnz)), eng.ctx)
        self.ia = np.empty(Xh.nDof() + 1, dtype=np.int64)
        self.ja = np.empty(nnz.value, dtype=np.int64)
        check(eng.lib.feelgpu_graph_get(self.g, ptr(self.ia, c_i64p), ptr(self.ja, c_i64p)), eng.ctx)

    def destroy(self):
        if self.g:
            self.eng.lib.feelgpu_graph_destroy(self.g)
            self.g = C.c_void_p()


class BilinearForm:
    """form2(_test=Xh,_trial=Xh): owns the CSR graph (host) and the device value array."""

    def __init__(self, test, trial, engine, pattern_flags=0, algorithm=ALGO_AUTO, device_graph=None):
        if test is not trial:
            raise NotImplementedError("test and trial spaces must be the same space on this path")
        self.Xh, self.eng, self.algorithm = test, engine, algorithm
        self.pat = C.c_void_p()
        if device_graph is not None:
            if device_graph.Xh is not test:
                raise ValueError("the device graph belongs to another space")
            self.ia, self.ja = device_graph.ia, device_graph.ja
            check(engine.lib.feelgpu_pattern_create_from_graph(device_graph.g, pattern_flags, C.byref(self.pat)), engine.ctx)
        else:
            self.ia, self.ja = test.graph()
            sp = engine.space(test)
            check(engine.lib.feelgpu_pattern_create(engine.ctx, sp, sp, test.nDof(), ptr(self.ia, c_i64p),
                                                    ptr(self.ja, c_i64p), pattern_flags, C.byref(self.pat)), engine.ctx)
        self.nnz = int(self.ia[-1])
        self.d_values = self._alloc_values()
        self._zeroed = False
        self.assign(None)

    def _alloc_values(self):
        return self.eng.malloc(8 * max(self.nnz, 1))

    def _desc(self, integ, keep):
        cl = vf.classify_form2(integ.expr, self.Xh.order)
        qo = integ.quad if integ.quad is not None else cl.quad_order
        d = Form2Desc()
        d.kind = cl.kind
        d.algorithm = self.algorithm
        fo = cl.coeff.field.space.order if cl.coeff.field is not None else None
        d.tables = _fill_tables(keep, self.Xh.order, qo, fo, isinstance(integ.range, hm.BoundaryFaceRange))
        _fill_coeff(self.eng, keep, cl, d.coeff)
        return d

    def assign(self, integ):
        """a = integrate(...): zero, then assemble (feelvf/bilinearform.hpp:1368-1371)."""
        if integ is None:
            check(self.eng.lib.feelgpu_memset_zero(self.eng.ctx, self.d_values, 8 * max(self.nnz, 1), None), self.eng.ctx)
            self.eng.sync()
            return self
        return self._assemble(integ, 0)

    def __iadd__(self, integ):
        """a += integrate(...) (feelvf/bilinearform.hpp:1458-1468); a += on(...) (feelvf/on.hpp:1126-1190)."""
        if isinstance(integ, vf.On):
            return self._on(integ)
        return self._assemble(integ, 1)

    def _on(self, o):
        dofs, vals = o.dofs_and_values()
        b = o.rhs.d_values if o.rhs is not None else None
        check(self.eng.lib.feelgpu_form2_zero_rows(self.pat, len(dofs), ptr(dofs, c_i32p), ptr(vals, c_dblp), o.strategy,
                                                   o.value_on_diagonal, self.d_values, b, None), self.eng.ctx)
        self.eng.sync()
        return self

    def _assemble(self, integ, accumulate):
        keep = _Keep()
        d = self._desc(integ, keep)
        if isinstance(integ.range, hm.BoundaryFaceRange):
            check(self.eng.lib.feelgpu_form2_assemble_faces(self.pat, self.eng.faces(integ.range), C.byref(d), accumulate,
                                                            self.d_values, None), self.eng.ctx)
        else:
            check(self.eng.lib.feelgpu_form2_assemble(self.pat, C.byref(d), accumulate, self.d_values, None), self.eng.ctx)
        self.eng.sync()
        for r in keep.refs:
            if isinstance(r, tuple):
                self.eng.free(r[1])
        return self

    def assemble_host(self, integ, accumulate=0, out=None):
        """end-to-end variant: values land in a host array (feelgpu_form2_assemble_host)."""
        keep = _Keep()
        d = self._desc(integ, keep)
        if out is None:
            out = np.zeros(max(self.nnz, 1))
        check(self.eng.lib.feelgpu_form2_assemble_host(self.pat, C.byref(d), accumulate, ptr(out, c_dblp)), self.eng.ctx)
        return out[:self.nnz]

    def values(self):
        return self.eng.to_host(self.d_values, max(self.nnz, 1))[:self.nnz]

    def multVector(self, x):
        """y = A x on the device (MatrixPetsc::multVector, feelalg/matrixpetsc.cpp:998-1062); x, y host arrays here"""
        dx = self.eng.to_device(np.ascontiguousarray(x, dtype=np.float64))
        dy = self.eng.malloc(8 * max(self.Xh.nDof(), 1))
        check(self.eng.lib.feelgpu_matvec(self.pat, self.d_values, dx, dy, None), self.eng.ctx)
        y = self.eng.to_host(dy, max(self.Xh.nDof(), 1))[:self.Xh.nDof()]
        self.eng.free(dx)
        self.eng.free(dy)
        return y

    def energy(self, v, u):
        """v^T A u (MatrixPetsc::energy, feelalg/matrixpetsc.cpp:2132-2170)"""
        return float(np.dot(v, self.multVector(u)))

    def close(self):
        """release the device mirrors of this form (feelgpu_pattern_destroy joins a background plan build first)"""
        if self.pat:
            self.eng.lib.feelgpu_pattern_destroy(self.pat)
            self.pat = C.c_void_p()

    def plans_ready(self, wait=False):
        """(ready, seconds the entity plan took to build): background plan builds of FEELGPU_PATTERN_ASYNC"""
        ready, secs = C.c_int(), C.c_double()
        check(self.eng.lib.feelgpu_pattern_plans_ready(self.pat, 1 if wait else 0, C.byref(ready), C.byref(secs)), self.eng.ctx)
        return bool(ready.value), secs.value

    def plan_stats(self):
        st = np.zeros(10, dtype=np.int64)
        check(self.eng.lib.feelgpu_pattern_plan_stats(self.pat, ptr(st, c_i64p)), self.eng.ctx)
        keys = ["has_plan", "ctas", "bundles", "units", "contributions", "sources_padded", "visits", "bytes",
                "max_chunks", "max_rounds"]
        out = dict(zip(keys, (int(x) for x in st)))
        check(self.eng.lib.feelgpu_pattern_eplan_stats(self.pat, ptr(st, c_i64p)), self.eng.ctx)
        ekeys = ["has_plan", "ctas", "tasks", "iterations", "iterations_padded", "visits", "outputs", "bytes",
                 "lds_wavefronts", "lds_wavefronts_ideal"]
        out["eplan"] = dict(zip(ekeys, (int(x) for x in st)))
        check(self.eng.lib.feelgpu_pattern_rplan_stats(self.pat, ptr(st, c_i64p)), self.eng.ctx)
        rkeys = ["has_plan", "patches", "units", "visits", "bytes", "max_slots", "max_visits", "max_rows", "build_ms"]
        out["rplan"] = dict(zip(rkeys, (int(x) for x in st)))
        return out

    def scipy(self):
        import scipy.sparse as sps
        return sps.csr_matrix((self.values(), self.ja, self.ia), shape=(self.Xh.nDof(), self.Xh.nDof()))


class LinearForm:
    """form1(_test=Xh)."""

    def __init__(self, test, engine):
        self.Xh, self.eng = test, engine
        self.sp = engine.space(test)
        self.d_values = engine.to_device(np.zeros(max(test.nDof(), 1)))

    def _desc(self, integ, keep):
        cl = vf.classify_form1(integ.expr, self.Xh.order)
        qo = integ.quad if integ.quad is not None else cl.quad_order
        d = Form1Desc()
        d.kind = cl.kind
        fo = cl.coeff.field.space.order if cl.coeff.field is not None else None
        d.tables = _fill_tables(keep, self.Xh.order, qo, fo, isinstance(integ.range, hm.BoundaryFaceRange))
        _fill_coeff(self.eng, keep, cl, d.coeff)
        return d

    def _assemble(self, integ, accumulate):
        keep = _Keep()
        d = self._desc(integ, keep)
        if isinstance(integ.range, hm.BoundaryFaceRange):
            check(self.eng.lib.feelgpu_form1_assemble_faces(self.sp, self.eng.faces(integ.range), C.byref(d), accumulate,
                                                            self.d_values, None), self.eng.ctx)
        else:
            check(self.eng.lib.feelgpu_form1_assemble(self.sp, C.byref(d), accumulate, self.d_values, None), self.eng.ctx)
        self.eng.sync()
        for r in keep.refs:
            if isinstance(r, tuple):
                self.eng.free(r[1])
        return self

    def assign(self, integ):
        return self._assemble(integ, 0)

    def __iadd__(self, integ):
        return self._assemble(integ, 1)

    def values(self):
        return self.eng.to_host(self.d_values, max(self.Xh.nDof(), 1))[:self.Xh.nDof()]


def form2(_test, _trial, engine, pattern_flags=0, algorithm=ALGO_AUTO, device_graph=None):
    """form2(_test=Xh,_trial=Xh) (feelvf/form.hpp:130-147): pattern COUPLED, matrix zeroed."""
    return BilinearForm(_test, _trial, engine, pattern_flags, algorithm, device_graph)


def form1(_test, engine):
    """form1(_test=Xh) (feelvf/form.hpp:92-106)."""
    return LinearForm(_test, engine)
