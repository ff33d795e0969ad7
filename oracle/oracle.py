"""ctypes wrapper of the CPU oracle (oracle/_build/libfeel_oracle.so).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs; never by feelpp_b200/."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libfeel_oracle.so")
c_i32p, c_i64p, c_dblp = C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_double)

MASS, STIFFNESS, ELASTICITY = 0, 1, 2
ABS_FLAG = 0x100
COEFF_CONST, COEFF_FIELD, COEFF_EXPR = 0, 1, 2


class OCoeff(C.Structure):
    _fields_ = [("kind", C.c_int), ("value", C.c_double * 3), ("field_order", C.c_int), ("field_l2g", c_i32p),
                ("field_values", c_dblp), ("expr_len", C.c_int), ("expr_ops", c_i32p), ("expr_consts", c_dblp),
                ("lambda_", C.c_double), ("mu", C.c_double)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
        L = C.CDLL(_SO)
        L.oracle_quadrature.argtypes = [C.c_int, c_dblp]
        L.oracle_lagrange_points.argtypes = [C.c_int, c_dblp]
        L.oracle_tabulate.argtypes = [C.c_int, C.c_int, c_dblp, c_dblp, c_dblp]
        L.oracle_doftable.argtypes = [C.c_int64, C.c_int64, c_i32p, C.c_int, C.c_int, c_i32p]
        L.oracle_doftable.restype = C.c_int64
        L.oracle_graph.argtypes = [C.c_int64, C.c_int, c_i32p, C.c_int64, c_i64p, C.POINTER(c_i64p)]
        L.oracle_graph.restype = C.c_int64
        L.oracle_free.argtypes = [C.c_void_p]
        f2 = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, c_dblp, C.c_int64, c_i32p, c_i32p, c_i64p, c_i64p,
              C.POINTER(OCoeff)]
        L.oracle_form2.argtypes = f2 + [C.c_int64, C.c_int64, c_dblp]
        L.oracle_form2_threads.argtypes = f2 + [C.c_int, C.c_int64, C.c_int64, c_dblp]
        L.oracle_form2_threads.restype = C.c_double
        L.oracle_form1.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, c_dblp, C.c_int64, c_i32p, c_i32p,
                                   C.POINTER(OCoeff), C.c_int64, C.c_int64, c_dblp]
        L.oracle_local_matrix.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, c_dblp, C.POINTER(OCoeff), c_dblp]
        L.oracle_face_quadrature.argtypes = [C.c_int, c_dblp, c_dblp]
        ff = [C.c_int, C.c_int, C.c_int, c_dblp, c_i32p, c_i32p, C.c_int64, c_i32p, c_i32p, C.POINTER(OCoeff)]
        L.oracle_form1_faces.argtypes = ff + [c_dblp]
        L.oracle_form2_faces.argtypes = ff + [c_i64p, c_i64p, c_dblp]
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(t)


def nloc(order):
    return (order + 1) * (order + 2) * (order + 3) // 6


def quadrature(order):
    n = lib().oracle_quadrature(order, None)
    q = np.empty(4 * n)
    lib().oracle_quadrature(order, _p(q, c_dblp))
    q = q.reshape(n, 4)
    return q[:, :3].copy(), q[:, 3].copy()


def lagrange_points(order):
    p = np.empty((nloc(order), 3))
    lib().oracle_lagrange_points(order, _p(p, c_dblp))
    return p


def tabulate(order, xi):
    xi = np.ascontiguousarray(xi, dtype=np.float64).reshape(-1, 3)
    n, Q = nloc(order), xi.shape[0]
    phi, dphi = np.empty((n, Q)), np.empty((n, Q, 3))
    lib().oracle_tabulate(order, Q, _p(xi, c_dblp), _p(phi, c_dblp), _p(dphi, c_dblp))
    return phi, dphi


def quad_order_form2(kind, order, coeff_deg=0):
    return lib().oracle_quad_order_form2(kind, order, coeff_deg)


def quad_order_form1(order, coeff_deg=0):
    return lib().oracle_quad_order_form1(order, coeff_deg)


def doftable(npts, conn, order, ncomp=1):
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    l2g = np.empty((conn.shape[0], nloc(order) * ncomp), dtype=np.int32)
    nd = lib().oracle_doftable(npts, conn.shape[0], _p(conn, c_i32p), order, ncomp, _p(l2g, c_i32p))
    return l2g, int(nd)


def graph(l2g, ndof):
    l2g = np.ascontiguousarray(l2g, dtype=np.int32)
    ia = np.empty(ndof + 1, dtype=np.int64)
    jap = c_i64p()
    nnz = lib().oracle_graph(l2g.shape[0], l2g.shape[1], _p(l2g, c_i32p), ndof, _p(ia, c_i64p), C.byref(jap))
    ja = np.ctypeslib.as_array(jap, shape=(max(nnz, 1),))[:nnz].copy()
    lib().oracle_free(jap)
    return ia, ja


def doftable_mpi(conn, part, nparts, rank, order, ncomp=1):
    """DofTable + DataMap of one rank in MPI mode: (l2g[nMine][nd], nOwned, p2c[nWithGhost], ghost_owner, first_dof[nparts+1])"""
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    part = np.ascontiguousarray(part, dtype=np.int32)
    L = lib()
    L.oracle_doftable_mpi.argtypes = [C.c_int64, c_i32p, c_i32p, C.c_int, C.c_int, C.c_int, C.c_int, c_i32p, c_i64p, c_i64p,
                                      c_i32p, c_i64p]
    L.oracle_doftable_mpi.restype = C.c_int64
    nd = nloc(order) * ncomp
    nmine = int((part == rank).sum())
    l2g = np.empty((nmine, nd), dtype=np.int32)
    p2c = np.empty(max(nmine * nd, 1), dtype=np.int64)
    gho = np.empty(max(nmine * nd, 1), dtype=np.int32)
    first = np.empty(nparts + 1, dtype=np.int64)
    nown = C.c_int64()
    nwg = L.oracle_doftable_mpi(conn.shape[0], _p(conn, c_i32p), _p(part, c_i32p), nparts, rank, order, ncomp, _p(l2g, c_i32p),
                                C.byref(nown), _p(p2c, c_i64p), _p(gho, c_i32p), _p(first, c_i64p))
    return l2g, int(nown.value), p2c[:nwg].copy(), gho[:nwg - nown.value].copy(), first


def graph_mpi(conn, part, nparts, rank, order, ncomp, l2g, nwithghost):
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    part = np.ascontiguousarray(part, dtype=np.int32)
    l2g = np.ascontiguousarray(l2g, dtype=np.int32)
    L = lib()
    L.oracle_graph_mpi.argtypes = [C.c_int64, c_i32p, c_i32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, c_i32p, c_i64p,
                                   C.POINTER(c_i64p)]
    L.oracle_graph_mpi.restype = C.c_int64
    ia = np.empty(nwithghost + 1, dtype=np.int64)
    jap = c_i64p()
    nnz = L.oracle_graph_mpi(conn.shape[0], _p(conn, c_i32p), _p(part, c_i32p), nparts, rank, order, ncomp, nwithghost,
                             _p(l2g, c_i32p), _p(ia, c_i64p), C.byref(jap))
    ja = np.ctypeslib.as_array(jap, shape=(max(nnz, 1),))[:nnz].copy()
    L.oracle_free(jap)
    return ia, ja


class Coeff:
    """keeps the numpy buffers of an oracle_coeff alive"""

    def __init__(self, kind=COEFF_CONST, value=1.0, vector=None, field_order=0, field_l2g=None, field_values=None,
                 ops=None, consts=None, lam=0.0, mu=0.0):
        self.c = OCoeff()
        self.c.kind = kind
        self.c.value[0] = value
        if vector is not None:
            for k in range(3):
                self.c.value[k] = vector[k]
        self.c.lambda_, self.c.mu = lam, mu
        self._keep = []
        if field_l2g is not None:
            a = np.ascontiguousarray(field_l2g, dtype=np.int32)
            b = np.ascontiguousarray(field_values, dtype=np.float64)
            self._keep += [a, b]
            self.c.field_order, self.c.field_l2g, self.c.field_values = field_order, _p(a, c_i32p), _p(b, c_dblp)
        if ops is not None:
            a = np.ascontiguousarray(ops, dtype=np.int32)
            b = np.ascontiguousarray(consts, dtype=np.float64)
            self._keep += [a, b]
            self.c.expr_len, self.c.expr_ops, self.c.expr_consts = len(a), _p(a, c_i32p), _p(b, c_dblp)


def form2(kind, order, ncomp, quad_order, xyz, conn, l2g, ia, ja, coeff=None, a=None, elt_range=None, threads=0):
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    l2g = np.ascontiguousarray(l2g, dtype=np.int32)
    ia = np.ascontiguousarray(ia, dtype=np.int64)
    ja = np.ascontiguousarray(ja, dtype=np.int64)
    if a is None:
        a = np.zeros(int(ia[-1]))
    coeff = coeff or Coeff()
    eb, ee = elt_range if elt_range is not None else (0, conn.shape[0])
    args = [kind, order, ncomp, quad_order, xyz.shape[0], _p(xyz, c_dblp), conn.shape[0], _p(conn, c_i32p),
            _p(l2g, c_i32p), _p(ia, c_i64p), _p(ja, c_i64p), C.byref(coeff.c)]
    if threads > 0:
        secs = lib().oracle_form2_threads(*args, threads, eb, ee, _p(a, c_dblp))
        if secs < 0:
            raise RuntimeError("oracle: entry outside the pattern")
        return a, secs
    rc = lib().oracle_form2(*args, eb, ee, _p(a, c_dblp))
    if rc:
        raise RuntimeError("oracle: entry outside the pattern")
    return a


def form1(order, ncomp, quad_order, xyz, conn, l2g, ndof, coeff=None, b=None):
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    l2g = np.ascontiguousarray(l2g, dtype=np.int32)
    if b is None:
        b = np.zeros(ndof)
    coeff = coeff or Coeff()
    lib().oracle_form1(0, order, ncomp, quad_order, xyz.shape[0], _p(xyz, c_dblp), conn.shape[0], _p(conn, c_i32p),
                       _p(l2g, c_i32p), C.byref(coeff.c), 0, conn.shape[0], _p(b, c_dblp))
    return b


def face_quadrature(quad_order):
    """(xi[4][Qf][3], w[4][Qf]): IMTriangle<quad_order> mapped onto the four faces of the reference tet"""
    qf = lib().oracle_face_quadrature(quad_order, None, None)
    assert qf > 0
    xi, w = np.empty((4, qf, 3)), np.empty((4, qf))
    lib().oracle_face_quadrature(quad_order, _p(xi, c_dblp), _p(w, c_dblp))
    return xi, w


def boundary_faces(conn, npts):
    """(element, local face) of every face that belongs to one element only, in element order"""
    conn = np.asarray(conn, dtype=np.int64)
    F2P = np.array([[1, 2, 3], [0, 2, 3], [0, 1, 3], [0, 1, 2]])   # feelmesh/simplexordering.hpp:341-347
    fv = np.sort(conn[:, F2P], axis=2)
    key = (fv[..., 0] * npts + fv[..., 1]) * npts + fv[..., 2]
    _, inv, cnt = np.unique(key.ravel(), return_inverse=True, return_counts=True)
    e, f = np.nonzero((cnt[inv] == 1).reshape(key.shape))
    return e.astype(np.int32), f.astype(np.int32)


def form1_faces(order, ncomp, quad_order, xyz, conn, l2g, ndof, felt, flocal, coeff=None, b=None):
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    l2g = np.ascontiguousarray(l2g, dtype=np.int32)
    felt = np.ascontiguousarray(felt, dtype=np.int32)
    flocal = np.ascontiguousarray(flocal, dtype=np.int32)
    if b is None:
        b = np.zeros(ndof)
    coeff = coeff or Coeff()
    rc = lib().oracle_form1_faces(order, ncomp, quad_order, _p(xyz, c_dblp), _p(conn, c_i32p), _p(l2g, c_i32p), len(felt),
                                  _p(felt, c_i32p), _p(flocal, c_i32p), C.byref(coeff.c), _p(b, c_dblp))
    assert rc == 0, rc
    return b


def form2_faces(order, ncomp, quad_order, xyz, conn, l2g, ia, ja, felt, flocal, coeff=None, a=None):
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    l2g = np.ascontiguousarray(l2g, dtype=np.int32)
    felt = np.ascontiguousarray(felt, dtype=np.int32)
    flocal = np.ascontiguousarray(flocal, dtype=np.int32)
    ia = np.ascontiguousarray(ia, dtype=np.int64)
    ja = np.ascontiguousarray(ja, dtype=np.int64)
    if a is None:
        a = np.zeros(int(ia[-1]))
    coeff = coeff or Coeff()
    rc = lib().oracle_form2_faces(order, ncomp, quad_order, _p(xyz, c_dblp), _p(conn, c_i32p), _p(l2g, c_i32p), len(felt),
                                  _p(felt, c_i32p), _p(flocal, c_i32p), C.byref(coeff.c), _p(ia, c_i64p), _p(ja, c_i64p),
                                  _p(a, c_dblp))
    assert rc == 0, rc
    return a


def local_matrix(kind, order, ncomp, quad_order, G, coeff=None):
    G = np.ascontiguousarray(G, dtype=np.float64).reshape(4, 3)
    nd = nloc(order) * ncomp
    Ae = np.empty((nd, nd))
    coeff = coeff or Coeff()
    lib().oracle_local_matrix(kind, order, ncomp, quad_order, _p(G, c_dblp), C.byref(coeff.c), _p(Ae, c_dblp))
    return Ae


ON_ELIMINATION, ON_KEEP_DIAGONAL, ON_SYMMETRIC = 1, 2, 4   # ContextOn, feelalg/enums.hpp:56-66


def zero_rows(ia, ja, a, b, rows, x, strategy=ON_ELIMINATION, value_on_diagonal=1.0):
    """MatrixPetsc::zeroRows on a copy of (a, b); returns the eliminated (a, b)."""
    ia = np.ascontiguousarray(ia, dtype=np.int64)
    ja = np.ascontiguousarray(ja, dtype=np.int64)
    a, b = np.array(a, dtype=np.float64), np.array(b, dtype=np.float64)
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    x = np.ascontiguousarray(x, dtype=np.float64)
    lib().oracle_zero_rows.argtypes = [C.c_int64, c_i64p, c_i64p, C.c_int64, c_i32p, c_dblp, C.c_int, C.c_double, c_dblp,
                                       c_dblp]
    rc = lib().oracle_zero_rows(len(ia) - 1, _p(ia, c_i64p), _p(ja, c_i64p), len(rows), _p(rows, c_i32p), _p(x, c_dblp),
                                strategy, value_on_diagonal, _p(a, c_dblp), _p(b, c_dblp))
    if rc:
        raise RuntimeError("oracle: row outside the matrix")
    return a, b
