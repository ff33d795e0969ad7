"""Host-side objects of the harness: mesh, function space (DofTable), graph, reference tables, expr("...").

They wrap libfeelhost.so (C++), and hold what a Feel++ build already owns when it reaches
Integrator::assemble (feelvf/integrator.hpp:975): Mesh (feeldiscr/mesh.hpp:162), Pch<k>/Pchv<k> spaces
(feeldiscr/pch.hpp:85-90, pchv.hpp:43-48), the CSR graph (feelalg/graphcsr.hpp:287-295).
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from ._capi import c_dblp, c_i32p, c_i64p, hcheck, ptr


class Mesh:
    def __init__(self, xyz, conn):
        self.xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
        self.conn = np.ascontiguousarray(conn, dtype=np.int32).reshape(-1, 4)

    def numPoints(self):
        return self.xyz.shape[0]

    def numElements(self):
        return self.conn.shape[0]

    numGlobalElements = numElements

    tri = None

    def markerId(self, m):
        """marker name ($PhysicalNames) or number -> number (MeshBase::markerId)"""
        if isinstance(m, str):
            if m not in getattr(self, "markerNames", {}):
                raise KeyError("unknown marker %r" % m)
            return self.markerNames[m][1]
        return int(m)


def unitCube(m, jitter=0.0, shuffle=False):
    """Synthetic unit-cube tet mesh, m^3 cells x 6 tets (stands in for gmsh's unitCube(), feelfilters/unitcube.cpp:38-52)."""
    h = capi.host()
    npts, nelt = C.c_int64(), C.c_int64()
    h.feelhost_unit_cube_sizes(m, C.byref(npts), C.byref(nelt))
    xyz = np.empty((npts.value, 3), dtype=np.float64)
    conn = np.empty((nelt.value, 4), dtype=np.int32)
    hcheck(h.feelhost_unit_cube(m, float(jitter), int(bool(shuffle)), ptr(xyz, c_dblp), ptr(conn, c_i32p)))
    return Mesh(xyz, conn)


class BoundaryFaceRange:
    """boundaryfaces(mesh) (feelmesh/filters.hpp): used by on() and by face integrals.  faces() lists the faces that belong
    to one element only as (element, local face) pairs in element order (face.element(0).id(), face.pos_first())."""

    def __init__(self, mesh):
        self.mesh = mesh
        self._faces = None

    def faces(self):
        if self._faces is None:
            mesh = self.mesh
            e = np.empty(4 * mesh.numElements(), dtype=np.int32)
            f = np.empty(4 * mesh.numElements(), dtype=np.int32)
            n = C.c_int64()
            hcheck(capi.host().feelhost_boundary_faces(mesh.numPoints(), mesh.numElements(), ptr(mesh.conn, c_i32p),
                                                       ptr(e, c_i32p), ptr(f, c_i32p), C.byref(n)))
            self._faces = (e[:n.value].copy(), f[:n.value].copy())
        return self._faces


def boundaryfaces(mesh):
    return BoundaryFaceRange(mesh)


class MarkedFaceRange(BoundaryFaceRange):
    """markedfaces(mesh, marker) (feelmesh/filters.hpp): the boundary faces that carry one of the given markers, as
    (element, local face) pairs in element order.  Markers come from the triangles of a .msh file (loadMesh)."""

    def __init__(self, mesh, markers):
        super().__init__(mesh)
        self.markers = markers

    def faces(self):
        if self._faces is None:
            mesh = self.mesh
            if getattr(mesh, "tri", None) is None:
                raise ValueError("markedfaces(): the mesh carries no marked faces (use loadMesh on a .msh with triangles)")
            ids = [mesh.markerId(m) for m in (self.markers if isinstance(self.markers, (list, tuple)) else [self.markers])]
            e, f = BoundaryFaceRange(mesh).faces()
            F2P = np.array([[1, 2, 3], [0, 2, 3], [0, 1, 3], [0, 1, 2]])
            # match on the three sorted point ids as one 24-byte record (no packing into one integer: npts^3 overflows
            # int64 beyond ~2 M points)
            rec = np.dtype([("a", np.int64), ("b", np.int64), ("c", np.int64)])
            fv = np.ascontiguousarray(np.sort(mesh.conn.astype(np.int64)[e[:, None], F2P[f]], axis=1)).view(rec).ravel()
            tv = np.ascontiguousarray(np.sort(mesh.tri[np.isin(mesh.triMarker, ids)].astype(np.int64), axis=1)).view(rec).ravel()
            sel = np.isin(fv, tv)
            self._faces = (np.ascontiguousarray(e[sel]), np.ascontiguousarray(f[sel]))
        return self._faces


def markedfaces(mesh, markers):
    return MarkedFaceRange(mesh, markers)


def loadMesh(path, elementaryAsPhysical=False):
    """loadMesh(_mesh=new Mesh<Simplex<3>>, _filename="x.msh") (feelfilters/loadmesh.hpp -> ImporterGmsh::readFromFile,
    feelfilters/importergmsh.hpp:983-1420): MSH 2.x and MSH 4.1, ASCII or binary, tetrahedra + marked triangles.
    Point indices follow the ascending node numbers of the file (mesh.pointId holds them).  A partitioned MSH 2.2 file
    (tags numPartitions, partition, -ghosts) gives mesh.eltPartition, the `part` input of parallel.LocalProblem: the
    reference keeps the elements of its own partition the same way (importergmsh.hpp:1272-1281)."""
    h = capi.host()
    hd = C.c_void_p()
    npts, nelt, ntri, nlen = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
    version, binary, nparts = C.c_double(), C.c_int(), C.c_int()
    hcheck(h.feelhost_gmsh_open(str(path).encode(), int(bool(elementaryAsPhysical)), C.byref(hd), C.byref(npts),
                                C.byref(nelt), C.byref(ntri), C.byref(nlen), C.byref(version), C.byref(binary),
                                C.byref(nparts)))
    try:
        xyz = np.empty((npts.value, 3), dtype=np.float64)
        pid = np.empty(npts.value, dtype=np.int64)
        conn = np.empty((nelt.value, 4), dtype=np.int32)
        em = np.empty(nelt.value, dtype=np.int32)
        ep = np.empty(nelt.value, dtype=np.int32)
        tri = np.empty((ntri.value, 3), dtype=np.int32)
        tm = np.empty(ntri.value, dtype=np.int32)
        names = C.create_string_buffer(max(nlen.value, 1))
        hcheck(h.feelhost_gmsh_get(hd, ptr(xyz, c_dblp), ptr(pid, c_i64p), ptr(conn, c_i32p), ptr(em, c_i32p),
                                   ptr(tri, c_i32p), ptr(tm, c_i32p), names, ptr(ep, c_i32p)))
    finally:
        h.feelhost_gmsh_free(hd)
    mesh = Mesh(xyz, conn)
    mesh.pointId, mesh.eltMarker, mesh.tri, mesh.triMarker = pid, em, tri, tm
    mesh.version, mesh.binary = version.value, bool(binary.value)
    mesh.eltPartition, mesh.numPartitions = ep, nparts.value     # partition tags of the file (0-based), input of LocalProblem
    mesh.markerNames = {}
    for line in names.raw[:nlen.value].decode().splitlines():
        dim, tag, name = line.split(" ", 2)
        mesh.markerNames[name] = (int(dim), int(tag))
    return mesh


class ElementRange:
    def __init__(self, mesh):
        self.mesh = mesh


def elements(mesh):
    """elements(mesh) (feelmesh/filters.hpp:201)."""
    return ElementRange(mesh)


def nLocalDof(order):
    return (order + 1) * (order + 2) * (order + 3) // 6


class FunctionSpace:
    """Continuous Lagrange Pk space on a tet mesh; holds the flattened DofTable."""

    def __init__(self, mesh, order, ncomp=1):
        h = capi.host()
        self.mesh, self.order, self.ncomp = mesh, order, ncomp
        self.n = nLocalDof(order)
        self.nd = self.n * ncomp
        self.l2g = np.empty((mesh.numElements(), self.nd), dtype=np.int32)
        nd = C.c_int64()
        hcheck(h.feelhost_doftable(mesh.numPoints(), ptr(mesh.xyz, c_dblp), mesh.numElements(),
                                   ptr(mesh.conn, c_i32p), order, ncomp, ptr(self.l2g, c_i32p), C.byref(nd)))
        self._ndof = nd.value
        self._graph = None
        self._dofpts = None

    def nDof(self):
        return self._ndof

    nLocalDofWithGhost = nDof
    nLocalDofWithoutGhost = nDof

    def graph(self, nthreads=0):
        """CSR graph of (this, this): (ia, ja) int64, ascending columns (Stencil::computeGraph + GraphCSR::close)."""
        if self._graph is None:
            h = capi.host()
            hd, nnz = C.c_void_p(), C.c_int64()
            hcheck(h.feelhost_graph_build(self.mesh.numElements(), self.order, self.ncomp, ptr(self.l2g, c_i32p),
                                          self._ndof, nthreads, C.byref(hd), C.byref(nnz)))
            ia = np.empty(self._ndof + 1, dtype=np.int64)
            ja = np.empty(nnz.value, dtype=np.int64)
            h.feelhost_graph_get(hd, ptr(ia, c_i64p), ptr(ja, c_i64p))
            h.feelhost_graph_free(hd)
            self._graph = (ia, ja)
        return self._graph

    def dofPoints(self):
        """physical coordinates of the scalar nodes, [nDof/ncomp][3] (for interpolation in tests)."""
        if self._dofpts is None:
            ref = lagrangePoints(self.order)
            lam = np.stack([-(1 + ref.sum(1)) / 2, (1 + ref[:, 0]) / 2, (1 + ref[:, 1]) / 2, (1 + ref[:, 2]) / 2], 1)
            P = self.mesh.xyz[self.mesh.conn]               # [nElt][4][3]
            X = np.einsum("nk,ekc->enc", lam, P)             # [nElt][n][3]
            pts = np.zeros((self._ndof // self.ncomp, 3))
            pts[self.l2g[:, :self.n] // self.ncomp] = X
            self._dofpts = pts
        return self._dofpts

    def interpolate(self, f):
        """nodal interpolant of a scalar callable f(x,y,z) -> array (scalar spaces)."""
        assert self.ncomp == 1
        p = self.dofPoints()
        return np.asarray(f(p[:, 0], p[:, 1], p[:, 2]), dtype=np.float64)

    def boundaryDofs(self, faces=None):
        """scalar node ids carried by the boundary faces (faces of exactly one element): what
        on(_range=boundaryfaces(mesh), ...) visits through the face dof table (feelvf/on.hpp:395-449).  With
        faces = (element, local face) arrays -- markedfaces(mesh, marker).faces() -- only those faces are visited."""
        F2P = np.array([[1, 2, 3], [0, 2, 3], [0, 1, 3], [0, 1, 2]])        # feelmesh/simplexordering.hpp:341-347
        E2P = np.array([[1, 2], [2, 0], [0, 1], [0, 3], [1, 3], [2, 3]])    # :283-291
        if faces is None:
            faces = BoundaryFaceRange(self.mesh).faces()
        onb = np.zeros((self.mesh.numElements(), 4), dtype=bool)             # [nElt][4] the faces visited
        onb[faces[0], faces[1]] = True
        nodes = self.l2g[:, :self.n] // self.ncomp
        out = []
        npe = self.order - 1
        for f in range(4):
            sel = onb[:, f]
            loc = list(F2P[f])
            for e, (a, b) in enumerate(E2P):
                if a in F2P[f] and b in F2P[f]:
                    loc += [4 + e * npe + l for l in range(npe)]
            if self.order >= 3:
                loc.append(4 + 6 * npe + f)
            out.append(nodes[sel][:, loc].ravel())
        return np.unique(np.concatenate(out)).astype(np.int32)


def Pch(mesh, order):
    """Pch<order>(mesh) (feeldiscr/pch.hpp:85-90)."""
    return FunctionSpace(mesh, order, 1)


def Pchv(mesh, order):
    """Pchv<order>(mesh) (feeldiscr/pchv.hpp:43-48)."""
    return FunctionSpace(mesh, order, 3)


def im(order):
    """im(3,order,simplex): (xi[nq][3], w[nq]) (feelpoly/imsimplex.cpp:8375-8396)."""
    h = capi.host()
    nq = C.c_int()
    hcheck(h.feelhost_quadrature(order, C.byref(nq), None, None))
    xi = np.empty((nq.value, 3))
    w = np.empty(nq.value)
    hcheck(h.feelhost_quadrature(order, C.byref(nq), ptr(xi, c_dblp), ptr(w, c_dblp)))
    return xi, w


def im_faces(order):
    """the triangle rule of that order carried onto the 4 faces of the reference tet: (xi[4*Qf][3], w[4*Qf]), block f =
    local face f, weights times the reference-face area factor (feelpoly/pointsetquadrature.hpp:696-760)."""
    h = capi.host()
    nq = C.c_int()
    hcheck(h.feelhost_face_quadrature(order, C.byref(nq), None, None))
    xi = np.empty((nq.value, 3))
    w = np.empty(nq.value)
    hcheck(h.feelhost_face_quadrature(order, C.byref(nq), ptr(xi, c_dblp), ptr(w, c_dblp)))
    return xi, w


def lagrangePoints(order):
    h = capi.host()
    pts = np.empty((nLocalDof(order), 3))
    hcheck(h.feelhost_lagrange_points(order, ptr(pts, c_dblp)))
    return pts


def tabulate(order, xi):
    """PreCompute tables: phi[n][nq], dphi[n][nq][3] (feelpoly/polynomialset.hpp:1052-1083)."""
    h = capi.host()
    xi = np.ascontiguousarray(xi, dtype=np.float64).reshape(-1, 3)
    n, nq = nLocalDof(order), xi.shape[0]
    phi = np.empty((n, nq))
    dphi = np.empty((n, nq, 3))
    hcheck(h.feelhost_tabulate(order, nq, ptr(xi, c_dblp), ptr(phi, c_dblp), ptr(dphi, c_dblp)))
    return phi, dphi


class GinacProgram:
    def __init__(self, source):
        h = capi.host()
        self.source = source
        ops = np.zeros(64, dtype=np.int32)
        consts = np.zeros(64)
        ln, deg = C.c_int(), C.c_int()
        hcheck(h.feelhost_parse_expr(source.encode(), 64, ptr(ops, c_i32p), ptr(consts, c_dblp), C.byref(ln),
                                     C.byref(deg)))
        self.ops = ops[:ln.value].copy()
        self.consts = consts[:ln.value].copy()
        self.polyDegree = deg.value

    def evaluate(self, p):
        """the postfix program at the points p[:,3] (host side: Dirichlet values of on())"""
        p = np.asarray(p, dtype=np.float64)
        st = []
        un = {8: np.negative, 9: np.sin, 10: np.cos, 11: np.exp, 12: np.log, 13: np.sqrt}
        bi = {4: np.add, 5: np.subtract, 6: np.multiply, 7: np.divide, 14: np.power}
        for op, c in zip(self.ops, self.consts):
            if op <= 2:
                st.append(p[:, op].copy())
            elif op == 3:
                st.append(np.full(p.shape[0], c))
            elif op in un:
                st.append(un[op](st.pop()))
            else:
                b = st.pop()
                a = st.pop()
                st.append(bi[op](a, b))
        return st[0]
