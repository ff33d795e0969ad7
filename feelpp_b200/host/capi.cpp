// capi.cpp -- plain-C entry points of the host-side mirror, for the Python test/bench harness (ctypes).
// These expose the pieces a Feel++ build already owns (mesh, DofTable, graph, reference tables); the engine itself
// is reached through include/feelgpu.h only.
#include <cstring>
#include <exception>
#include <stdexcept>
#include <string>

#include "doftable.hpp"
#include "ginac.hpp"
#include "gmsh.hpp"
#include "lagrange.hpp"
#include "mesh.hpp"
#include "stencil.hpp"

using namespace feelpp_b200;

static thread_local std::string g_err;
#define GUARD(body)                              \
    try { body; return 0; }                      \
    catch (std::exception const& e) { g_err = e.what(); return 1; }

extern "C" {

const char* feelhost_last_error() { return g_err.c_str(); }

// ---- mesh ----
int feelhost_unit_cube_sizes(int m, int64_t* nPts, int64_t* nElt)
{
    *nPts = (int64_t)(m + 1) * (m + 1) * (m + 1);
    *nElt = (int64_t)m * m * m * 6;
    return 0;
}
int feelhost_unit_cube(int m, double jitter, int shuffle, double* xyz, int32_t* conn)
{
    GUARD({
        auto mesh = unitCube(m, jitter, shuffle != 0);
        std::memcpy(xyz, mesh->xyz.data(), sizeof(double) * mesh->xyz.size());
        std::memcpy(conn, mesh->conn.data(), sizeof(int32_t) * mesh->conn.size());
    })
}

// boundaryfaces(mesh): elt/face have room for 4*nElt entries, *n receives the count
int feelhost_boundary_faces(int64_t nPts, int64_t nElt, const int32_t* conn, int32_t* elt, int32_t* face, int64_t* n)
{
    GUARD({
        std::vector<int32_t> e;
        std::vector<int32_t> f;
        boundaryFaces(nPts, nElt, conn, e, f);
        *n = (int64_t)e.size();
        if (!e.empty()) {
            std::memcpy(elt, e.data(), sizeof(int32_t) * e.size());
            std::memcpy(face, f.data(), sizeof(int32_t) * f.size());
        }
    })
}

// ---- Gmsh .msh reader (feelfilters/importergmsh.hpp:983-1420) ----
struct feelhost_gmsh_s { GmshMesh m; };
int feelhost_gmsh_open(const char* path, int elementaryAsPhysical, feelhost_gmsh_s** out, int64_t* nPts, int64_t* nElt,
                       int64_t* nTri, int64_t* namesLen, double* version, int* binary, int* numPartitions)
{
    GUARD({
        auto* h = new feelhost_gmsh_s;
        try { h->m = readGmsh(path, elementaryAsPhysical != 0); }
        catch (...) { delete h; throw; }
        *out = h;
        *nPts = (int64_t)h->m.pointId.size();
        *nElt = (int64_t)h->m.eltMarker.size();
        *nTri = (int64_t)h->m.triMarker.size();
        *namesLen = (int64_t)h->m.physicalNames.size();
        *version = h->m.version;
        *binary = h->m.binary ? 1 : 0;
        *numPartitions = h->m.numPartitions;
    })
}
int feelhost_gmsh_get(feelhost_gmsh_s* h, double* xyz, int64_t* pointId, int32_t* conn, int32_t* eltMarker, int32_t* tri,
                      int32_t* triMarker, char* names, int32_t* eltPartition)
{
    GUARD({
        GmshMesh const& m = h->m;
        std::memcpy(xyz, m.xyz.data(), sizeof(double) * m.xyz.size());
        std::memcpy(pointId, m.pointId.data(), sizeof(int64_t) * m.pointId.size());
        std::memcpy(conn, m.conn.data(), sizeof(int32_t) * m.conn.size());
        std::memcpy(eltMarker, m.eltMarker.data(), sizeof(int32_t) * m.eltMarker.size());
        std::memcpy(eltPartition, m.eltPartition.data(), sizeof(int32_t) * m.eltPartition.size());
        if (!m.tri.empty()) std::memcpy(tri, m.tri.data(), sizeof(int32_t) * m.tri.size());
        if (!m.triMarker.empty()) std::memcpy(triMarker, m.triMarker.data(), sizeof(int32_t) * m.triMarker.size());
        if (!m.physicalNames.empty()) std::memcpy(names, m.physicalNames.data(), m.physicalNames.size());
    })
}
int feelhost_gmsh_free(feelhost_gmsh_s* h) { delete h; return 0; }

// ---- DofTable ----
int feelhost_doftable(int64_t nPts, const double* xyz, int64_t nElt, const int32_t* conn, int order, int ncomp,
                      int32_t* l2g, int64_t* nDof)
{
    GUARD({
        Mesh mesh;
        mesh.xyz.assign(xyz, xyz + 3 * nPts);
        mesh.conn.assign(conn, conn + 4 * nElt);
        DofTable dof;
        dof.build(mesh, order, ncomp);
        std::memcpy(l2g, dof.l2g.data(), sizeof(int32_t) * dof.l2g.size());
        *nDof = dof.nDof;
    })
}

// ---- graph: two calls (sizes, then fill) so that the caller owns the arrays ----
struct feelhost_graph_s { GraphCSR g; };
int feelhost_graph_build(int64_t nElt, int order, int ncomp, const int32_t* l2g, int64_t nDof, int nthreads,
                         feelhost_graph_s** out, int64_t* nnz)
{
    GUARD({
        DofTable dof;
        dof.order = order;
        dof.ncomp = ncomp;
        dof.n = nLocalDofTet(order);
        dof.nd = dof.n * ncomp;
        dof.nDof = nDof;
        dof.l2g.assign(l2g, l2g + nElt * dof.nd);
        auto* h = new feelhost_graph_s;
        computeGraph(dof, nElt, h->g, nthreads);
        *out = h;
        *nnz = h->g.nnz();
    })
}
int feelhost_graph_get(feelhost_graph_s* h, int64_t* ia, int64_t* ja)
{
    std::memcpy(ia, h->g.ia.data(), sizeof(int64_t) * h->g.ia.size());
    std::memcpy(ja, h->g.ja.data(), sizeof(int64_t) * h->g.ja.size());
    return 0;
}
int feelhost_graph_free(feelhost_graph_s* h) { delete h; return 0; }

// ---- reference tables ----
int feelhost_quadrature(int order, int* nq, double* xi, double* w)
{
    GUARD({
        Quadrature q = im(order);
        *nq = q.nq;
        if (xi) std::memcpy(xi, q.xi.data(), sizeof(double) * 3 * q.nq);
        if (w) std::memcpy(w, q.w.data(), sizeof(double) * q.nq);
    })
}
int feelhost_face_quadrature(int order, int* nq, double* xi, double* w)
{
    GUARD({
        Quadrature q = imFaces(order);
        *nq = q.nq;
        if (xi) std::memcpy(xi, q.xi.data(), sizeof(double) * 3 * q.nq);
        if (w) std::memcpy(w, q.w.data(), sizeof(double) * q.nq);
    })
}
int feelhost_nlocaldof(int order) { return nLocalDofTet(order); }
int feelhost_lagrange_points(int order, double* pts)
{
    GUARD({
        auto p = lagrangePoints(order);
        std::memcpy(pts, p.data(), sizeof(double) * p.size());
    })
}
int feelhost_tabulate(int order, int nq, const double* xi, double* phi, double* dphi)
{
    GUARD({
        BasisTables T = tabulate(order, nq, xi);
        std::memcpy(phi, T.phi.data(), sizeof(double) * T.phi.size());
        std::memcpy(dphi, T.dphi.data(), sizeof(double) * T.dphi.size());
    })
}

// ---- expr("...") ----
int feelhost_parse_expr(const char* s, int cap, int32_t* ops, double* consts, int* len, int* polyDegree)
{
    GUARD({
        GinacProgram g = parseExpr(s);
        if ((int)g.ops.size() > cap) throw std::length_error("parse_expr: program longer than the caller's buffer");
        *len = (int)g.ops.size();
        *polyDegree = g.polyDegree;
        std::memcpy(ops, g.ops.data(), sizeof(int32_t) * g.ops.size());
        std::memcpy(consts, g.consts.data(), sizeof(double) * g.consts.size());
    })
}

} // extern "C"

// ---- distributed setup helper: build the local CSR rows of one rank -------------------------------------------
// Row k of the result is row rid[k] of graph A (k < nFromA: owned rows, complete pattern) or of graph B (ghost rows,
// pattern of the local elements only); columns are mapped through colmap (process-local -> global-cluster ids,
// feelalg/datamap.hpp:467) and sorted ascending as GraphCSR::close does (feelalg/graphcsr.cpp:994-1019).
#include <algorithm>
#include <thread>
#include <vector>
extern "C" int feelhost_csr_select_remap_sort(int64_t nNew, int64_t nFromA, const int64_t* rid, const int64_t* iaA,
                                               const int64_t* jaA, const int64_t* iaB, const int64_t* jaB,
                                               const int64_t* colmap, int64_t* iaOut, int64_t* jaOut /*may be NULL: count only*/)
{
    iaOut[0] = 0;
    for (int64_t k = 0; k < nNew; ++k) {
        const int64_t* ia = k < nFromA ? iaA : iaB;
        iaOut[k + 1] = iaOut[k] + (ia[rid[k] + 1] - ia[rid[k]]);
    }
    if (!jaOut) return 0;
    int nth = (int)std::max(1u, std::thread::hardware_concurrency());
    nth = std::min(nth, 32);
    auto work = [&](int t) {
        for (int64_t k = nNew * t / nth; k < nNew * (t + 1) / nth; ++k) {
            const int64_t* ia = k < nFromA ? iaA : iaB;
            const int64_t* ja = k < nFromA ? jaA : jaB;
            int64_t* out = jaOut + iaOut[k];
            const int64_t b = ia[rid[k]], e = ia[rid[k] + 1];
            for (int64_t q = b; q < e; ++q) out[q - b] = colmap[ja[q]];
            std::sort(out, out + (e - b));
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nth; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& x : th) x.join();
    return 0;
}
