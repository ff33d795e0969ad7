// adapter_mock.cpp -- the Feel++-side adapter of INTEGRATION.md section 1, compiled WITHOUT Feel++ (g++ only).
//
// Feel++ cannot be built in this image, so the adapter stub of INTEGRATION.md would otherwise stay prose.  This file holds
// mock objects with the accessors the real ones offer on the path -- Mesh/GeoElement (elt.point(i).id(), elt.G():
// feelmesh/geond.hpp:626), DofTable (localToGlobalIndices(eltId): feeldiscr/doftable.hpp:732-737, nLocalDofWithGhost /
// WithoutGhost, mapGlobalProcessToGlobalCluster: feelalg/datamap.hpp:226,251,467), GraphCSR (ia()/ja() as size_t:
// feelalg/graphcsr.hpp:287-295), IMGeneral points/weights and PreCompute phi/grad -- and the SAME flattening code as the
// stub: flatten_points, flatten_connectivity, flatten_l2g, ia64/ja64, fill_tables, then the C ABI calls
// (mesh_create, space_create, pattern_create, form2_assemble_host, form1_assemble_host).  It assembles a P1 Laplacian
// and a load vector on a small Kuhn cube and writes ia / ja / a / b to a file that the test compares with the oracle.
//
// usage: adapter_mock <cells per edge> <output file>      exit codes: 0 ok, 2 bad usage, 3 engine error (message on stderr)
#include <feelgpu.h>

#include <array>
#include <cstdio>
#include <cstdlib>
#include <set>
#include <vector>

namespace mock {

struct Point { int id_; std::array<double, 3> x; int id() const { return id_; } };
struct Element {
    int id_; std::array<const Point*, 4> p;
    int id() const { return id_; }
    const Point& point(int i) const { return *p[i]; }
    std::array<double, 12> G() const {                 // 3 x 4, column = vertex (geond.hpp:626)
        std::array<double, 12> g{};
        for (int v = 0; v < 4; ++v)
            for (int c = 0; c < 3; ++c) g[c * 4 + v] = p[v]->x[c];
        return g;
    }
};
struct Mesh {
    std::vector<Point> points;
    std::vector<Element> elements;
    size_t numPoints() const { return points.size(); }
    size_t numElements() const { return elements.size(); }
};
// Pch<1>: one dof per point, numbered by first touch in element order (feeldiscr/doftable.hpp:1213-1338)
struct DofTable {
    std::vector<std::array<int, 4>> el_l2g;
    size_t ndof = 0;
    explicit DofTable(const Mesh& m) {
        std::vector<int> of(m.numPoints(), -1);
        for (const Element& e : m.elements) {
            std::array<int, 4> r;
            for (int i = 0; i < 4; ++i) {
                int& d = of[e.point(i).id()];
                if (d < 0) d = (int)ndof++;
                r[i] = d;
            }
            el_l2g.push_back(r);
        }
    }
    std::array<int, 4> const& localToGlobalIndices(size_t eltId) const { return el_l2g[eltId]; }
    size_t nLocalDofWithGhost() const { return ndof; }
    size_t nLocalDofWithoutGhost() const { return ndof; }
    std::vector<size_t> mapGlobalProcessToGlobalCluster() const {
        std::vector<size_t> m(ndof);
        for (size_t k = 0; k < ndof; ++k) m[k] = k;     // sequential: identity (doftable.hpp:1963-1971)
        return m;
    }
};
// Stencil::computeGraph (COUPLED) + GraphCSR::close: std::set rows, diagonal always present
struct GraphCSR {
    std::vector<size_t> ia_, ja_;
    GraphCSR(const Mesh& m, const DofTable& d) {
        std::vector<std::set<size_t>> rows(d.ndof);
        for (size_t e = 0; e < m.numElements(); ++e)
            for (int i : d.localToGlobalIndices(e))
                for (int j : d.localToGlobalIndices(e)) rows[i].insert(j);
        ia_.push_back(0);
        for (size_t r = 0; r < d.ndof; ++r) {
            rows[r].insert(r);
            ja_.insert(ja_.end(), rows[r].begin(), rows[r].end());
            ia_.push_back(ja_.size());
        }
    }
    std::vector<size_t> const& ia() const { return ia_; }
    std::vector<size_t> const& ja() const { return ja_; }
};
// IMTetrahedra<1> (feelpoly/imsimplex.cpp:1808-1827) and the P1 PreCompute on the reference tet (-1,-1,-1),(1,-1,-1),...
struct IM { std::vector<double> points{-0.5, -0.5, -0.5}, weights{1.33333333333333}; };
struct PreCompute {
    std::vector<double> phi{0.25, 0.25, 0.25, 0.25};                                   // [n][Q]
    std::vector<double> grad{-0.5, -0.5, -0.5, 0.5, 0, 0, 0, 0.5, 0, 0, 0, 0.5};       // [n][Q][3]
};

Mesh unitCube(int m) {   // Kuhn split, point id = lexicographic grid index
    Mesh M;
    const int np = m + 1;
    for (int k = 0; k < np; ++k)
        for (int j = 0; j < np; ++j)
            for (int i = 0; i < np; ++i) M.points.push_back(Point{(int)M.points.size(), {double(i) / m, double(j) / m, double(k) / m}});
    static const int perm[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 0, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}};
    auto pid = [&](int i, int j, int k) { return (k * np + j) * np + i; };
    for (int k = 0; k < m; ++k)
        for (int j = 0; j < m; ++j)
            for (int i = 0; i < m; ++i)
                for (auto& pm : perm) {
                    int c[3] = {i, j, k};
                    Element e;
                    e.id_ = (int)M.elements.size();
                    int v[4];
                    v[0] = pid(c[0], c[1], c[2]);
                    for (int s = 0; s < 3; ++s) { c[pm[s]]++; v[s + 1] = pid(c[0], c[1], c[2]); }
                    for (int s = 0; s < 4; ++s) e.p[s] = nullptr;
                    M.elements.push_back(e);
                    for (int s = 0; s < 4; ++s) M.elements.back().p[s] = reinterpret_cast<const Point*>((size_t)v[s]);   // fixed up below
                }
    for (Element& e : M.elements)
        for (int s = 0; s < 4; ++s) e.p[s] = &M.points[(size_t)e.p[s]];
    return M;
}

} // namespace mock

// ---- the flattening of INTEGRATION.md section 1, verbatim in shape ----
template <class MeshT> std::vector<double> flatten_points(MeshT const& mesh) {
    std::vector<double> xyz(3 * mesh.numPoints());
    for (auto const& elt : mesh.elements) {
        auto G = elt.G();
        for (int v = 0; v < 4; ++v)
            for (int c = 0; c < 3; ++c) xyz[3 * (size_t)elt.point(v).id() + c] = G[c * 4 + v];
    }
    return xyz;
}
template <class MeshT> std::vector<int32_t> flatten_connectivity(MeshT const& mesh) {
    std::vector<int32_t> conn;
    for (auto const& elt : mesh.elements)
        for (int v = 0; v < 4; ++v) conn.push_back((int32_t)elt.point(v).id());
    return conn;
}
template <class MeshT, class DofT> std::vector<int32_t> flatten_l2g(MeshT const& mesh, DofT const& dof) {
    std::vector<int32_t> l2g;
    for (auto const& elt : mesh.elements)
        for (int g : dof.localToGlobalIndices(elt.id())) l2g.push_back((int32_t)g);
    return l2g;
}
static std::vector<int64_t> to64(std::vector<size_t> const& v) { return std::vector<int64_t>(v.begin(), v.end()); }

#define ENGINE(call)                                                                 \
    do {                                                                             \
        int rc__ = (call);                                                           \
        if (rc__) {                                                                  \
            std::fprintf(stderr, "%s -> %d: %s\n", #call, rc__, feelgpu_last_error(ctx)); \
            return 3;                                                                \
        }                                                                            \
    } while (0)

int main(int argc, char** argv)
{
    if (argc != 3) { std::fprintf(stderr, "usage: adapter_mock <m> <out>\n"); return 2; }
    const int m = std::atoi(argv[1]);
    mock::Mesh mesh = mock::unitCube(m);
    mock::DofTable dof(mesh);
    mock::GraphCSR graph(mesh, dof);
    mock::IM im;
    mock::PreCompute pc;

    feelgpu_context_t ctx = nullptr;
    ENGINE(feelgpu_context_create(0, &ctx));
    std::vector<double> xyz = flatten_points(mesh);
    std::vector<int32_t> conn = flatten_connectivity(mesh);
    feelgpu_mesh_t gm = nullptr;
    ENGINE(feelgpu_mesh_create(ctx, (int64_t)mesh.numPoints(), xyz.data(), (int64_t)mesh.numElements(), conn.data(), &gm));
    std::vector<int32_t> l2g = flatten_l2g(mesh, dof);
    std::vector<int64_t> p2c = to64(dof.mapGlobalProcessToGlobalCluster());
    feelgpu_space_t sp = nullptr;
    ENGINE(feelgpu_space_create(ctx, gm, 1, 1, l2g.data(), (int64_t)dof.nLocalDofWithGhost(), (int64_t)dof.nLocalDofWithoutGhost(),
                                p2c.data(), &sp));
    std::vector<int64_t> ia = to64(graph.ia()), ja = to64(graph.ja());
    feelgpu_pattern_t pat = nullptr;
    ENGINE(feelgpu_pattern_create(ctx, sp, sp, (int64_t)ia.size() - 1, ia.data(), ja.data(), 0, &pat));

    feelgpu_form2_desc d2{};
    d2.kind = FEELGPU_FORM2_STIFFNESS;                 // classify<gradt(u)*trans(grad(v))>
    d2.coeff.kind = FEELGPU_COEFF_CONST;
    d2.coeff.value[0] = 1.0;
    d2.tables.nq = 1;
    d2.tables.w = im.weights.data();
    d2.tables.xi = im.points.data();
    d2.tables.n = 4;
    d2.tables.phi = pc.phi.data();
    d2.tables.dphi = pc.grad.data();
    d2.algorithm = FEELGPU_ALGO_AUTO;
    std::vector<double> a(ja.size(), -1.0);            // stands for MatSeqAIJGetArray(petsc_mat(form), &a)
    ENGINE(feelgpu_form2_assemble_host(pat, &d2, 0, a.data()));

    feelgpu_form1_desc d1{};
    d1.kind = FEELGPU_FORM1_SOURCE;                    // classify<cst(2.)*id(v)>
    d1.coeff.kind = FEELGPU_COEFF_CONST;
    d1.coeff.value[0] = 2.0;
    d1.tables = d2.tables;
    std::vector<double> b(dof.nLocalDofWithGhost(), -1.0);   // VecGetArray
    ENGINE(feelgpu_form1_assemble_host(sp, &d1, 0, b.data()));

    FILE* f = std::fopen(argv[2], "wb");
    if (!f) return 2;
    const int64_t hdr[4] = {(int64_t)mesh.numPoints(), (int64_t)mesh.numElements(), (int64_t)ia.size() - 1, (int64_t)ja.size()};
    std::fwrite(hdr, 8, 4, f);
    std::fwrite(xyz.data(), 8, xyz.size(), f);
    std::fwrite(conn.data(), 4, conn.size(), f);
    std::fwrite(l2g.data(), 4, l2g.size(), f);
    std::fwrite(ia.data(), 8, ia.size(), f);
    std::fwrite(ja.data(), 8, ja.size(), f);
    std::fwrite(a.data(), 8, a.size(), f);
    std::fwrite(b.data(), 8, b.size(), f);
    std::fclose(f);
    feelgpu_pattern_destroy(pat);
    feelgpu_space_destroy(sp);
    feelgpu_mesh_destroy(gm);
    feelgpu_context_destroy(ctx);
    std::printf("adapter_mock ok: %zu elements, %zu dofs, %zu nonzeros\n", mesh.numElements(), dof.nLocalDofWithGhost(), ja.size());
    return 0;
}
