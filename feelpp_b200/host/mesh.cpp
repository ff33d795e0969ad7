#include "mesh.hpp"

#include <algorithm>
#include <random>
#include <stdexcept>

namespace feelpp_b200 {

mesh_ptrtype unitCube(int m, double jitter, bool shuffle)
{
    if (m < 1) throw std::invalid_argument("unitCube: m must be >= 1");
    auto mesh = std::make_shared<Mesh>();
    const int64_t np1 = m + 1;
    const int64_t nPts = np1 * np1 * np1;
    if (nPts >= ((int64_t)1 << 31)) throw std::invalid_argument("unitCube: too many points for int32 ids");
    const double h = 1.0 / m;
    mesh->xyz.resize(3 * nPts);
    std::mt19937_64 rng(20261104ULL);
    std::uniform_real_distribution<double> U(-1.0, 1.0);
    for (int64_t k = 0; k < np1; ++k)
        for (int64_t j = 0; j < np1; ++j)
            for (int64_t i = 0; i < np1; ++i) {
                const int64_t id = (k * np1 + j) * np1 + i;
                double x = i * h, y = j * h, z = k * h;
                if (jitter > 0.0) {
                    // draw for every vertex so that the stream does not depend on which ones are interior
                    double dx = U(rng), dy = U(rng), dz = U(rng);
                    const bool interior = i > 0 && i < m && j > 0 && j < m && k > 0 && k < m;
                    if (interior) {
                        x += jitter * h * dx;
                        y += jitter * h * dy;
                        z += jitter * h * dz;
                    }
                }
                mesh->xyz[3 * id + 0] = x;
                mesh->xyz[3 * id + 1] = y;
                mesh->xyz[3 * id + 2] = z;
            }
    // Kuhn split: one tet per permutation (p0,p1,p2) of the axes: v0, v0+e_p0, v0+e_p0+e_p1, v0+(1,1,1)
    static const int perms[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 0, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}};
    const int64_t nElt = (int64_t)m * m * m * 6;
    mesh->conn.resize(4 * nElt);
    int64_t e = 0;
    for (int64_t k = 0; k < m; ++k)
        for (int64_t j = 0; j < m; ++j)
            for (int64_t i = 0; i < m; ++i)
                for (int p = 0; p < 6; ++p, ++e) {
                    int64_t c[3] = {i, j, k};
                    int32_t v[4];
                    v[0] = (int32_t)((c[2] * np1 + c[1]) * np1 + c[0]);
                    c[perms[p][0]] += 1;
                    v[1] = (int32_t)((c[2] * np1 + c[1]) * np1 + c[0]);
                    c[perms[p][1]] += 1;
                    v[2] = (int32_t)((c[2] * np1 + c[1]) * np1 + c[0]);
                    c[perms[p][2]] += 1;
                    v[3] = (int32_t)((c[2] * np1 + c[1]) * np1 + c[0]);
                    // orientation of the unperturbed tet = parity of the permutation; make every element positive
                    const bool odd = (p == 1 || p == 2 || p == 5);
                    if (odd) std::swap(v[2], v[3]);
                    for (int a = 0; a < 4; ++a) mesh->conn[4 * e + a] = v[a];
                }
    if (shuffle) {
        std::mt19937_64 r2(7ULL);
        for (int64_t a = nElt - 1; a > 0; --a) {
            const int64_t b = (int64_t)(r2() % (uint64_t)(a + 1));
            for (int c = 0; c < 4; ++c) std::swap(mesh->conn[4 * a + c], mesh->conn[4 * b + c]);
        }
    }
    return mesh;
}

void boundaryFaces(int64_t nPts, int64_t nElt, const int32_t* conn, std::vector<int32_t>& elt, std::vector<int32_t>& face)
{
    static const int F2P[4][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}};
    // faces bucketed by their smallest point id (counting sort: one pass to count, one to place), then each small bucket
    // (about a dozen faces) is sorted by the two other ids: linear in the number of elements
    struct Rest { int32_t b, c; int64_t ef; };
    std::vector<int64_t> first((size_t)nPts + 1, 0);
    auto sorted3 = [&](int64_t e, int f, int32_t* v) {
        for (int k = 0; k < 3; ++k) {
            v[k] = conn[4 * e + F2P[f][k]];
            if (v[k] < 0 || v[k] >= nPts) throw std::invalid_argument("boundaryFaces: point id out of range");
        }
        if (v[0] > v[1]) std::swap(v[0], v[1]);
        if (v[1] > v[2]) std::swap(v[1], v[2]);
        if (v[0] > v[1]) std::swap(v[0], v[1]);
    };
    int32_t v[3];
    for (int64_t e = 0; e < nElt; ++e)
        for (int f = 0; f < 4; ++f) {
            sorted3(e, f, v);
            ++first[(size_t)v[0] + 1];
        }
    for (int64_t p = 0; p < nPts; ++p) first[(size_t)p + 1] += first[(size_t)p];
    std::vector<Rest> rest((size_t)nElt * 4);
    {
        std::vector<int64_t> at(first.begin(), first.end() - 1);
        for (int64_t e = 0; e < nElt; ++e)
            for (int f = 0; f < 4; ++f) {
                sorted3(e, f, v);
                rest[(size_t)at[(size_t)v[0]]++] = Rest{v[1], v[2], e * 4 + f};
            }
    }
    std::vector<uint8_t> single((size_t)nElt * 4, 0);
    for (int64_t p = 0; p < nPts; ++p) {
        Rest* b = rest.data() + first[(size_t)p];
        Rest* e = rest.data() + first[(size_t)p + 1];
        std::sort(b, e, [](Rest const& x, Rest const& y) { return x.b != y.b ? x.b < y.b : x.c < y.c; });
        for (Rest* i = b; i < e;) {
            Rest* k = i + 1;
            while (k < e && k->b == i->b && k->c == i->c) ++k;
            if (k == i + 1) single[(size_t)i->ef] = 1;
            i = k;
        }
    }
    elt.clear();
    face.clear();
    for (size_t k = 0; k < single.size(); ++k)
        if (single[k]) {
            elt.push_back((int32_t)(k / 4));
            face.push_back((int32_t)(k % 4));
        }
}

} // namespace feelpp_b200
