#include "doftable.hpp"

#include <algorithm>
#include <stdexcept>

namespace feelpp_b200 {

const int kEdgeToPoint[6][2] = {{1, 2}, {2, 0}, {0, 1}, {0, 3}, {1, 3}, {2, 3}};
const int kFaceToPoint[4][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}};

namespace {

// open-addressing hash: 128-bit key (three 32-bit ids packed in two words) -> first dof of the entity
struct EntityTable {
    struct Slot { uint64_t k0, k1; int64_t v; };
    std::vector<Slot> slots;
    uint64_t mask = 0;
    explicit EntityTable(size_t expected)
    {
        size_t cap = 16;
        while (cap < expected * 2) cap <<= 1;
        slots.assign(cap, Slot{~0ULL, ~0ULL, -1});
        mask = cap - 1;
    }
    static uint64_t mix(uint64_t a, uint64_t b)
    {
        uint64_t h = a * 0x9E3779B97F4A7C15ULL ^ (b + 0x7F4A7C15ULL) * 0xC2B2AE3D27D4EB4FULL;
        h ^= h >> 29;
        h *= 0xBF58476D1CE4E5B9ULL;
        h ^= h >> 32;
        return h;
    }
    // returns reference to value; inserted=true when the key is new
    int64_t& find_or_insert(uint64_t k0, uint64_t k1, bool& inserted)
    {
        uint64_t p = mix(k0, k1) & mask;
        for (;;) {
            Slot& s = slots[p];
            if (s.v < 0 && s.k0 == ~0ULL && s.k1 == ~0ULL) {
                s.k0 = k0;
                s.k1 = k1;
                inserted = true;
                return s.v;
            }
            if (s.k0 == k0 && s.k1 == k1) {
                inserted = false;
                return s.v;
            }
            p = (p + 1) & mask;
        }
    }
};

} // namespace

void DofTable::build(Mesh const& mesh, int order_, int ncomp_)
{
    if (order_ < 1 || order_ > 3) throw std::invalid_argument("DofTable: order must be 1, 2 or 3");
    if (ncomp_ < 1) throw std::invalid_argument("DofTable: ncomp must be >= 1");
    order = order_;
    ncomp = ncomp_;
    n = nLocalDofTet(order);
    nd = n * ncomp;
    const int64_t nElt = mesh.numElements(), nPts = mesh.numPoints();
    const int nPerEdge = order - 1;
    const int nPerFace = (order - 1) * (order - 2) / 2;
    l2g.assign((size_t)nElt * nd, -1);
    std::vector<int64_t> vertexDof(nPts, -1);
    // Euler-type bounds for a tet mesh: edges <= 6*nElt, faces <= 4*nElt; the tables grow from a tighter guess
    EntityTable edges(nPerEdge ? (size_t)(nElt * 1.3 + nPts + 64) : 1);
    EntityTable faces(nPerFace ? (size_t)(nElt * 2.2 + 64) : 1);
    int64_t next = 0;
    auto put = [&](int64_t e, int ldof, int64_t g) {
        for (int c = 0; c < ncomp; ++c) l2g[e * nd + n * c + ldof] = (int32_t)(g + c);
    };
    for (int64_t e = 0; e < nElt; ++e) {
        const int32_t* v = &mesh.conn[4 * e];
        int lc = 0;
        for (int i = 0; i < 4; ++i, ++lc) {
            int64_t& d = vertexDof[v[i]];
            if (d < 0) {
                d = next;
                next += ncomp;
            }
            put(e, lc, d);
        }
        if (nPerEdge)
            for (int ed = 0; ed < 6; ++ed) {
                const uint64_t i1 = (uint32_t)v[kEdgeToPoint[ed][0]], i2 = (uint32_t)v[kEdgeToPoint[ed][1]];
                const bool reverse = i1 > i2;
                bool ins;
                int64_t& base = edges.find_or_insert(std::min(i1, i2) << 32 | std::max(i1, i2), 0, ins);
                if (ins) {
                    // the dofs of a new edge are numbered in the order this element visits them (l = 0,1,..), and
                    // local dof l addresses edge slot l (or nPerEdge-1-l when reversed): keep base + visiting order
                    base = next;
                    next += (int64_t)ncomp * nPerEdge;
                    // first-touch order: slot visited first gets the smaller number
                    base |= (int64_t)(reverse ? 1 : 0) << 62;
                }
                const int64_t b0 = base & ~((int64_t)1 << 62);
                const bool firstReversed = (base >> 62) & 1;
                for (int l = 0; l < nPerEdge; ++l, ++lc) {
                    const int slot = reverse ? nPerEdge - 1 - l : l;            // key of the reference: edge*nPerEdge + slot
                    const int rank = firstReversed ? nPerEdge - 1 - slot : slot; // order in which slots were first inserted
                    put(e, lc, b0 + (int64_t)ncomp * rank);
                }
            }
        if (nPerFace)
            for (int f = 0; f < 4; ++f) {
                uint64_t a = (uint32_t)v[kFaceToPoint[f][0]], b = (uint32_t)v[kFaceToPoint[f][1]],
                         c = (uint32_t)v[kFaceToPoint[f][2]];
                if (a > b) std::swap(a, b);
                if (b > c) std::swap(b, c);
                if (a > b) std::swap(a, b);
                bool ins;
                int64_t& base = faces.find_or_insert(a << 32 | b, c, ins);
                if (ins) {
                    base = next;
                    next += (int64_t)ncomp * nPerFace;
                }
                for (int l = 0; l < nPerFace; ++l, ++lc) put(e, lc, base + (int64_t)ncomp * l);
            }
        // no volume dofs for k <= 3
    }
    if (next >= ((int64_t)1 << 31)) throw std::overflow_error("DofTable: more than 2^31 dofs");
    nDof = next;
}

} // namespace feelpp_b200
