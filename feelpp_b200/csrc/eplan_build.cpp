// eplan_build.cpp -- host-side construction of the entity-task plan (see eplan.h).
#include "eplan.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace {

// f(k0, k1, t) on contiguous pieces of [0, n), one per thread
template <class F>
void parallel_for(int64_t n, int nthreads, F f, int64_t serial_below = 65536)
{
    nthreads = (int)std::min<int64_t>(nthreads, n);
    if (nthreads <= 1 || n < serial_below) { f((int64_t)0, n, 0); return; }
    std::vector<std::thread> th;
    for (int t = 1; t < nthreads; ++t) th.emplace_back([=] { f(n * t / nthreads, n * (t + 1) / nthreads, t); });
    f((int64_t)0, n / nthreads, 0);
    for (auto& x : th) x.join();
}

// local entity numbering of the reference tetrahedron (feelmesh/simplexordering.hpp:283-291)
const int E2P[6][2] = {{1, 2}, {2, 0}, {0, 1}, {0, 3}, {1, 3}, {2, 3}};

inline int edge_of(int a, int b)
{
    for (int e = 0; e < 6; ++e)
        if ((E2P[e][0] == a && E2P[e][1] == b) || (E2P[e][0] == b && E2P[e][1] == a)) return e;
    return -1;
}

const EPlanSpec SPEC_P1 = {1, 4, 2,
                           {{1, 1, 1, {{0, 0}}, 1, {0}, 0},
                            {1, 2, 1, {{0, 1}}, 1, {1}, 0}}};
// P2 local dofs: vertices 0-3, edge dofs 4+e; re-labelled edges used below: (0,1) -> e2 -> dof 6, (1,2) -> e0 -> dof 4,
// (2,0) -> e1 -> dof 5, (2,3) -> e5 -> dof 9
const EPlanSpec SPEC_P2 = {2, 10, 6,
                           {{1, 1, 1, {{0, 0}}, 1, {0}, 0},
                            {1, 2, 2, {{0, 0}, {0, 1}}, 2, {1, 6}, 0},
                            {1, 3, 2, {{0, 1}, {0, 2}}, 1, {4}, 0},
                            {2, 2, 3, {{0, 0}, {0, 1}, {1, 1}}, 3, {0, 1, 6}, 6},
                            {2, 3, 5, {{0, 0}, {0, 1}, {0, 2}, {1, 1}, {1, 2}}, 3, {2, 5, 4}, 6},
                            {2, 4, 4, {{0, 2}, {0, 3}, {1, 2}, {1, 3}}, 1, {9}, 6, 1}}};

// position of the grid point (x,y,z), 10 bits each, along the 3-D Hilbert curve (Skilling's transpose algorithm): consecutive
// curve positions are face-adjacent cells, so a run of rows is a compact patch (a Morton run can jump across the domain)
inline uint32_t hilbert3(uint32_t x, uint32_t y, uint32_t z)
{
    uint32_t X[3] = {x & 0x3FF, y & 0x3FF, z & 0x3FF};
    const uint32_t M = 1u << 9;
    for (uint32_t Q = M; Q > 1; Q >>= 1) {       // inverse undo
        const uint32_t P = Q - 1;
        for (int i = 0; i < 3; ++i) {
            if (X[i] & Q) X[0] ^= P;
            else { const uint32_t t = (X[0] ^ X[i]) & P; X[0] ^= t; X[i] ^= t; }
        }
    }
    for (int i = 1; i < 3; ++i) X[i] ^= X[i - 1];   // Gray encode
    uint32_t t = 0;
    for (uint32_t Q = M; Q > 1; Q >>= 1)
        if (X[2] & Q) t ^= Q - 1;
    for (int i = 0; i < 3; ++i) X[i] ^= t;
    uint32_t h = 0;                                  // interleave, X[0] most significant
    for (int b = 9; b >= 0; --b)
        for (int i = 0; i < 3; ++i) h = (h << 1) | ((X[i] >> b) & 1);
    return h;
}

inline uint32_t spread3(uint32_t x)
{
    x &= 0x3FF;
    x = (x | (x << 16)) & 0x30000FF;
    x = (x | (x << 8)) & 0x300F00F;
    x = (x | (x << 4)) & 0x30C30C3;
    x = (x | (x << 2)) & 0x9249249;
    return x;
}

struct Entry {                      // one (task, element) incidence of a row
    uint32_t type;
    int32_t k1, k2;                 // global vertex ids completing sigma (ascending), -1 if unused
    int64_t q;                      // unit index (element, local dof)
    uint8_t p[4];                   // re-labelling: new local vertex k = original local vertex p[k]
    bool operator<(const Entry& o) const
    {
        if (type != o.type) return type < o.type;
        if (k1 != o.k1) return k1 < o.k1;
        if (k2 != o.k2) return k2 < o.k2;
        return q < o.q;
    }
};

struct Task {
    uint8_t type;
    uint16_t niter;
    uint32_t first;                 // into the per-CTA iteration word array
    uint16_t out[EPLAN_MAXOUT];
    uint8_t aux;                    // tasks done in phase A: original local edge of the row dof
};

struct CtaSeg {
    int64_t nb, ne;                 // node range along the curve
    uint32_t vert_off, visit_off;
    uint16_t nverts, nvisit;
    uint32_t nout, npad = 0;
};

struct Local {                      // per-thread output, concatenated at the end
    std::vector<EPlanChunk> chunks;
    std::vector<uint32_t> blob;
    std::vector<EPlanRow> rows;
    struct PerCta { uint32_t chunk0, nchunk, row0, nrows, q0, nq, needzero; };
    std::vector<PerCta> per;
    int64_t ntasks = 0, niters = 0, niters_padded = 0;
    double wf = 0, wf_ideal = 0, a1 = 0, a1_before = 0;
    std::string error;
};

struct Builder {
    const EPlanSpec* spec;
    int order, ncomp, n, nd;
    int64_t nElt, nNode;
    const int32_t* l2g;
    const double* xyz;
    const int32_t* conn;
    const int64_t* ia;
    const uint16_t* pos;
    std::vector<int64_t> nptr;      // node -> units
    std::vector<int32_t> uelt;
    std::vector<uint8_t> uloc;
    std::vector<uint16_t> uvis;     // visit slot of every unit inside its CTA
    std::vector<int32_t> order_;    // nodes along the curve
    bool use_hilbert = true;

    inline int64_t node_of(int64_t e, int i) const { return l2g[e * nd + i] / ncomp; }
    inline int row_len(int64_t node) const { return (int)((ia[node * ncomp + 1] - ia[node * ncomp]) / ncomp); }

    // units of every node in ascending element order (every task sums its elements in mesh order).  Counting and filling run
    // on all threads with atomic cursors; the few units of a node are then put back in element order (an element holds a
    // node once, so the order is unique and the result does not depend on the thread count).
    void node_units(int nthreads)
    {
        nptr.assign(nNode + 1, 0);
        int64_t* cnt = nptr.data() + 1;
        parallel_for(nElt, nthreads, [&](int64_t e0, int64_t e1, int) {
            for (int64_t e = e0; e < e1; ++e)
                for (int i = 0; i < n; ++i) __atomic_fetch_add(&cnt[node_of(e, i)], (int64_t)1, __ATOMIC_RELAXED);
        });
        for (int64_t k = 0; k < nNode; ++k) nptr[k + 1] += nptr[k];
        uelt.resize(nptr[nNode]);
        uloc.resize(nptr[nNode]);
        std::vector<int64_t> cur(nptr.begin(), nptr.end() - 1);
        parallel_for(nElt, nthreads, [&](int64_t e0, int64_t e1, int) {
            for (int64_t e = e0; e < e1; ++e)
                for (int i = 0; i < n; ++i) {
                    const int64_t q = __atomic_fetch_add(&cur[node_of(e, i)], (int64_t)1, __ATOMIC_RELAXED);
                    uelt[q] = (int32_t)e;
                    uloc[q] = (uint8_t)i;
                }
        });
        parallel_for(nNode, nthreads, [&](int64_t k0, int64_t k1, int) {
            for (int64_t k = k0; k < k1; ++k)
                for (int64_t a = nptr[k] + 1; a < nptr[k + 1]; ++a) {      // insertion sort by element
                    const int32_t e = uelt[a];
                    const uint8_t l = uloc[a];
                    int64_t b = a;
                    for (; b > nptr[k] && uelt[b - 1] > e; --b) { uelt[b] = uelt[b - 1]; uloc[b] = uloc[b - 1]; }
                    uelt[b] = e;
                    uloc[b] = l;
                }
        });
    }

    void curve_order(int nthreads)
    {
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        {
            std::vector<double> plo((size_t)std::max(nthreads, 1) * 3, 1e300), phi((size_t)std::max(nthreads, 1) * 3, -1e300);
            parallel_for(nElt, nthreads, [&](int64_t e0, int64_t e1, int t) {
                double l[3] = {1e300, 1e300, 1e300}, h[3] = {-1e300, -1e300, -1e300};
                for (int64_t e = e0; e < e1; ++e)
                    for (int a = 0; a < 4; ++a)
                        for (int c = 0; c < 3; ++c) {
                            const double x = xyz[3 * (int64_t)conn[4 * e + a] + c];
                            l[c] = std::min(l[c], x);
                            h[c] = std::max(h[c], x);
                        }
                for (int c = 0; c < 3; ++c) { plo[(size_t)t * 3 + c] = l[c]; phi[(size_t)t * 3 + c] = h[c]; }
            });
            for (size_t t = 0; t < plo.size() / 3; ++t)
                for (int c = 0; c < 3; ++c) { lo[c] = std::min(lo[c], plo[t * 3 + c]); hi[c] = std::max(hi[c], phi[t * 3 + c]); }
        }
        // one scale for the three axes: the curve's cells stay cubes on a partition whose bounding box is not one
        double sc[3], ext = 0.0;
        for (int c = 0; c < 3; ++c) ext = std::max(ext, hi[c] - lo[c]);
        for (int c = 0; c < 3; ++c) sc[c] = ext > 0.0 ? 1023.999 / ext : 0.0;
        std::vector<uint64_t> key(nNode);
        parallel_for(nNode, nthreads, [&](int64_t k0, int64_t k1, int) {
            for (int64_t k = k0; k < k1; ++k) {
                if (nptr[k] == nptr[k + 1]) { key[k] = ((uint64_t)0x3FFFFFFF << 32) | (uint32_t)k; continue; }
                const int64_t e = uelt[nptr[k]];
                const int i = uloc[nptr[k]];
                int v[2], nv = 1;
                if (i < 4) v[0] = i;
                else { v[0] = E2P[i - 4][0]; v[1] = E2P[i - 4][1]; nv = 2; }
                double p[3] = {0, 0, 0};
                for (int a = 0; a < nv; ++a)
                    for (int c = 0; c < 3; ++c) p[c] += xyz[3 * (int64_t)conn[4 * e + v[a]] + c];
                uint32_t m = 0;
                if (use_hilbert)
                    m = hilbert3((uint32_t)((p[0] / nv - lo[0]) * sc[0]), (uint32_t)((p[1] / nv - lo[1]) * sc[1]),
                                 (uint32_t)((p[2] / nv - lo[2]) * sc[2]));
                else
                    for (int c = 0; c < 3; ++c) m |= spread3((uint32_t)((p[c] / nv - lo[c]) * sc[c])) << c;
                key[k] = ((uint64_t)m << 32) | (uint32_t)k;
            }
        });
        {   // the keys are distinct (node id in the low word): chunks sorted by the threads, then merged pairwise
            int nch = 1;
            while (nch * 2 <= nthreads && (int64_t)key.size() / (nch * 2) > 65536) nch *= 2;
            std::vector<size_t> cut((size_t)nch + 1);
            for (int c = 0; c <= nch; ++c) cut[c] = key.size() * (size_t)c / nch;
            {
                std::vector<std::thread> th;
                for (int c = 1; c < nch; ++c) th.emplace_back([&, c] { std::sort(key.begin() + cut[c], key.begin() + cut[c + 1]); });
                std::sort(key.begin() + cut[0], key.begin() + cut[1]);
                for (auto& x : th) x.join();
            }
            for (int w = 1; w < nch; w *= 2) {
                std::vector<std::thread> th;
                for (int c = 0; c + w < nch; c += 2 * w)
                    th.emplace_back([&, c, w] {
                        std::inplace_merge(key.begin() + cut[c], key.begin() + cut[c + w], key.begin() + cut[std::min(c + 2 * w, nch)]);
                    });
                for (auto& x : th) x.join();
            }
        }
        order_.resize(nNode);
        parallel_for(nNode, nthreads, [&](int64_t k0, int64_t k1, int) {
            for (int64_t k = k0; k < k1; ++k) order_[k] = (int32_t)(key[k] & 0xFFFFFFFFu);
        });
    }

    // entries of one unit (row node, element e, local dof i)
    void unit_entries(int64_t q, std::vector<Entry>& out) const
    {
        const int64_t e = uelt[q];
        const int i = uloc[q];
        const int32_t* c = conn + 4 * e;
        Entry en;
        en.q = q;
        if (i < 4) {
            int o[3], no = 0;
            for (int v = 0; v < 4; ++v)
                if (v != i) o[no++] = v;
            std::sort(o, o + 3, [&](int a, int b) { return c[a] < c[b]; });
            en.type = 0; en.k1 = en.k2 = -1;
            en.p[0] = (uint8_t)i; en.p[1] = (uint8_t)o[0]; en.p[2] = (uint8_t)o[1]; en.p[3] = (uint8_t)o[2];
            out.push_back(en);
            for (int b = 0; b < 3; ++b) {                       // (1,2): sigma = {v, o[b]}
                en.type = 1; en.k1 = c[o[b]]; en.k2 = -1;
                en.p[1] = (uint8_t)o[b];
                int r = 2;
                for (int x = 0; x < 3; ++x)
                    if (x != b) en.p[r++] = (uint8_t)o[x];
                out.push_back(en);
            }
            if (spec->ntypes > 2)
                for (int d = 0; d < 3; ++d) {                   // (1,3): sigma = {v} U (o minus o[d])
                    int b0 = d == 0 ? 1 : 0, b1 = d == 2 ? 1 : 2;
                    en.type = 2; en.k1 = c[o[b0]]; en.k2 = c[o[b1]];
                    en.p[1] = (uint8_t)o[b0]; en.p[2] = (uint8_t)o[b1]; en.p[3] = (uint8_t)o[d];
                    out.push_back(en);
                }
        } else {
            int r0 = E2P[i - 4][0], r1 = E2P[i - 4][1];
            if (c[r0] > c[r1]) std::swap(r0, r1);
            int o[2], no = 0;
            for (int v = 0; v < 4; ++v)
                if (v != r0 && v != r1) o[no++] = v;
            if (c[o[0]] > c[o[1]]) std::swap(o[0], o[1]);
            en.p[0] = (uint8_t)r0; en.p[1] = (uint8_t)r1;
            en.type = 3; en.k1 = en.k2 = -1; en.p[2] = (uint8_t)o[0]; en.p[3] = (uint8_t)o[1];
            out.push_back(en);
            en.type = 4; en.k1 = c[o[0]];
            out.push_back(en);
            en.k1 = c[o[1]]; en.p[2] = (uint8_t)o[1]; en.p[3] = (uint8_t)o[0];
            out.push_back(en);
            en.type = 5; en.k1 = c[o[0]]; en.k2 = c[o[1]]; en.p[2] = (uint8_t)o[0]; en.p[3] = (uint8_t)o[1];
            out.push_back(en);
        }
    }

    // original local dof of the re-labelled local dof d under the re-labelling p
    inline int orig_dof(int d, const uint8_t* p) const
    {
        if (d < 4) return p[d];
        const int e = d - 4;
        return 4 + edge_of(p[E2P[e][0]], p[E2P[e][1]]);
    }
};

// modelled shared-memory wavefronts of one chunk's S loads: per iteration, per load, per half-warp the conflict degree
// is the largest number of distinct addresses that fall into one 8-byte bank pair (bank pair = visit slot mod 16)
void model_chunk(const EPlanType& T, const std::vector<uint32_t>& words, int niter, double& wf, double& ideal)
{
    for (int t = 0; t < niter; ++t)
        for (int k = 0; k < T.nld; ++k)
            for (int half = 0; half < 2; ++half) {
                uint32_t seen[16][16];
                int cnt[16];
                for (int b = 0; b < 16; ++b) cnt[b] = 0;
                for (int l = half * 16; l < half * 16 + 16; ++l) {
                    const uint32_t w = words[(size_t)t * 32 + l];
                    const uint32_t slot = w & 0x3FF, idx = (w >> (EPLAN_IDX_SHIFT + 4 * k)) & 15;
                    const uint32_t addr = idx * EPLAN_VS + slot, b = slot & 15;
                    bool dup = false;
                    for (int s = 0; s < cnt[b]; ++s)
                        if (seen[b][s] == addr) dup = true;
                    if (!dup) seen[b][cnt[b]++] = addr;
                }
                int deg = 1;
                for (int b = 0; b < 16; ++b) deg = std::max(deg, cnt[b]);
                wf += deg;
                ideal += 1;
            }
}

// Phase A1 reads the coordinates of corner k of 16 consecutive visits at a time (one half-warp, thread = visit slot); the
// read is served in as many passes as the fullest bank pair (vertex slot mod 16) holds distinct vertices.  The slot of a
// vertex is free: deal the vertices to the 16 bank pairs (16 slots each), most used first, each to the pair that raises the
// fewest group maxima, then let every vertex reconsider once.  Returns the modelled passes; rewrites verts[] and the
// corner bytes of the visit words; nverts becomes the highest slot in use + 1.
double level_vertex_banks(uint32_t nv, uint32_t np, uint32_t* vw, int vstride, int32_t* verts, uint16_t& nverts,
                          double& before)
{
    const uint32_t ngrp = (nv + 15) / 16 * 4;
    std::vector<uint32_t> gptr(np + 1, 0), gl;
    std::vector<uint8_t> cnt((size_t)ngrp * 16, 0), gmax(ngrp, 0);
    auto group_of = [&](uint32_t v, int k) { return (v / 16) * 4 + (uint32_t)k; };
    // distinct (vertex, group) pairs
    std::vector<uint32_t> pairs;
    pairs.reserve((size_t)nv * 4);
    for (uint32_t v = 0; v < nv; ++v)
        for (int k = 0; k < 4; ++k) pairs.push_back((((vw[(size_t)v * vstride] >> (8 * k)) & 255u) << 16) | group_of(v, k));
    std::sort(pairs.begin(), pairs.end());
    pairs.erase(std::unique(pairs.begin(), pairs.end()), pairs.end());
    for (uint32_t pr : pairs) gptr[(pr >> 16) + 1]++;
    for (uint32_t u = 0; u < np; ++u) gptr[u + 1] += gptr[u];
    gl.resize(pairs.size());
    for (size_t i = 0; i < pairs.size(); ++i) gl[i] = pairs[i] & 0xFFFF;   // sorted by vertex: already in CSR order
    auto passes = [&]() {
        double t = 0;
        for (uint32_t g = 0; g < ngrp; ++g) t += std::max<int>(1, gmax[g]);
        return t;
    };
    // the numbering found by the packing pass
    for (uint32_t u = 0; u < np; ++u)
        for (uint32_t q = gptr[u]; q < gptr[u + 1]; ++q) {
            uint8_t& c = cnt[(size_t)gl[q] * 16 + (u & 15)];
            gmax[gl[q]] = std::max<uint8_t>(gmax[gl[q]], ++c);
        }
    before = passes();
    if (np <= 16) return before;
    std::fill(cnt.begin(), cnt.end(), 0);
    std::fill(gmax.begin(), gmax.end(), 0);
    std::vector<uint32_t> order(np);
    for (uint32_t u = 0; u < np; ++u) order[u] = u;
    std::stable_sort(order.begin(), order.end(),
                     [&](uint32_t x, uint32_t y) { return gptr[x + 1] - gptr[x] > gptr[y + 1] - gptr[y]; });
    std::vector<int> bank(np, -1);
    int load[16] = {0};
    auto best_bank = [&](uint32_t u) {
        int bb = -1;
        long bcost = 0;
        for (int b = 0; b < 16; ++b) {
            if (load[b] >= EPLAN_VERTS / 16) continue;
            long raise = 0, crowd = 0;
            for (uint32_t q = gptr[u]; q < gptr[u + 1]; ++q) {
                const uint8_t c = cnt[(size_t)gl[q] * 16 + b];
                raise += c + 1 > gmax[gl[q]];
                crowd += c;
            }
            const long cost = (raise << 20) + (crowd << 8) + load[b];
            if (bb < 0 || cost < bcost) { bb = b; bcost = cost; }
        }
        return bb;
    };
    auto put = [&](uint32_t u, int b) {
        bank[u] = b;
        ++load[b];
        for (uint32_t q = gptr[u]; q < gptr[u + 1]; ++q) {
            uint8_t& c = cnt[(size_t)gl[q] * 16 + b];
            gmax[gl[q]] = std::max<uint8_t>(gmax[gl[q]], ++c);
        }
    };
    auto take = [&](uint32_t u) {
        const int b = bank[u];
        --load[b];
        for (uint32_t q = gptr[u]; q < gptr[u + 1]; ++q) {
            const uint32_t g = gl[q];
            --cnt[(size_t)g * 16 + b];
            uint8_t m = 0;
            for (int x = 0; x < 16; ++x) m = std::max(m, cnt[(size_t)g * 16 + x]);
            gmax[g] = m;
        }
    };
    for (uint32_t u : order) put(u, best_bank(u));
    for (int pass = 0; pass < 2; ++pass)
        for (uint32_t u : order) {
            take(u);
            put(u, best_bank(u));
        }
    // slots: the j-th vertex of bank pair b sits at b + 16 j
    int fill[16] = {0};
    std::vector<uint8_t> slot(np);
    std::vector<int32_t> old(verts, verts + np);
    uint32_t hi = 0;
    std::fill(verts, verts + EPLAN_VERTS, 0);
    for (uint32_t u = 0; u < np; ++u) {
        const uint32_t sl = (uint32_t)bank[u] + 16u * (uint32_t)fill[bank[u]]++;
        slot[u] = (uint8_t)sl;
        verts[sl] = old[u];
        hi = std::max(hi, sl + 1);
    }
    nverts = (uint16_t)hi;
    for (uint32_t v = 0; v < nv; ++v) {
        uint32_t& w = vw[(size_t)v * vstride];
        uint32_t r = 0;
        for (int k = 0; k < 4; ++k) r |= (uint32_t)slot[(w >> (8 * k)) & 255u] << (8 * k);
        w = r;
    }
    return passes();
}

} // namespace

const EPlanSpec* eplan_spec(int order)
{
    if (order == 1) return &SPEC_P1;
    if (order == 2) return &SPEC_P2;
    return nullptr;
}

bool eplan_build_host(int order, int ncomp, int omax, int64_t nElt, const int32_t* l2g, int nd, int64_t nPts,
                      const double* xyz, const int32_t* conn, int64_t nRows, const int64_t* ia, const uint16_t* pos,
                      int nthreads, EPlanHost& plan)
{
    // FEELGPU_EPLAN_TIMING=1: phase times of the builder on stderr
    const bool timing = std::getenv("FEELGPU_EPLAN_TIMING") != nullptr;
    auto tprev = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[eplan build] %-28s %8.3f s\n", what, std::chrono::duration<double>(now - tprev).count());
        tprev = now;
    };
    Builder B;
    B.spec = eplan_spec(order);
    if (!B.spec) { plan.error = "entity plan: order not covered"; return false; }
    B.order = order;
    B.ncomp = ncomp;
    B.n = B.spec->n;
    B.nd = nd;
    B.nElt = nElt;
    B.nNode = nRows / ncomp;
    B.l2g = l2g;
    B.xyz = xyz;
    B.conn = conn;
    B.ia = ia;
    B.pos = pos;
    if (nRows % ncomp) { plan.error = "entity plan: rows not a multiple of ncomp"; return false; }
    if (omax < 512 || omax > 32767) { plan.error = "entity plan: bad out-buffer size"; return false; }
    if (nthreads <= 0) {
        const char* bt = std::getenv("FEELGPU_BUILD_THREADS");
        nthreads = bt && std::atoi(bt) > 0 ? std::atoi(bt) : (int)std::max(1u, std::thread::hardware_concurrency());
    }
    nthreads = std::min(nthreads, 64);
    {
        std::vector<int> bad((size_t)nthreads, 0);
        parallel_for(nElt, nthreads, [&](int64_t e0, int64_t e1, int t) {
            int b = 0;
            for (int64_t e = e0; e < e1 && !b; ++e) {
                for (int i = 0; i < B.n; ++i)
                    if (l2g[e * nd + i] < 0 || l2g[e * nd + i] >= nRows) b = 1;
                for (int a = 0; a < 4; ++a)
                    for (int c = a + 1; c < 4; ++c)
                        if (conn[4 * e + a] == conn[4 * e + c]) b = std::max(b, 2);
            }
            bad[t] = b;
        });
        for (int b : bad) {
            if (b == 1) { plan.error = "entity plan: a test dof has no CSR row"; return false; }
            if (b == 2) { plan.error = "entity plan: degenerate element"; return false; }
        }
    }
    lap("input checks");
    B.node_units(nthreads);
    lap("node -> element units");
    if (const char* cv = std::getenv("FEELGPU_EPLAN_CURVE")) B.use_hilbert = std::strcmp(cv, "morton") != 0;
    B.curve_order(nthreads);
    lap("space-filling curve order");

    // ---- CTA packing along the curve: consecutive rows while visits, unique vertices and outputs fit ----
    // The curve is cut into pieces of PACK_PIECE rows that are packed independently (a piece ends with a CTA that may be short:
    // 1 in ~240 for P2); the pieces are a property of the mesh, not of the thread count, so the plan does not depend on it.
    std::vector<CtaSeg> segs;
    B.uvis.resize(B.uelt.size());
    plan.verts.clear();
    plan.visits.clear();
    const int vstride = order == 2 ? 4 : 1;   // words per visit: vertex bytes [+ 6 u16 row-block slots of the tet-type outputs]
    {
        constexpr int64_t PACK_PIECE = 1 << 16;
        struct Piece { std::vector<CtaSeg> segs; std::vector<int32_t> verts; std::vector<uint32_t> visits; std::string error; };
        const int64_t npiece = std::max<int64_t>(1, (B.nNode + PACK_PIECE - 1) / PACK_PIECE);
        std::vector<Piece> pieces((size_t)npiece);
        int64_t next_piece = 0;
        auto pack = [&]() {
            std::vector<int32_t> estamp, eslot, pstamp, pslot, newpts;
            int32_t id = 0;                       // stamp of the CTA being packed (unique inside this thread's arrays)
            for (;;) {
                const int64_t pi = __atomic_fetch_add(&next_piece, (int64_t)1, __ATOMIC_RELAXED);
                if (pi >= npiece) break;
                if (estamp.empty()) {
                    estamp.assign((size_t)nElt, -1);
                    eslot.assign((size_t)nElt, 0);
                    pstamp.assign((size_t)nPts, -1);
                    pslot.assign((size_t)nPts, 0);
                }
                Piece& P = pieces[(size_t)pi];
                const int64_t kend = std::min(B.nNode, (pi + 1) * PACK_PIECE);
                int64_t k = pi * PACK_PIECE;
                while (k < kend) {
                    CtaSeg sg;
                    sg.nb = k;
                    sg.vert_off = (uint32_t)P.verts.size();       // inside the piece; rebased below
                    sg.visit_off = (uint32_t)P.visits.size();
                    ++id;
                    uint32_t nv = 0, np = 0, nout = 1;      // slot 0 of the out buffer is the trash cell
                    int64_t rows = 0;
                    for (; k < kend; ++k) {
                        const int64_t node = B.order_[k];
                        uint32_t addv = 0, addp = 0;
                        newpts.clear();
                        for (int64_t q = B.nptr[node]; q < B.nptr[node + 1]; ++q) {
                            const int32_t e = B.uelt[q];
                            if (estamp[e] != id) {
                                ++addv;
                                for (int a = 0; a < 4; ++a) {
                                    const int32_t pt = conn[4 * e + a];
                                    if (pstamp[pt] != id && std::find(newpts.begin(), newpts.end(), pt) == newpts.end()) newpts.push_back(pt);
                                }
                            }
                        }
                        addp = (uint32_t)newpts.size();
                        const uint32_t len = (uint32_t)B.row_len(node);
                        if (rows > 0 && (nv + addv > EPLAN_VS - 1 || np + addp > EPLAN_VERTS ||
                                         nout + len > (uint32_t)omax - (B.ncomp == 1 ? EPLAN_PADRES : 0) || rows >= EPLAN_ROWS))
                            break;
                        if (addv > EPLAN_VS - 1 || addp > EPLAN_VERTS || 1 + len > (uint32_t)omax) {
                            P.error = "entity plan: a single row exceeds the CTA limits (elements around a node / row length)";
                            return;
                        }
                        for (int64_t q = B.nptr[node]; q < B.nptr[node + 1]; ++q) {
                            const int32_t e = B.uelt[q];
                            if (estamp[e] != id) {
                                estamp[e] = id;
                                eslot[e] = (int32_t)nv++;
                                uint32_t w = 0;
                                for (int a = 0; a < 4; ++a) {
                                    const int32_t pt = conn[4 * e + a];
                                    if (pstamp[pt] != id) {
                                        pstamp[pt] = id;
                                        pslot[pt] = (int32_t)np++;
                                        P.verts.push_back(pt);
                                    }
                                    w |= (uint32_t)pslot[pt] << (8 * a);
                                }
                                P.visits.push_back(w);
                                for (int z = 1; z < vstride; ++z) P.visits.push_back(0u);   // filled per CTA below
                            }
                            B.uvis[q] = (uint16_t)eslot[e];
                        }
                        nout += len;
                        ++rows;
                    }
                    sg.ne = k;
                    sg.nverts = (uint16_t)np;
                    P.verts.resize((size_t)sg.vert_off + EPLAN_VERTS, 0);   // fixed stride: the kernel loads ids before it has the CTA header
                    sg.nvisit = (uint16_t)nv;
                    sg.nout = nout;
                    P.segs.push_back(sg);
                }
            }
        };
        {
            // every packing thread holds stamp arrays over the elements and points: a handful of threads is enough here
            const int npack = (int)std::max<int64_t>(1, std::min<int64_t>(std::min(nthreads, 16), npiece));
            std::vector<std::thread> th;
            for (int t = 1; t < npack; ++t) th.emplace_back(pack);
            pack();
            for (auto& x : th) x.join();
        }
        size_t nsg = 0, nvt = 0, nvs = 0;
        std::vector<size_t> sgb((size_t)npiece), vtb((size_t)npiece), vsb((size_t)npiece);
        for (int64_t pi = 0; pi < npiece; ++pi) {
            const Piece& P = pieces[(size_t)pi];
            if (!P.error.empty()) { plan.error = P.error; return false; }
            sgb[(size_t)pi] = nsg; vtb[(size_t)pi] = nvt; vsb[(size_t)pi] = nvs;
            nsg += P.segs.size(); nvt += P.verts.size(); nvs += P.visits.size();
        }
        if (nvt > 0xFFFFFFFFull || nvs > 0xFFFFFFFFull) { plan.error = "entity plan: visit offsets exceed 32 bits"; return false; }
        segs.resize(nsg);
        plan.verts.resize(nvt);
        plan.visits.resize(nvs);
        parallel_for(npiece, nthreads, [&](int64_t p0, int64_t p1, int) {
            for (int64_t pi = p0; pi < p1; ++pi) {
                Piece& P = pieces[(size_t)pi];
                for (size_t c = 0; c < P.segs.size(); ++c) {
                    CtaSeg sg = P.segs[c];
                    sg.vert_off += (uint32_t)vtb[(size_t)pi];
                    sg.visit_off += (uint32_t)vsb[(size_t)pi];
                    segs[sgb[(size_t)pi] + c] = sg;
                }
                if (!P.verts.empty()) std::memcpy(plan.verts.data() + vtb[(size_t)pi], P.verts.data(), P.verts.size() * sizeof(int32_t));
                if (!P.visits.empty()) std::memcpy(plan.visits.data() + vsb[(size_t)pi], P.visits.data(), P.visits.size() * sizeof(uint32_t));
                P = Piece();
            }
        }, 4);
    }

    lap("CTA packing");
    // ---- tasks and chunks, in parallel over CTAs ----
    const int64_t nCta = (int64_t)segs.size();
    const EPlanSpec& S = *B.spec;
    std::vector<Local> locals(nthreads);
    auto work = [&](int t) {
        Local& L = locals[t];
        std::vector<Entry> ent;
        std::vector<Task> tasks[EPLAN_MAXTYPES];
        std::vector<uint32_t> iterw;          // iteration words of the CTA's tasks
        std::vector<uint32_t> cw;             // one chunk's words
        std::vector<uint8_t> bankcnt;
        uint8_t outcnt[EPLAN_MAXOUT * 16];
        std::vector<uint8_t> covered;
        std::vector<uint32_t> usage, sorted, perm, vtmp;
        std::vector<int32_t> cnodes;
        std::vector<uint32_t> cell[256];      // tasks with <= 2 elements by (bank pair of element 0, of element 1)
        uint32_t csz[256];                    // their sizes, both orientations (csz[a * 16 + b] == csz[b * 16 + a])
        std::vector<uint32_t> cell1[16];
        int cnt2[32];
        std::vector<uint32_t> pool, lanes;    // remaining tasks of a type; tasks of every chunk lane
        std::vector<uint16_t> tmask, zmask;   // window search: bank pairs of a task; bank pairs unused per iteration
        struct ChunkTmp { uint8_t type, niter; uint32_t first, count; int cost; };
        std::vector<ChunkTmp> ctmp;
        std::vector<uint32_t> noff, segoff, pads;
        double sect[6] = {0, 0, 0, 0, 0, 0};
        auto sprev = std::chrono::steady_clock::now();
        auto mark = [&](int k) {
            if (!timing || t != 0) return;
            auto now = std::chrono::steady_clock::now();
            sect[k] += std::chrono::duration<double>(now - sprev).count();
            sprev = now;
        };
        for (int64_t c = nCta * t / nthreads; c < nCta * (t + 1) / nthreads; ++c) {
            mark(4);
            CtaSeg& sg = segs[c];
            for (int ty = 0; ty < S.ntypes; ++ty) tasks[ty].clear();
            iterw.clear();
            Local::PerCta pc;
            pc.row0 = (uint32_t)L.rows.size();
            // rows in CSR order: runs that are contiguous in memory are written as one segment (scalar spaces; pieces of
            // <= 256 values).  Phase C moves a segment with one bulk copy (TMA) when the segment starts at the same 16-byte
            // phase in the row block as in the CSR array, i.e. when slot and base have the same parity; the order of the
            // segments inside the row block is free, so they are dealt greedily to keep the parities equal (a segment of odd
            // length flips the phase); the few that cannot be matched are copied by the warps.
            cnodes.assign(B.order_.begin() + sg.nb, B.order_.begin() + sg.ne);
            std::sort(cnodes.begin(), cnodes.end(), [&](int32_t x, int32_t y) { return B.ia[(int64_t)x * B.ncomp] < B.ia[(int64_t)y * B.ncomp]; });
            noff.assign(cnodes.size(), 0);
            {
                struct SegTmp { int64_t base; uint32_t len, k0, k1; };
                std::vector<SegTmp> st;
                for (size_t k = 0; k < cnodes.size(); ++k) {
                    const int64_t node = cnodes[k];
                    const uint32_t len = (uint32_t)B.row_len(node);
                    const int64_t base = B.ia[node * B.ncomp];
                    if (B.ncomp == 1 && !st.empty() && st.back().base + (int64_t)st.back().len == base &&
                        st.back().len + len <= 256 && base < plan.nomerge_from) {
                        st.back().len += len;
                        st.back().k1 = (uint32_t)k + 1;
                    } else
                        st.push_back(SegTmp{base, len, (uint32_t)k, (uint32_t)k + 1});
                }
                // a walk on the two phases: a segment of even length stays, one of odd length crosses over; when the phase has
                // nothing left a one-slot pad crosses (EPLAN_PADRES slots are held back for that by the packing pass)
                std::vector<uint32_t> queue[2][2];      // [base parity][length parity]
                for (size_t i = 0; i < st.size(); ++i) queue[st[i].base & 1][st[i].len & 1].push_back((uint32_t)i);
                size_t qh[2][2] = {{0, 0}, {0, 0}};
                auto left = [&](int b, int l) { return queue[b][l].size() - qh[b][l]; };
                uint32_t off = 1;      // slot 0 of the row block is the trash cell
                segoff.assign(st.size(), 0);
                pads.clear();
                for (size_t placed = 0; placed < st.size(); ++placed) {
                    int p = (int)(off & 1), b, l;
                    if (left(p, 0)) { b = p; l = 0; }
                    else if (left(p, 1)) { b = p; l = 1; }
                    else if (B.ncomp == 1 && sg.nout + (uint32_t)pads.size() + 1 <= (uint32_t)omax) {
                        pads.push_back(off++);           // (the values still to place fit behind the pad)
                        --placed;
                        continue;
                    } else { b = p ^ 1; l = left(b, 1) ? 1 : 0; }
                    const uint32_t i = queue[b][l][qh[b][l]++];
                    segoff[i] = off;
                    off += st[i].len;
                }
                sg.nout = off;
                sg.npad = (uint32_t)pads.size();
                for (size_t i = 0; i < st.size(); ++i) {
                    EPlanRow row;
                    row.base = st[i].base;
                    row.off = segoff[i];
                    row.len = st[i].len;
                    L.rows.push_back(row);
                    uint32_t o = segoff[i];
                    for (uint32_t k = st[i].k0; k < st[i].k1; ++k) {
                        noff[k] = o;
                        o += (uint32_t)B.row_len(cnodes[k]);
                    }
                }
            }
            for (size_t k = 0; k < cnodes.size(); ++k) {
                const int64_t node = cnodes[k];
                const uint32_t len = (uint32_t)B.row_len(node);
                const uint32_t rowoff = noff[k];
                ent.clear();
                for (int64_t q = B.nptr[node]; q < B.nptr[node + 1]; ++q) B.unit_entries(q, ent);
                std::sort(ent.begin(), ent.end());
                size_t g0 = 0;
                while (g0 < ent.size()) {
                    size_t g1 = g0 + 1;
                    while (g1 < ent.size() && ent[g1].type == ent[g0].type && ent[g1].k1 == ent[g0].k1 && ent[g1].k2 == ent[g0].k2) ++g1;
                    const EPlanType& T = S.type[ent[g0].type];
                    Task tk;
                    tk.type = (uint8_t)ent[g0].type;
                    tk.niter = (uint16_t)(g1 - g0);
                    tk.first = (uint32_t)iterw.size();
                    for (int o = 0; o < EPLAN_MAXOUT; ++o) tk.out[o] = 0;
                    tk.aux = (uint8_t)(B.uloc[ent[g0].q] >= 4 ? B.uloc[ent[g0].q] - 4 : 0);
                    if (g1 - g0 > 255) { L.error = "entity plan: more than 255 elements around an entity"; return; }
                    for (size_t g = g0; g < g1; ++g) {
                        const Entry& en = ent[g];
                        const int64_t e = B.uelt[en.q];
                        const int i = B.uloc[en.q];
                        uint32_t w = B.uvis[en.q];
                        for (int kk = 0; kk < T.nld; ++kk)
                            w |= (uint32_t)eplan_sidx(en.p[T.ld[kk][0]], en.p[T.ld[kk][1]]) << (EPLAN_IDX_SHIFT + 4 * kk);
                        iterw.push_back(w);
                        if (B.orig_dof(T.rowdof, en.p) != i) { L.error = "entity plan: re-labelling does not map the row dof"; return; }
                        for (int o = 0; o < T.nout; ++o) {
                            const int j = B.orig_dof(T.outdof[o], en.p);
                            const uint32_t p = B.pos[(e * B.n + i) * B.n + j] / (uint32_t)B.ncomp;
                            if (p >= len) { L.error = "entity plan: position outside the row"; return; }
                            const uint16_t slot = (uint16_t)(rowoff + p);
                            if (g == g0) tk.out[o] = slot;
                            else if (tk.out[o] != slot) { L.error = "entity plan: elements of one task disagree on an output column"; return; }
                        }
                    }
                    tasks[tk.type].push_back(tk);
                    g0 = g1;
                }
            }
            pc.nrows = (uint32_t)L.rows.size() - pc.row0;
            mark(0);
            // ---- renumber the visit slots: every half-warp wants 16 distinct bank pairs (slot mod 16) per iteration, which
            // needs the load-weighted use of the 16 residues to be level: deal the slots, most used first, in snake order
            {
                const uint32_t nv = sg.nvisit;
                usage.assign(nv, 0);
                for (int ty = 0; ty < S.ntypes; ++ty)
                    for (const Task& tk : tasks[ty])
                        for (int it = 0; it < tk.niter; ++it)
                            usage[iterw[tk.first + it] & 0x3FF] += (uint32_t)(S.type[ty].phaseA ? 0 : S.type[ty].nld);
                sorted.resize(nv);
                for (uint32_t v = 0; v < nv; ++v) sorted[v] = v;
                std::stable_sort(sorted.begin(), sorted.end(), [&](uint32_t x, uint32_t y) { return usage[x] > usage[y]; });
                perm.resize(nv);
                const uint32_t full = nv / 16 * 16;
                for (uint32_t i2 = 0; i2 < nv; ++i2) {
                    const uint32_t g = i2 / 16, j = i2 % 16;
                    perm[sorted[i2]] = (i2 < full && (g & 1)) ? g * 16 + 15 - j : i2;
                }
                for (uint32_t& w : iterw) w = (w & ~0x3FFu) | perm[w & 0x3FF];
                uint32_t* vw = plan.visits.data() + sg.visit_off;      // this CTA's segment: no other thread touches it
                vtmp.assign(vw, vw + (size_t)nv * vstride);
                for (uint32_t v = 0; v < nv; ++v)
                    for (int z = 0; z < vstride; ++z) vw[(size_t)perm[v] * vstride + z] = vtmp[(size_t)v * vstride + z];
                // outputs whose simplex is the whole tet have one contribution: the thread that computes the element's S in
                // phase A stores them straight into the row block (slot per original local edge of the row; 0 = not ours)
                for (int ty = 0; ty < S.ntypes; ++ty) {
                    if (!S.type[ty].phaseA) continue;
                    for (const Task& tk : tasks[ty]) {
                        const uint32_t slot = iterw[tk.first] & 0x3FF;
                        vw[(size_t)slot * vstride + 1 + tk.aux / 2] |= (uint32_t)tk.out[0] << (16 * (tk.aux & 1));
                    }
                }
            }
            mark(1);
            if (!getenv("FEELGPU_EPLAN_NO_VERTEX_BANKS")) {
                double before = 0;
                L.a1 += level_vertex_banks(sg.nvisit, sg.nverts, plan.visits.data() + sg.visit_off, vstride,
                                           plan.verts.data() + sg.vert_off, sg.nverts, before);
                L.a1_before += before;
            }
            mark(2);
            // ---- chunks: 32 tasks of one type, longest first ----
            // All S loads of one iteration hit the bank pair (visit slot mod 16); a half-warp is served in one pass only
            // if its 16 lanes hit distinct bank pairs.  Which tasks share a half-warp and the order in which a task walks
            // its elements are both free (the latter only fixes the summation order): lanes are filled one by one with the
            // candidate (from a window of the remaining tasks) whose greedy element order adds the fewest bank conflicts.
            ctmp.clear();
            lanes.clear();
            for (int ty = 0; ty < S.ntypes; ++ty) {
                mark(5);
                std::vector<Task>& tv = tasks[ty];
                L.ntasks += (int64_t)tv.size();
                if (S.type[ty].phaseA) continue;
                std::stable_sort(tv.begin(), tv.end(), [](const Task& a, const Task& b) { return a.niter > b.niter; });
                const int nto = S.type[ty].nout;
                size_t nlong = 0;
                while (nlong < tv.size() && tv[nlong].niter > 2) ++nlong;
                // ---- tasks with 3 or more elements: window search ----
                pool.resize(nlong);
                for (size_t k = 0; k < nlong; ++k) pool[k] = (uint32_t)k;
                // bank pairs a task touches (the set does not change when its elements are re-ordered): a candidate none of
                // whose bank pairs is free in iteration `it` pays at least one conflict there -- a lower bound that discards a
                // third of the candidates before their greedy order is worked out (same choice, less work)
                tmask.resize(nlong);
                for (size_t k = 0; k < nlong; ++k) {
                    uint16_t mk = 0;
                    for (int j = 0; j < tv[k].niter; ++j) mk |= (uint16_t)(1u << (iterw[tv[k].first + j] & 15));
                    tmask[k] = mk;
                }
                size_t head = 0;
                while (head < pool.size()) {
                    ChunkTmp ch;
                    ch.type = (uint8_t)ty;
                    {   // the window below displaces tasks by up to 40 places: bring the longest of them to the front
                        size_t mx = head;
                        for (size_t c = head + 1; c < std::min(pool.size(), head + 128); ++c)
                            if (tv[pool[c]].niter > tv[pool[mx]].niter) mx = c;
                        std::swap(pool[head], pool[mx]);
                    }
                    ch.niter = (uint8_t)tv[pool[head]].niter;
                    ch.first = (uint32_t)lanes.size();
                    ch.count = 0;
                    ch.cost = 45 + ch.niter * (5 * S.type[ty].nld + 8) + nto * (S.type[ty].nld + 3);
                    for (int l = 0; l < 32 && head < pool.size(); ++l) {
                        if ((l & 15) == 0) {
                            bankcnt.assign((size_t)ch.niter * 16, 0);
                            zmask.assign((size_t)ch.niter, (uint16_t)0xFFFF);      // bank pairs still unused per iteration
                            for (int z = 0; z < EPLAN_MAXOUT * 16; ++z) outcnt[z] = 0;
                        }
                        const int hl = l & 15;
                        size_t K = l == 0 ? 1 : 8 + 4 * hl;
                        K = std::min(K, pool.size() - head);
                        size_t best = head;
                        int bestcost = 1 << 30;
                        for (size_t c = head; c < head + K; ++c) {
                            const Task& tk = tv[pool[c]];
                            const int nt = tk.niter;
                            if (nt > ch.niter) continue;
                            if (bestcost < (1 << 30)) {
                                const uint16_t mk = tmask[pool[c]];
                                int lb = 0;
                                for (int it = 0; it < nt; ++it) lb += (mk & zmask[it]) == 0;
                                if (lb * S.type[ty].nld >= bestcost) continue;
                            }
                            const uint32_t* src = &iterw[tk.first];
                            int cost = 0;
                            uint8_t res[256];
                            for (int j = 0; j < nt; ++j) res[j] = (uint8_t)(src[j] & 15);
                            for (int it = 0; it < nt; ++it) {
                                int bj = it;
                                uint8_t bc = 255;
                                for (int j = it; j < nt; ++j) {
                                    const uint8_t c2 = bankcnt[(size_t)it * 16 + res[j]];
                                    if (c2 < bc) { bc = c2; bj = j; if (!c2) break; }
                                }
                                std::swap(res[it], res[bj]);
                                cost += bc * S.type[ty].nld;
                                if (cost >= bestcost) break;   // costs only grow: this candidate cannot win any more
                            }
                            if (cost >= bestcost) continue;
                            for (int o = 0; o < nto; ++o) cost += outcnt[o * 16 + (tk.out[o] & 15)];   // stores into the row block
                            if (cost < bestcost) { bestcost = cost; best = c; if (!cost) break; }
                        }
                        std::swap(pool[head], pool[best]);
                        // commit: greedy element order of the chosen task
                        const Task& tk = tv[pool[head]];
                        const int nt = tk.niter;
                        uint32_t* src = &iterw[tk.first];
                        for (int it = 0; it < nt; ++it) {
                            int bj = it;
                            uint8_t bc = 255;
                            for (int j = it; j < nt; ++j) {
                                const uint8_t c2 = bankcnt[(size_t)it * 16 + (src[j] & 15)];
                                if (c2 < bc) { bc = c2; bj = j; if (!c2) break; }
                            }
                            std::swap(src[it], src[bj]);
                            bankcnt[(size_t)it * 16 + (src[it] & 15)]++;
                            zmask[it] &= (uint16_t)~(1u << (src[it] & 15));
                        }
                        for (int o = 0; o < nto; ++o) outcnt[o * 16 + (tk.out[o] & 15)]++;
                        lanes.push_back(pool[head]);
                        ++head;
                        ++ch.count;
                    }
                    ctmp.push_back(ch);
                }
                mark(3);
                // ---- tasks with two elements, then with one: buckets by bank pair.  A lane takes a task whose bank pairs
                // are still free in both iterations of its half-warp (the order of the two elements is chosen to fit).
                {
                    for (auto& cvec : cell) cvec.clear();
                    for (auto& cvec : cell1) cvec.clear();
                    size_t rem2 = 0, rem1 = 0;
                    uint16_t ne[16];      // ne[a] bit b: the cell of bank pairs {a, b} holds tasks (the scan below skips the rest)
                    for (int z = 0; z < 16; ++z) ne[z] = 0;
                    for (int z = 0; z < 256; ++z) csz[z] = 0;
                    for (size_t k = nlong; k < tv.size(); ++k) {
                        const uint32_t* src = &iterw[tv[k].first];
                        const int r0 = src[0] & 15;
                        if (tv[k].niter == 2) {
                            const int r1 = src[1] & 15;
                            cell[std::min(r0, r1) * 16 + std::max(r0, r1)].push_back((uint32_t)k);
                            ++csz[r0 * 16 + r1];
                            if (r0 != r1) ++csz[r1 * 16 + r0];
                            ne[r0] |= (uint16_t)(1u << r1);
                            ne[r1] |= (uint16_t)(1u << r0);
                            ++rem2;
                        } else {
                            cell1[r0].push_back((uint32_t)k);
                            ++rem1;
                        }
                    }
                    while (rem2 + rem1) {
                        const int cn = rem2 ? 2 : 1;                 // a chunk of two-element tasks is topped up with one-element tasks
                        ChunkTmp ch;
                        ch.type = (uint8_t)ty;
                        ch.niter = (uint8_t)cn;
                        ch.first = (uint32_t)lanes.size();
                        ch.count = 0;
                        ch.cost = 45 + cn * (5 * S.type[ty].nld + 8) + nto * (S.type[ty].nld + 3);
                        for (int l = 0; l < 32 && rem2 + rem1; ++l) {
                            if ((l & 15) == 0) {
                                for (int z = 0; z < 32; ++z) cnt2[z] = 0;
                                for (int z = 0; z < EPLAN_MAXOUT * 16; ++z) outcnt[z] = 0;
                            }
                            const int nt = rem2 ? 2 : 1;
                            // best cell: fewest S-load conflicts, then the fullest (keeps the buckets level)
                            int ba = 0, bb = 0, bcost = 1 << 30;
                            size_t bsize = 0;
                            std::vector<uint32_t>* bcv = nullptr;
                            if (nt == 2) {
                                // cells whose bank pairs are unused in both iterations cost nothing: the fullest of them wins
                                // (first in (a2, b2) order among equals, as in the full scan below)
                                uint32_t FA = 0, FB = 0, bs = 0;
                                for (int z = 0; z < 16; ++z) { FA |= (uint32_t)(cnt2[z] == 0) << z; FB |= (uint32_t)(cnt2[16 + z] == 0) << z; }
                                for (uint32_t ma = FA; ma; ma &= ma - 1) {
                                    const int a2 = __builtin_ctz(ma);
                                    const uint32_t* row = csz + a2 * 16;
                                    for (uint32_t mb = ne[a2] & FB; mb; mb &= mb - 1) {
                                        const int b2 = __builtin_ctz(mb);
                                        if (row[b2] > bs) { bs = row[b2]; ba = a2; bb = b2; }
                                    }
                                }
                                if (bs) {
                                    bcost = 0;
                                    bsize = bs;
                                    bcv = &cell[std::min(ba, bb) * 16 + std::max(ba, bb)];
                                }
                            }
                            if (nt == 2 && !bcv) {
                                // (for a given a2 only the b2 of least count can win: the others are skipped through the masks
                                // of the bank pairs by their count in iteration 1 -- same choice as the full scan, a third of the work)
                                uint32_t bycnt[18];
                                for (int z = 0; z < 18; ++z) bycnt[z] = 0;
                                for (int b2 = 0; b2 < 16; ++b2) bycnt[std::min(cnt2[16 + b2], 17)] |= 1u << b2;
                                for (int a2 = 0; a2 < 16; ++a2) {    // a2: bank pair of iteration 0, b2: of iteration 1
                                    const uint32_t na = ne[a2];
                                    if (!na || cnt2[a2] > bcost) continue;
                                    int c1 = 0;
                                    while (c1 < 18 && !(na & bycnt[c1])) ++c1;
                                    const int cost = cnt2[a2] + c1;
                                    if (c1 >= 17) {                   // (never with 16 lanes per half-warp; kept exact all the same)
                                        for (uint32_t mb = na; mb; mb &= mb - 1) {
                                            const int b2 = __builtin_ctz(mb);
                                            std::vector<uint32_t>& cv2 = cell[std::min(a2, b2) * 16 + std::max(a2, b2)];
                                            const int co = cnt2[a2] + cnt2[16 + b2];
                                            if (co < bcost || (co == bcost && cv2.size() > bsize)) { bcost = co; bsize = cv2.size(); bcv = &cv2; ba = a2; bb = b2; }
                                        }
                                        continue;
                                    }
                                    if (cost > bcost) continue;
                                    for (uint32_t mb = na & bycnt[c1]; mb; mb &= mb - 1) {
                                        const int b2 = __builtin_ctz(mb);
                                        std::vector<uint32_t>& cv2 = cell[std::min(a2, b2) * 16 + std::max(a2, b2)];
                                        const size_t sz = cv2.size();
                                        if (cost < bcost || (cost == bcost && sz > bsize)) { bcost = cost; bsize = sz; bcv = &cv2; ba = a2; bb = b2; }
                                    }
                                }
                            } else if (nt == 1) {
                                for (int a2 = 0; a2 < 16; ++a2) {
                                    const size_t sz = cell1[a2].size();
                                    if (!sz) continue;
                                    const int cost = cnt2[a2];
                                    if (cost < bcost || (cost == bcost && sz > bsize)) { bcost = cost; bsize = sz; bcv = &cell1[a2]; ba = a2; }
                                }
                            }
                            // inside the cell: the task whose stores into the row block collide the least
                            std::vector<uint32_t>& cv = *bcv;
                            size_t bk = 0;
                            int bo = 1 << 30;
                            for (size_t k = 0; k < cv.size() && k < 12; ++k) {
                                int oc = 0;
                                for (int o = 0; o < nto; ++o) oc += outcnt[o * 16 + (tv[cv[k]].out[o] & 15)];
                                if (oc < bo) { bo = oc; bk = k; if (!oc) break; }
                            }
                            const uint32_t ti = cv[bk];
                            cv[bk] = cv.back();
                            cv.pop_back();
                            if (nt == 2) {
                                --csz[ba * 16 + bb];
                                if (ba != bb) --csz[bb * 16 + ba];
                            }
                            if (nt == 2 && cv.empty()) {
                                ne[ba] &= (uint16_t)~(1u << bb);
                                ne[bb] &= (uint16_t)~(1u << ba);
                            }
                            if (nt == 2) --rem2; else --rem1;
                            const Task& tk = tv[ti];
                            uint32_t* src = &iterw[tk.first];
                            if (nt == 2 && (int)(src[0] & 15) != ba) std::swap(src[0], src[1]);
                            cnt2[ba]++;
                            if (nt == 2) cnt2[16 + bb]++;
                            for (int o = 0; o < nto; ++o) outcnt[o * 16 + (tk.out[o] & 15)]++;
                            lanes.push_back(ti);
                            ++ch.count;
                        }
                        ctmp.push_back(ch);
                    }
                }
            }
            mark(5);
            std::stable_sort(ctmp.begin(), ctmp.end(), [](const ChunkTmp& a, const ChunkTmp& b) { return a.cost > b.cost; });
            // does every slot of the row block receive a value?  (always, unless the graph holds entries no element produces)
            covered.assign(sg.nout, 0);
            for (int ty = 0; ty < S.ntypes; ++ty)
                for (const Task& tk : tasks[ty])
                    for (int o = 0; o < S.type[ty].nout; ++o) covered[tk.out[o]] = 1;
            for (uint32_t z : pads) covered[z] = 1;
            pc.needzero = 0;
            for (uint32_t z = 1; z < sg.nout; ++z)
                if (!covered[z]) pc.needzero = 1;
            pc.chunk0 = (uint32_t)L.chunks.size();
            pc.q0 = (uint32_t)(L.blob.size() / 4);
            // Short chunks (<= 2 iterations) cost more in per-chunk overhead than in work: up to EPLAN_NSUB of them (same
            // type and length) are run back to back under one header.  Data of a chunk: words u32 [nsub*niter][32], then
            // row-block slots u16 [nsub*nout][32].
            for (size_t c0 = 0; c0 < ctmp.size();) {
                const ChunkTmp& ch = ctmp[c0];
                const EPlanType& T = S.type[ch.type];
                const std::vector<Task>& tv = tasks[ch.type];
                size_t nsub = 1;
                if (ch.niter <= 2)
                    while (nsub < EPLAN_NSUB && c0 + nsub < ctmp.size() && ctmp[c0 + nsub].type == ch.type && ctmp[c0 + nsub].niter == ch.niter) ++nsub;
                EPlanChunk hc;
                hc.q0 = (uint32_t)(L.blob.size() / 4);
                hc.type = ch.type;
                hc.niter = ch.niter;
                hc.pad = (uint16_t)nsub;
                L.chunks.push_back(hc);
                for (size_t j = 0; j < nsub; ++j) {
                    const ChunkTmp& cj = ctmp[c0 + j];
                    cw.assign((size_t)cj.niter * 32, (uint32_t)(EPLAN_VS - 1));   // padding: the all-zero visit record
                    for (uint32_t l = 0; l < cj.count; ++l) {
                        const Task& tk = tv[lanes[cj.first + l]];
                        for (int it = 0; it < tk.niter; ++it) cw[(size_t)it * 32 + l] = iterw[tk.first + it];
                        L.niters += tk.niter;
                    }
                    model_chunk(T, cw, cj.niter, L.wf, L.wf_ideal);
                    // as the kernel reads them: the visit slot already times 8 (bits 3-11), and the words of two
                    // consecutive iterations side by side per lane ([ceil(niter/2)][32] pairs: one 64-bit load brings both;
                    // an odd chunk ends with a word of the all-zero record)
                    const size_t npair = ((size_t)cj.niter + 1) / 2;
                    L.niters_padded += (int64_t)npair * 64;
                    const size_t wb = L.blob.size();
                    L.blob.resize(wb + npair * 64, (uint32_t)(EPLAN_VS - 1) << 3);
                    for (int it = 0; it < cj.niter; ++it)
                        for (int l = 0; l < 32; ++l) {
                            const uint32_t w = cw[(size_t)it * 32 + l];
                            L.blob[wb + ((size_t)(it >> 1) * 32 + l) * 2 + (it & 1)] = ((w & 0x1FFu) << 3) | (w & 0xFFFFF000u);
                        }
                }
                for (size_t j = 0; j < nsub; ++j) {
                    const ChunkTmp& cj = ctmp[c0 + j];
                    const size_t ob = L.blob.size();
                    L.blob.resize(ob + (size_t)T.nout * 16, 0u);              // u16 [nout][32], 0 = trash cell
                    uint16_t* op = reinterpret_cast<uint16_t*>(L.blob.data() + ob);
                    for (uint32_t l = 0; l < cj.count; ++l)
                        for (int o = 0; o < T.nout; ++o) op[o * 32 + l] = tv[lanes[cj.first + l]].out[o];
                }
                c0 += nsub;
            }
            pc.nchunk = (uint32_t)L.chunks.size() - pc.chunk0;
            pc.nq = (uint32_t)(L.blob.size() / 4) - pc.q0;
            if (pc.nchunk > 65535 || pc.nq > 65535) { L.error = "entity plan: too many chunks in one CTA"; return; }
            L.per.push_back(pc);
            mark(4);
        }
        if (timing && t == 0)
            std::fprintf(stderr, "[eplan build] thread 0: rows/tasks %.3f  slots %.3f  vertex banks %.3f  window search %.3f  buckets %.3f  emission %.3f s\n",
                         sect[0], sect[1], sect[2], sect[3], sect[5], sect[4]);
    };
    {
        std::vector<std::thread> th;
        for (int t = 1; t < nthreads; ++t) th.emplace_back(work, t);
        work(0);
        for (auto& x : th) x.join();
    }
    for (auto& L : locals)
        if (!L.error.empty()) { plan.error = L.error; return false; }
    lap("tasks and chunks (threads)");

    // ---- concatenate ----
    plan.order = order;
    plan.n = B.n;
    plan.ncomp = ncomp;
    plan.omax = omax;
    size_t nq = 0, nch = 0, nrw = 0;
    for (auto& L : locals) { nq += L.blob.size() / 4; nch += L.chunks.size(); nrw += L.rows.size(); }
    if (nq + 64 > 0xFFFFFFFFull) { plan.error = "entity plan: blob offsets exceed 32 bits"; return false; }
    plan.ctas.resize(nCta);
    plan.chunks.resize(nch);
    plan.rows.resize(nrw);
    plan.blob.resize(nq * 4 + 64);             // (uninitialised: EPlanNoInit; the threads below touch their pieces first)
    std::memset(plan.blob.data() + nq * 4, 0, 64 * sizeof(uint32_t));
    {
        // thread t built the CTAs [nCta t / nthreads, nCta (t+1) / nthreads): its pieces go behind those of the threads before it
        std::vector<size_t> cb((size_t)nthreads), rb((size_t)nthreads), qb((size_t)nthreads);
        size_t c0 = 0, r0 = 0, q0 = 0;
        for (int t = 0; t < nthreads; ++t) {
            cb[t] = c0; rb[t] = r0; qb[t] = q0;
            c0 += locals[t].chunks.size(); r0 += locals[t].rows.size(); q0 += locals[t].blob.size() / 4;
        }
        auto copy = [&](int t) {
            Local& L = locals[t];
            for (size_t k = 0; k < L.chunks.size(); ++k) {
                EPlanChunk h = L.chunks[k];
                h.q0 += (uint32_t)qb[t];
                plan.chunks[cb[t] + k] = h;
            }
            if (!L.rows.empty()) std::memcpy(plan.rows.data() + rb[t], L.rows.data(), L.rows.size() * sizeof(EPlanRow));
            if (!L.blob.empty()) std::memcpy(plan.blob.data() + qb[t] * 4, L.blob.data(), L.blob.size() * sizeof(uint32_t));
            int64_t c = nCta * t / nthreads;
            for (auto& pc : L.per) {
                EPlanCta& ct = plan.ctas[c];
                const CtaSeg& sg = segs[c];
                ct.vert_off = sg.vert_off;
                ct.visit_off = sg.visit_off;
                ct.chunk_off = (uint32_t)cb[t] + pc.chunk0;
                ct.row_off = (uint32_t)rb[t] + pc.row0;
                ct.nverts = sg.nverts;
                ct.nvisit = sg.nvisit;
                ct.nchunk = (uint16_t)pc.nchunk;
                ct.nrows = (uint16_t)pc.nrows;
                ct.nout = (uint16_t)(sg.nout | (pc.needzero ? 0x8000u : 0u));
                ct.blob_q0 = (uint32_t)qb[t] + pc.q0;
                ct.blob_nq = (uint16_t)pc.nq;
                ++c;
            }
        };
        std::vector<std::thread> th;
        for (int t = 1; t < nthreads; ++t) th.emplace_back(copy, t);
        copy(0);
        for (auto& x : th) x.join();
        for (int64_t c = 0; c < nCta; ++c) plan.nouts += segs[c].nout - 1 - segs[c].npad;
        for (auto& L : locals) {
            plan.ntasks += L.ntasks;
            plan.niters += L.niters;
            plan.niters_padded += L.niters_padded;
            plan.lds_wavefronts += L.wf;
            plan.lds_ideal += L.wf_ideal;
            plan.a1_passes += L.a1;
            plan.a1_passes_before += L.a1_before;
            L = Local();
        }
    }
    plan.nvisits = (int64_t)plan.visits.size() / vstride;
    lap("concatenate");
    return true;
}
