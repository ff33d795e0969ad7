// rowplan_build.cpp -- host-side construction of the owner-computes plan (see rowplan.h).
#include <string>
#include "rowplan.h"

#include <algorithm>
#include <cstring>
#include <numeric>
#include <thread>

namespace {

// local entity numbering of the reference tetrahedron (feelmesh/simplexordering.hpp:283-291, 341-347)
const int E2P[6][2] = {{1, 2}, {2, 0}, {0, 1}, {0, 3}, {1, 3}, {2, 3}};
const int F2P[4][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}};

inline int nloc(int k) { return (k + 1) * (k + 2) * (k + 3) / 6; }

// vertices of the entity carrying local dof i (for the row's position on the space-filling curve)
int dof_entity_vertices(int order, int i, int v[3])
{
    if (i < 4) { v[0] = i; return 1; }
    const int npe = order - 1;
    if (i < 4 + 6 * npe) {
        const int e = (i - 4) / npe;
        v[0] = E2P[e][0];
        v[1] = E2P[e][1];
        return 2;
    }
    const int f = i - 4 - 6 * npe;
    v[0] = F2P[f][0];
    v[1] = F2P[f][1];
    v[2] = F2P[f][2];
    return 3;
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

} // namespace

void rowplan_rotation_tables(int order, RowPlanRot* rot, int canon[3])
{
    const int n = nloc(order), npe = order - 1;
    canon[0] = 0;
    canon[1] = npe ? 4 + 2 * npe : -1;       // first dof of local edge 2 = (0,1)
    canon[2] = order >= 3 ? 4 + 6 * npe + 3 : -1;  // dof of local face 3 = (0,1,2)
    for (int i = 0; i < n; ++i) {
        RowPlanRot& R = rot[i];
        int p[4];
        bool used[4] = {false, false, false, false};
        int np = 0;
        if (i < 4) {
            R.kind = 0;
            p[np++] = i;
        } else if (i < 4 + 6 * npe) {
            R.kind = 1;
            const int e = (i - 4) / npe, l = (i - 4) % npe;
            // dof l sits at the l-th interior node counted from the first vertex; the nodal set is symmetric, so the
            // dof nearest to the second vertex is the first dof of the reversed edge
            const bool fromSecond = (npe == 2 && l == 1);
            p[np++] = E2P[e][fromSecond ? 1 : 0];
            p[np++] = E2P[e][fromSecond ? 0 : 1];
        } else {
            R.kind = 2;
            const int f = i - 4 - 6 * npe;
            p[np++] = F2P[f][0];
            p[np++] = F2P[f][1];
            p[np++] = F2P[f][2];
        }
        for (int k = 0; k < np; ++k) used[p[k]] = true;
        for (int v = 0; v < 4; ++v)
            if (!used[v]) p[np++] = v;
        for (int k = 0; k < 4; ++k) R.perm[k] = (uint8_t)p[k];
        // rotated-local dof -> original-local dof
        for (int j = 0; j < 4; ++j) R.rmap[j] = (uint8_t)p[j];
        for (int e2 = 0; e2 < 6 && npe; ++e2) {
            const int a = p[E2P[e2][0]], b = p[E2P[e2][1]];
            int eo = -1;
            bool same = true;
            for (int e = 0; e < 6; ++e) {
                if (E2P[e][0] == a && E2P[e][1] == b) { eo = e; same = true; }
                if (E2P[e][0] == b && E2P[e][1] == a) { eo = e; same = false; }
            }
            for (int l = 0; l < npe; ++l) R.rmap[4 + e2 * npe + l] = (uint8_t)(4 + eo * npe + (same ? l : npe - 1 - l));
        }
        if (order >= 3)
            for (int f2 = 0; f2 < 4; ++f2) {
                int m = (1 << p[F2P[f2][0]]) | (1 << p[F2P[f2][1]]) | (1 << p[F2P[f2][2]]);
                for (int f = 0; f < 4; ++f)
                    if (((1 << F2P[f][0]) | (1 << F2P[f][1]) | (1 << F2P[f][2])) == m) R.rmap[4 + 6 * npe + f2] = (uint8_t)(4 + 6 * npe + f);
            }
    }
}

namespace {

struct Builder {
    int order, ncomp, n, nd;
    int64_t nElt, nNode;
    const int32_t* l2g;
    const double* xyz;
    const int32_t* conn;
    const int64_t* ia;
    const uint16_t* pos;
    RowPlanRot rot[20];
    int canon[3];
    std::vector<int64_t> nptr;     // node -> units
    std::vector<int32_t> uelt;
    std::vector<uint8_t> uloc;
    std::vector<int32_t> order_;   // nodes along the curve
    std::vector<uint8_t> kind;

    inline int64_t node_of(int64_t e, int i) const { return l2g[e * nd + i] / ncomp; }

    void node_units()
    {
        nptr.assign(nNode + 1, 0);
        for (int64_t e = 0; e < nElt; ++e)
            for (int i = 0; i < n; ++i) nptr[node_of(e, i) + 1]++;
        for (int64_t k = 0; k < nNode; ++k) nptr[k + 1] += nptr[k];
        uelt.resize(nptr[nNode]);
        uloc.resize(nptr[nNode]);
        std::vector<int64_t> cur(nptr.begin(), nptr.end() - 1);
        for (int64_t e = 0; e < nElt; ++e)          // ascending e: every row sums its elements in mesh order
            for (int i = 0; i < n; ++i) {
                int64_t q = cur[node_of(e, i)]++;
                uelt[q] = (int32_t)e;
                uloc[q] = (uint8_t)i;
            }
    }

    void curve_order()
    {
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        // bounding box from the vertices actually used
        for (int64_t e = 0; e < nElt; ++e)
            for (int a = 0; a < 4; ++a)
                for (int c = 0; c < 3; ++c) {
                    double x = xyz[3 * (int64_t)conn[4 * e + a] + c];
                    lo[c] = std::min(lo[c], x);
                    hi[c] = std::max(hi[c], x);
                }
        double sc[3];
        for (int c = 0; c < 3; ++c) sc[c] = hi[c] > lo[c] ? 1023.999 / (hi[c] - lo[c]) : 0.0;
        std::vector<uint64_t> key(nNode);
        kind.assign(nNode, 0);
        for (int64_t k = 0; k < nNode; ++k) {
            if (nptr[k] == nptr[k + 1]) { key[k] = ((uint64_t)0x3FFFFFFF << 32) | (uint32_t)k; continue; }
            const int64_t e = uelt[nptr[k]];
            const int i = uloc[nptr[k]];
            kind[k] = (uint8_t)rot[i].kind;
            int v[3];
            const int nv = dof_entity_vertices(order, i, v);
            double p[3] = {0, 0, 0};
            for (int a = 0; a < nv; ++a)
                for (int c = 0; c < 3; ++c) p[c] += xyz[3 * (int64_t)conn[4 * e + v[a]] + c];
            uint32_t m = 0;
            for (int c = 0; c < 3; ++c) m |= spread3((uint32_t)((p[c] / nv - lo[c]) * sc[c])) << c;
            key[k] = ((uint64_t)m << 32) | (uint32_t)k;
        }
        std::sort(key.begin(), key.end());
        order_.resize(nNode);
        for (int64_t k = 0; k < nNode; ++k) order_[k] = (int32_t)(key[k] & 0xFFFFFFFFu);
    }
};

} // namespace

namespace {

struct CtaSeg { int64_t nb, ne; uint32_t visit_off, nvisit; };

struct Local {                      // per-thread output, concatenated at the end
    std::vector<RowPlanHeader> hdrs;
    std::vector<uint32_t> blob;     // 4 words per quad
    std::vector<std::pair<uint32_t, uint32_t>> cta_bundles;   // (first bundle, count) per CTA handled
    int maxchunks = 1, maxrounds = 1;
    int64_t nunits = 0, ncontrib = 0, nsrc = 0;
    std::string error;
};

inline int pair_idx(int x, int y) { return x + y - 1 + (std::min(x, y) >= 1 ? 1 : 0); }

// builds one bundle from nodes[0..cnt) (same kind); returns false if a capacity limit is hit
bool make_bundle(const Builder& B, const std::vector<uint16_t>& uvis, const int32_t* nodes, int cnt, Local& L,
                 std::vector<std::vector<uint16_t>>& srcOfOut, std::vector<int>& taskOrder, bool& hardFail)
{
    const int n = B.n;
    int64_t totalUnits = 0;
    int nout = 0;
    for (int k = 0; k < cnt; ++k) {
        const int64_t nd = nodes[k];
        totalUnits += B.nptr[nd + 1] - B.nptr[nd];
        nout += (int)((B.ia[nd * B.ncomp + 1] - B.ia[nd * B.ncomp]) / B.ncomp);
    }
    const int nchunks = (int)std::max<int64_t>(1, (totalUnits + 31) / 32);
    if (cnt > 1 && totalUnits > 32) return false;
    if (cnt > 31) return false;
    if (nchunks > ROWPLAN_MAXCHUNKS || nout > 32 * ROWPLAN_MAXROUNDS) {
        if (cnt == 1) hardFail = true;
        return false;
    }
    const int kind = B.kind[nodes[0]];
    const bool warpDiag = (cnt == 1 && totalUnits > ROWPLAN_BIGROW);
    std::vector<uint32_t> unitw((size_t)nchunks * 32, 0u);
    if ((int)srcOfOut.size() < nout) srcOfOut.resize(nout);
    for (int o = 0; o < nout; ++o) srcOfOut[o].clear();
    int diagOut = -1;
    int lane = 0, outoff = 0;
    std::vector<int> rowOff(cnt + 1, 0);
    for (int k = 0; k < cnt; ++k) {
        const int64_t nd = nodes[k];
        const int len = (int)((B.ia[nd * B.ncomp + 1] - B.ia[nd * B.ncomp]) / B.ncomp);
        for (int64_t q = B.nptr[nd]; q < B.nptr[nd + 1]; ++q, ++lane) {
            const int64_t e = B.uelt[q];
            const int i = B.uloc[q];
            const RowPlanRot& R = B.rot[i];
            const int p0 = R.perm[0], p1 = R.perm[1], p2 = R.perm[2], p3 = R.perm[3];
            const uint32_t pairs = pair_idx(p0, p1) | (pair_idx(p0, p2) << 3) | (pair_idx(p0, p3) << 6) |
                                   (pair_idx(p1, p2) << 9) | (pair_idx(p1, p3) << 12) | (pair_idx(p2, p3) << 15);
            unitw[lane] = (uint32_t)uvis[q] | (pairs << 10) | ROWPLAN_UNIT_ACTIVE;
            const uint16_t* pe = B.pos + (e * n + i) * n;     // positions of the element's columns in this row
            const int chunk = lane >> 5, ln = lane & 31;
            for (int jr = 0; jr < n; ++jr) {
                const int jo = R.rmap[jr];
                const int p = pe[jo];
                if (p >= len) { L.error = "row plan: position outside the row"; hardFail = true; return false; }
                if (jo == i && warpDiag) { diagOut = outoff + p; continue; }   // diagonal: warp-reduced in phase 1
                srcOfOut[outoff + p].push_back((uint16_t)(ROWPLAN_ZOFF + (chunk * n + jr) * 32 + ln));
            }
        }
        outoff += len;
        rowOff[k + 1] = outoff;
    }
    // LPT dealing of the outputs over the 32 lanes
    taskOrder.clear();
    for (int o = 0; o < nout; ++o)
        if (!srcOfOut[o].empty()) taskOrder.push_back(o);
    std::stable_sort(taskOrder.begin(), taskOrder.end(),
                     [&](int a, int b) { return srcOfOut[a].size() > srcOfOut[b].size(); });
    int load[32], ntask[32];
    std::vector<int> laneTasks[32];
    for (int l = 0; l < 32; ++l) { load[l] = 0; ntask[l] = 0; }
    // phase 3 reads result slot (task*32 + owner lane) of 16 consecutive outputs per half-warp: owners of an aligned
    // group of 16 outputs should sit in distinct shared-memory banks (owner lane mod 16)
    std::vector<uint16_t> groupMask((nout + 15) / 16, 0);
    for (int o : taskOrder) {
        int minload = load[0];
        for (int l = 1; l < 32; ++l) minload = std::min(minload, load[l]);
        const uint16_t gm = groupMask[o >> 4];
        int best = -1, bestFree = -1;
        for (int l = 0; l < 32; ++l) {
            if (load[l] > minload + 1) continue;
            const bool freeBank = !((gm >> (l & 15)) & 1);
            if (freeBank && (bestFree < 0 || load[l] < load[bestFree])) bestFree = l;
            if (best < 0 || load[l] < load[best]) best = l;
        }
        if (bestFree >= 0) best = bestFree;
        groupMask[o >> 4] |= (uint16_t)(1u << (best & 15));
        load[best] += (int)srcOfOut[o].size();
        laneTasks[best].push_back(o);
        ntask[best]++;
    }
    int niter = 0, nrounds = 0;
    for (int l = 0; l < 32; ++l) { niter = std::max(niter, load[l]); nrounds = std::max(nrounds, ntask[l]); }
    if (nrounds > ROWPLAN_MAXROUNDS) {
        if (cnt == 1) hardFail = true;
        return false;
    }
    const int niter4 = (niter + 3) / 4;
    // ---- emit the blob ----
    RowPlanHeader h;
    std::memset(&h, 0, sizeof(h));
    h.q0 = (uint32_t)(L.blob.size() / 4);
    h.niter4 = (uint16_t)niter4;
    h.nout = (uint16_t)nout;
    h.nrows = (uint8_t)cnt;
    h.nchunks = (uint8_t)nchunks;
    h.kind = (uint8_t)kind;
    h.flags = warpDiag ? ROWPLAN_FLAG_WARPDIAG : 0;
    const size_t qUnits = (size_t)nchunks * 8, qRows = (size_t)cnt, qSrc = (size_t)niter4 * 16, qOmap = (size_t)(nout + 7) / 8;
    const size_t base = L.blob.size();
    L.blob.resize(base + 4 * (qUnits + qRows + qSrc + qOmap), 0u);
    uint32_t* w = L.blob.data() + base;
    std::memcpy(w, unitw.data(), sizeof(uint32_t) * unitw.size());
    w += 4 * qUnits;
    for (int k = 0; k < cnt; ++k) {
        const int64_t nd = nodes[k];
        const int64_t b0 = B.ia[nd * B.ncomp];
        const int32_t len = (int32_t)((B.ia[nd * B.ncomp + 1] - b0) / B.ncomp);
        std::memcpy(w + 4 * k, &b0, 8);
        std::memcpy(w + 4 * k + 2, &len, 4);
        w[4 * k + 3] = (uint32_t)rowOff[k];
    }
    w += 4 * qRows;
    uint16_t* sp = reinterpret_cast<uint16_t*>(w);            // lane-major, 4*niter4 entries per lane, null = 0
    std::vector<uint16_t> om(qOmap * 8, (uint16_t)ROWPLAN_OMAP_ZERO);
    {
        // row slot of every output
        int k = 0;
        for (int o = 0; o < nout; ++o) {
            while (o >= rowOff[k + 1]) ++k;
            om[o] = (uint16_t)(ROWPLAN_OMAP_ZERO | (k << 11));
        }
    }
    // Phase 2 reads one source per lane and iteration; the bank of a source is its unit lane mod 16 and a half-warp is
    // served in one pass only if its 16 lanes hit distinct banks.  The order of the sources inside an output is free
    // (it only fixes the summation order), so pick, iteration by iteration, a source whose bank is still unused.
    {
        int curTask[32], remaining[32];
        std::vector<uint16_t> pool[32];
        for (int l = 0; l < 32; ++l) {
            curTask[l] = 0;
            remaining[l] = 0;
            for (size_t t = 0; t < laneTasks[l].size(); ++t) om[laneTasks[l][t]] = (uint16_t)((om[laneTasks[l][t]] & 0xF800u) | (uint16_t)(32 + t * 32 + l));
            if (!laneTasks[l].empty()) { pool[l] = srcOfOut[laneTasks[l][0]]; remaining[l] = (int)pool[l].size(); }
        }
        for (int it = 0; it < niter; ++it)
            for (int half = 0; half < 2; ++half) {
                uint32_t used = 0;
                for (int l = half * 16; l < half * 16 + 16; ++l) {
                    if (curTask[l] >= (int)laneTasks[l].size()) continue;   // stream exhausted: stays null (zero cell)
                    int pick = 0;
                    for (int k = 0; k < remaining[l]; ++k)
                        if (!((used >> (pool[l][k] & 15)) & 1)) { pick = k; break; }
                    const uint16_t sv = pool[l][pick];
                    used |= 1u << (sv & 15);
                    pool[l][pick] = pool[l][remaining[l] - 1];
                    --remaining[l];
                    uint16_t ent = sv;
                    if (remaining[l] == 0) {
                        ent |= ROWPLAN_SRC_END;
                        ++curTask[l];
                        if (curTask[l] < (int)laneTasks[l].size()) {
                            pool[l] = srcOfOut[laneTasks[l][curTask[l]]];
                            remaining[l] = (int)pool[l].size();
                        }
                    }
                    sp[(size_t)l * 4 * niter4 + it] = ent;
                }
            }
    }
    if (diagOut >= 0) om[diagOut] = (uint16_t)((om[diagOut] & 0xF800u) | ROWPLAN_OMAP_DIAG);   // result slot 1
    w += 4 * qSrc;
    std::memcpy(w, om.data(), sizeof(uint16_t) * om.size());
    L.hdrs.push_back(h);
    L.maxchunks = std::max(L.maxchunks, nchunks);
    L.maxrounds = std::max(L.maxrounds, nrounds);
    L.nunits += totalUnits;
    L.nsrc += (int64_t)niter4 * 4 * 32;
    for (int o = 0; o < nout; ++o) L.ncontrib += (int64_t)srcOfOut[o].size();
    return true;
}

} // namespace

bool rowplan_build_host(int order, int ncomp, int64_t nElt, const int32_t* l2g, int nd, int64_t nPts,
                        const double* xyz, const int32_t* conn, int64_t nRows, const int64_t* ia,
                        const uint16_t* pos, int nthreads, RowPlanHost& plan)
{
    (void)nPts;
    Builder B;
    B.order = order;
    B.ncomp = ncomp;
    B.n = nloc(order);
    B.nd = nd;
    B.nElt = nElt;
    B.nNode = nRows / ncomp;
    B.l2g = l2g;
    B.xyz = xyz;
    B.conn = conn;
    B.ia = ia;
    B.pos = pos;
    rowplan_rotation_tables(order, B.rot, B.canon);
    for (int i = 0; i < B.n; ++i)
        if (B.rot[i].rmap[B.canon[B.rot[i].kind]] != i) { plan.error = "row plan: rotation table inconsistent"; return false; }
    if (nRows % ncomp) { plan.error = "row plan: rows not a multiple of ncomp"; return false; }
    for (int64_t e = 0; e < nElt; ++e)
        for (int i = 0; i < B.n; ++i)
            if (l2g[e * nd + i] >= nRows) { plan.error = "row plan: a test dof has no CSR row"; return false; }
    B.node_units();
    B.curve_order();
    if (nthreads <= 0) nthreads = (int)std::max(1u, std::thread::hardware_concurrency());
    nthreads = std::min(nthreads, 64);

    // ---- CTA packing along the curve: as many consecutive rows as keep the visit set within ROWPLAN_VMAX ----
    std::vector<CtaSeg> segs;
    std::vector<uint16_t> uvis(B.uelt.size());
    std::vector<int32_t> stamp(nElt, -1), vidx(nElt, 0);
    plan.visits.clear();
    {
        int64_t k = 0;
        while (k < B.nNode) {
            CtaSeg sg;
            sg.nb = k;
            sg.visit_off = (uint32_t)plan.visits.size();
            const int32_t id = (int32_t)segs.size();
            uint32_t nv = 0;
            int64_t rows = 0;
            for (; k < B.nNode; ++k) {
                const int64_t node = B.order_[k];
                uint32_t add = 0;
                for (int64_t q = B.nptr[node]; q < B.nptr[node + 1]; ++q)
                    if (stamp[B.uelt[q]] != id) ++add;
                if (rows > 0 && (nv + add > ROWPLAN_VMAX || rows >= 4096)) break;
                if (add > ROWPLAN_VMAX) { plan.error = "row plan: a node has more elements than ROWPLAN_VMAX"; return false; }
                for (int64_t q = B.nptr[node]; q < B.nptr[node + 1]; ++q) {
                    const int32_t e = B.uelt[q];
                    if (stamp[e] != id) {
                        stamp[e] = id;
                        vidx[e] = (int32_t)nv++;
                        plan.visits.push_back(e);
                    }
                    uvis[q] = (uint16_t)vidx[e];
                }
                ++rows;
            }
            sg.ne = k;
            sg.nvisit = nv;
            segs.push_back(sg);
        }
    }
    stamp = std::vector<int32_t>();
    vidx = std::vector<int32_t>();

    // ---- bundles, in parallel over CTAs ----
    const int64_t nCta = (int64_t)segs.size();
    std::vector<Local> locals(nthreads);
    auto work = [&](int t) {
        Local& L = locals[t];
        std::vector<std::vector<uint16_t>> srcOfOut;
        std::vector<int> taskOrder;
        std::vector<int32_t> byKind[3];
        for (int64_t c = nCta * t / nthreads; c < nCta * (t + 1) / nthreads; ++c) {
            const CtaSeg& sg = segs[c];
            for (int kd = 0; kd < 3; ++kd) byKind[kd].clear();
            for (int64_t k = sg.nb; k < sg.ne; ++k) byKind[B.kind[B.order_[k]]].push_back(B.order_[k]);
            const uint32_t first = (uint32_t)L.hdrs.size();
            for (int kd = 0; kd < 3; ++kd) {
                const std::vector<int32_t>& nodes = byKind[kd];
                size_t s = 0;
                while (s < nodes.size()) {
                    // greedy: the longest run of small rows whose units fit one warp; a big row goes alone
                    size_t t2 = s;
                    int64_t units = 0;
                    while (t2 < nodes.size() && t2 - s < 31) {
                        const int64_t u = B.nptr[nodes[t2] + 1] - B.nptr[nodes[t2]];
                        if (u > ROWPLAN_BIGROW) {
                            if (t2 == s) ++t2;
                            break;
                        }
                        if (units + u > 32) break;
                        units += u;
                        ++t2;
                    }
                    bool hard = false;
                    while (!make_bundle(B, uvis, &nodes[s], (int)(t2 - s), L, srcOfOut, taskOrder, hard)) {
                        if (hard || t2 - s <= 1) {
                            if (L.error.empty()) L.error = "row plan: a row exceeds the plan limits (elements around a node / row length)";
                            return;
                        }
                        t2 = s + (t2 - s + 1) / 2;
                    }
                    s = t2;
                }
            }
            L.cta_bundles.emplace_back(first, (uint32_t)L.hdrs.size() - first);
        }
    };
    {
        std::vector<std::thread> th;
        for (int t = 1; t < nthreads; ++t) th.emplace_back(work, t);
        work(0);
        for (auto& x : th) x.join();
    }
    for (auto& L : locals)
        if (!L.error.empty()) { plan.error = L.error; return false; }

    // ---- concatenate ----
    plan.order = order;
    plan.n = B.n;
    plan.ncomp = ncomp;
    size_t nb = 0, nq = 0;
    for (auto& L : locals) { nb += L.hdrs.size(); nq += L.blob.size() / 4; }
    if (nq + 64 > 0xFFFFFFFFull) { plan.error = "row plan: blob offsets exceed 32 bits"; return false; }
    plan.hdrs.reserve(nb);
    plan.blob.reserve(nq * 4 + 256);
    plan.ctas.resize(nCta);
    int64_t c = 0;
    for (auto& L : locals) {
        const uint32_t hb = (uint32_t)plan.hdrs.size(), qb = (uint32_t)(plan.blob.size() / 4);
        for (auto h : L.hdrs) {
            h.q0 += qb;
            plan.hdrs.push_back(h);
        }
        plan.blob.insert(plan.blob.end(), L.blob.begin(), L.blob.end());
        for (auto& cb : L.cta_bundles) {
            plan.ctas[c].visit_off = segs[c].visit_off;
            plan.ctas[c].nvisit = segs[c].nvisit;
            plan.ctas[c].bundle_off = hb + cb.first;
            plan.ctas[c].nbundle = cb.second;
            ++c;
        }
        plan.maxchunks = std::max(plan.maxchunks, L.maxchunks);
        plan.maxrounds = std::max(plan.maxrounds, L.maxrounds);
        plan.nunits += L.nunits;
        plan.ncontrib += L.ncontrib;
        plan.nsrc_padded += L.nsrc;
        L = Local();
    }
    plan.blob.resize(plan.blob.size() + 256, 0u);   // the kernel may read a few quads past the last blob
    RowPlanHeader tail;
    std::memset(&tail, 0, sizeof(tail));
    for (int k = 0; k < ROWPLAN_WARPS; ++k) plan.hdrs.push_back(tail);   // header prefetch of "the next bundle"
    plan.nvisits = (int64_t)plan.visits.size();
    return true;
}
