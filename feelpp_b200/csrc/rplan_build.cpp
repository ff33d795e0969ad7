// rplan_build.cpp -- host builder of the row-thread plan (see rplan.h).  One pass over the node -> element adjacency:
// node rows ordered along a 3-D Hilbert curve, cut greedily into patches that respect the row-block, visit and row limits.
#include "rplan.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace {

// position of the grid point (x,y,z), 10 bits each, along the 3-D Hilbert curve (Skilling's transpose algorithm)
inline uint32_t hilbert3(uint32_t x, uint32_t y, uint32_t z)
{
    uint32_t X[3] = {x & 0x3FF, y & 0x3FF, z & 0x3FF};
    const uint32_t M = 1u << 9;
    for (uint32_t Q = M; Q > 1; Q >>= 1) {
        const uint32_t P = Q - 1;
        for (int i = 0; i < 3; ++i) {
            if (X[i] & Q) X[0] ^= P;
            else { const uint32_t t = (X[0] ^ X[i]) & P; X[0] ^= t; X[i] ^= t; }
        }
    }
    for (int i = 1; i < 3; ++i) X[i] ^= X[i - 1];
    uint32_t t = 0;
    for (uint32_t Q = M; Q > 1; Q >>= 1)
        if (X[2] & Q) t ^= Q - 1;
    for (int i = 0; i < 3; ++i) X[i] ^= t;
    uint32_t h = 0;
    for (int b = 9; b >= 0; --b)
        for (int i = 0; i < 3; ++i) h = (h << 1) | ((X[i] >> b) & 1);
    return h;
}

template <class F>
void parallel_for(int64_t n, int nthreads, F f)
{
    if (nthreads <= 1 || n < 4096) { f(0, n, 0); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back([=] { f(n * t / nthreads, n * (t + 1) / nthreads, t); });
    for (auto& x : th) x.join();
}

} // namespace

bool rplan_build_host(int order, int ncomp, int64_t nElt, const int32_t* l2g, int nd, int64_t nPts, const double* xyz,
                      const int32_t* conn, int64_t nRows, const int64_t* ia, int maxslots, int maxvisits, int maxrows,
                      int nthreads, RPlanHost& plan)
{
    (void)nPts;
    const bool timing = std::getenv("FEELGPU_RPLAN_TIMING") != nullptr;
    auto tprev = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "rplan build: %-12s %.3f s\n", what, std::chrono::duration<double>(now - tprev).count());
        tprev = now;
    };
    if (nthreads <= 0) {
        const char* bt = std::getenv("FEELGPU_BUILD_THREADS");
        nthreads = bt && std::atoi(bt) > 0 ? std::atoi(bt) : (int)std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
    }
    nthreads = std::min(nthreads, 64);
    const int n = (order + 1) * (order + 2) * (order + 3) / 6;
    plan.order = order;
    plan.n = n;
    plan.ncomp = ncomp;
    plan.maxslots = maxslots;
    plan.maxvisits = maxvisits;
    plan.maxrows = maxrows;
    if (nd != n * ncomp) { plan.error = "row-thread plan: nd != n * ncomp"; return false; }
    if (maxvisits > 4096 || n > 32) { plan.error = "row-thread plan: unit word overflow"; return false; }
    maxrows = std::min(maxrows, 32767);
    const int64_t nNode = nRows / ncomp;
    // ---- node -> (element, local dof) adjacency ----
    // (every thread owns a range of nodes and streams over the whole element -> dof table for it: no atomics, the units of a
    // node stay in element order, and the sequential re-reads cost less than the scattered increments they spread out)
    std::vector<int64_t> nptr((size_t)nNode + 1, 0);
    const int nadj = (nthreads > 1 && nElt >= 65536) ? nthreads : 1;
    auto node_range = [&](int t, int64_t& lo, int64_t& hi) { lo = nNode * t / nadj; hi = nNode * (t + 1) / nadj; };
    auto for_ranges = [&](auto body) {
        if (nadj == 1) { body(0); return; }
        std::vector<std::thread> th;
        for (int t = 1; t < nadj; ++t) th.emplace_back(body, t);
        body(0);
        for (auto& x : th) x.join();
    };
    for_ranges([&](int t) {
        int64_t lo, hi;
        node_range(t, lo, hi);
        int64_t* cnt = nptr.data() + 1;
        // (the division by the component count is kept out of the loop: 1 and 3 are the cases that occur)
        auto count = [&](auto nodeof) {
            for (int64_t e = 0; e < nElt; ++e)
                for (int i = 0; i < n; ++i) { const int64_t node = nodeof(l2g[e * nd + i]); if (node >= lo && node < hi) ++cnt[node]; }
        };
        if (ncomp == 1) count([](int32_t d) { return (int64_t)d; });
        else if (ncomp == 3) count([](int32_t d) { return (int64_t)(d / 3); });
        else count([ncomp](int32_t d) { return (int64_t)(d / ncomp); });
    });
    for (int64_t k = 0; k < nNode; ++k) nptr[(size_t)k + 1] += nptr[(size_t)k];
    std::vector<int32_t> uelt((size_t)nptr[(size_t)nNode]);
    std::vector<uint8_t> uloc((size_t)nptr[(size_t)nNode]);
    {
        std::vector<int64_t> cur(nptr.begin(), nptr.end() - 1);
        for_ranges([&](int t) {
            int64_t lo, hi;
            node_range(t, lo, hi);
            auto fill = [&](auto nodeof) {
                for (int64_t e = 0; e < nElt; ++e)
                    for (int i = 0; i < n; ++i) {
                        const int64_t node = nodeof(l2g[e * nd + i]);
                        if (node < lo || node >= hi) continue;
                        const int64_t q = cur[(size_t)node]++;
                        uelt[(size_t)q] = (int32_t)e;
                        uloc[(size_t)q] = (uint8_t)i;
                    }
            };
            if (ncomp == 1) fill([](int32_t d) { return (int64_t)d; });
            else if (ncomp == 3) fill([](int32_t d) { return (int64_t)(d / 3); });
            else fill([ncomp](int32_t d) { return (int64_t)(d / ncomp); });
        });
    }
    lap("adjacency");
    // ---- curve order of the nodes (position = centroid of the node's entity in its first element) ----
    static const int E2P[6][2] = {{1, 2}, {2, 0}, {0, 1}, {0, 3}, {1, 3}, {2, 3}};
    static const int F2P[4][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}};
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int64_t e = 0; e < nElt; ++e)
        for (int a = 0; a < 4; ++a)
            for (int c = 0; c < 3; ++c) {
                const double x = xyz[3 * (int64_t)conn[4 * e + a] + c];
                lo[c] = std::min(lo[c], x);
                hi[c] = std::max(hi[c], x);
            }
    double ext = 0.0;
    for (int c = 0; c < 3; ++c) ext = std::max(ext, hi[c] - lo[c]);
    const double sc = ext > 0.0 ? 1023.999 / ext : 0.0;
    const int npe = order - 1;
    std::vector<uint64_t> key((size_t)nNode);
    parallel_for(nNode, nthreads, [&](int64_t k0, int64_t k1, int) {
        for (int64_t k = k0; k < k1; ++k) {
            if (nptr[(size_t)k] == nptr[(size_t)k + 1]) { key[(size_t)k] = ((uint64_t)0x3FFFFFFF << 32) | (uint32_t)k; continue; }
            const int64_t e = uelt[(size_t)nptr[(size_t)k]];
            const int i = uloc[(size_t)nptr[(size_t)k]];
            int v[3], nv;
            if (i < 4) { v[0] = i; nv = 1; }
            else if (i < 4 + 6 * npe) { const int ed = (i - 4) / npe; v[0] = E2P[ed][0]; v[1] = E2P[ed][1]; nv = 2; }
            else { const int f = std::min(3, i - 4 - 6 * npe); v[0] = F2P[f][0]; v[1] = F2P[f][1]; v[2] = F2P[f][2]; nv = 3; }
            double p[3] = {0, 0, 0};
            for (int a = 0; a < nv; ++a)
                for (int c = 0; c < 3; ++c) p[c] += xyz[3 * (int64_t)conn[4 * e + v[a]] + c];
            const uint32_t m = hilbert3((uint32_t)((p[0] / nv - lo[0]) * sc), (uint32_t)((p[1] / nv - lo[1]) * sc),
                                        (uint32_t)((p[2] / nv - lo[2]) * sc));
            key[(size_t)k] = ((uint64_t)m << 32) | (uint32_t)k;
        }
    });
    if (nNode >= ((int64_t)1 << 32)) { plan.error = "row-thread plan: more than 2^32 nodes"; return false; }
    lap("keys");
    {   // parallel sort: chunks sorted by the threads, then merged pairwise
        int nch = 1;
        while (nch * 2 <= nthreads && (int64_t)key.size() / (nch * 2) > 65536) nch *= 2;
        std::vector<size_t> cut(nch + 1);
        for (int c = 0; c <= nch; ++c) cut[c] = key.size() * (size_t)c / nch;
        {
            std::vector<std::thread> th;
            for (int c = 0; c < nch; ++c) th.emplace_back([&, c] { std::sort(key.begin() + cut[c], key.begin() + cut[c + 1]); });
            for (auto& x : th) x.join();
        }
        for (int w = 1; w < nch; w *= 2) {
            std::vector<std::thread> th;
            for (int c = 0; c + w < nch; c += 2 * w)
                th.emplace_back([&, c, w] { std::inplace_merge(key.begin() + cut[c], key.begin() + cut[c + w], key.begin() + cut[std::min(c + 2 * w, nch)]); });
            for (auto& x : th) x.join();
        }
    }
    lap("sort");
    // ---- greedy cut into patches ----
    // The curve is cut into pieces of CUT_PIECE rows that are handled independently (a piece ends with a patch that may be
    // short, 1 in ~300); the pieces depend on the mesh only, so the plan does not depend on the thread count.
    struct Cut { int64_t k0, k1; uint32_t nvisit, nslots, nunits; };
    std::vector<Cut> cuts;
    {
        constexpr int64_t CUT_PIECE = 1 << 16;
        const int64_t npiece = std::max<int64_t>(1, (nNode + CUT_PIECE - 1) / CUT_PIECE);
        std::vector<std::vector<Cut>> pcuts((size_t)npiece);
        std::vector<int> perr((size_t)npiece, 0);
        int64_t next_piece = 0;
        auto cutter = [&]() {
            std::vector<int32_t> stamp;
            int32_t idnext = 0;                    // stamps are unique inside this thread's array
            for (;;) {
                const int64_t pi = __atomic_fetch_add(&next_piece, (int64_t)1, __ATOMIC_RELAXED);
                if (pi >= npiece) break;
                if (stamp.empty()) stamp.assign((size_t)nElt, -1);
                std::vector<Cut>& lc = pcuts[(size_t)pi];
                const int32_t idbase = idnext;
                const int64_t kbeg = pi * CUT_PIECE, kend = std::min(nNode, (pi + 1) * CUT_PIECE);
                int64_t k0 = kbeg;
                uint32_t nvis = 0, nslots = 0, nun = 0;
                int nrw = 0;
                for (int64_t k = kbeg; k < kend; ++k) {
                    const int64_t node = (int64_t)(key[(size_t)k] & 0xFFFFFFFFu);
                    const int64_t len = ia[node * ncomp + 1] - ia[node * ncomp];
                    const int64_t need = len * ncomp;
                    const int64_t u0 = nptr[(size_t)node], u1 = nptr[(size_t)node + 1];
                    if (need > maxslots || u1 - u0 > maxvisits) { perr[(size_t)pi] = 1; return; }
                    for (int pass = 0; pass < 2; ++pass) {
                        const int32_t id = idbase + (int32_t)lc.size();
                        uint32_t fresh = 0;
                        for (int64_t q = u0; q < u1; ++q)
                            if (stamp[(size_t)uelt[(size_t)q]] != id) ++fresh;
                        if (pass == 0 && k > k0 && (nslots + need > (uint32_t)maxslots || nvis + fresh > (uint32_t)maxvisits || nrw + 1 > maxrows)) {
                            lc.push_back(Cut{k0, k, nvis, nslots, nun});
                            k0 = k;
                            nvis = nslots = nun = 0;
                            nrw = 0;
                            continue;        // recount against the new patch
                        }
                        for (int64_t q = u0; q < u1; ++q) stamp[(size_t)uelt[(size_t)q]] = id;
                        nvis += fresh;
                        nslots += (uint32_t)need;
                        nun += (uint32_t)(u1 - u0);
                        ++nrw;
                        break;
                    }
                }
                if (kend > k0) lc.push_back(Cut{k0, kend, nvis, nslots, nun});
                idnext = idbase + (int32_t)lc.size() + 1;
            }
        };
        {
            const int ncut = (int)std::max<int64_t>(1, std::min<int64_t>(std::min(nthreads, 16), npiece));
            std::vector<std::thread> th;
            for (int t = 1; t < ncut; ++t) th.emplace_back(cutter);
            cutter();
            for (auto& x : th) x.join();
        }
        for (int64_t pi = 0; pi < npiece; ++pi) {
            if (perr[(size_t)pi]) { plan.error = "row-thread plan: one node row does not fit a patch"; return false; }
            cuts.insert(cuts.end(), pcuts[(size_t)pi].begin(), pcuts[(size_t)pi].end());
        }
    }
    lap("cut");
    // ---- emit ----
    const int64_t nP = (int64_t)cuts.size();
    plan.patches.resize((size_t)nP);
    std::vector<int64_t> voff((size_t)nP + 1, 0), roff((size_t)nP + 1, 0), uoff((size_t)nP + 1, 0);
    for (int64_t p = 0; p < nP; ++p) {
        voff[(size_t)p + 1] = voff[(size_t)p] + cuts[(size_t)p].nvisit;
        roff[(size_t)p + 1] = roff[(size_t)p] + (cuts[(size_t)p].k1 - cuts[(size_t)p].k0);
        uoff[(size_t)p + 1] = uoff[(size_t)p] + cuts[(size_t)p].nunits;
        plan.maxpatch_visits = std::max(plan.maxpatch_visits, (int)cuts[(size_t)p].nvisit);
        plan.maxpatch_rows = std::max(plan.maxpatch_rows, (int)(cuts[(size_t)p].k1 - cuts[(size_t)p].k0));
        plan.maxpatch_slots = std::max(plan.maxpatch_slots, (int)cuts[(size_t)p].nslots);
    }
    if (voff[(size_t)nP] >= ((int64_t)1 << 32) || uoff[(size_t)nP] >= ((int64_t)1 << 32) || roff[(size_t)nP] >= ((int64_t)1 << 32)) {
        plan.error = "row-thread plan: 32-bit offsets overflow";
        return false;
    }
    plan.visits.resize((size_t)voff[(size_t)nP]);
    plan.rows.resize((size_t)roff[(size_t)nP]);
    plan.units.resize((size_t)uoff[(size_t)nP]);
    plan.nunits = uoff[(size_t)nP];
    plan.nvisits = voff[(size_t)nP];

    std::vector<std::vector<int32_t>> slot_of(nthreads);       // per thread: element -> visit slot of the current patch
    std::vector<std::vector<int32_t>> slot_stamp(nthreads);
    parallel_for(nP, nthreads, [&](int64_t p0, int64_t p1, int t) {
        auto& slot = slot_of[t];
        auto& stp = slot_stamp[t];
        slot.assign((size_t)nElt, 0);
        stp.assign((size_t)nElt, -1);
        std::vector<std::pair<uint32_t, int64_t>> byn;
        for (int64_t p = p0; p < p1; ++p) {
            const Cut& c = cuts[(size_t)p];
            RPlanPatch& P = plan.patches[(size_t)p];
            P.visit_off = (uint32_t)voff[(size_t)p];
            P.nvisit = c.nvisit;
            P.row_off = (uint32_t)roff[(size_t)p];
            P.nrows = (uint32_t)(c.k1 - c.k0);
            P.unit_off = (uint32_t)uoff[(size_t)p];
            P.nunits = c.nunits;
            P.nslots = c.nslots;
            P.pad = 0;
            // rows sorted by descending unit count: the lanes of a warp walk lists of similar length
            byn.clear();
            for (int64_t k = c.k0; k < c.k1; ++k) {
                const int64_t node = (int64_t)(key[(size_t)k] & 0xFFFFFFFFu);
                byn.emplace_back((uint32_t)(nptr[(size_t)node + 1] - nptr[(size_t)node]), node);
            }
            std::stable_sort(byn.begin(), byn.end(), [](auto const& a, auto const& b) { return a.first > b.first; });
            uint32_t nv = 0, off = 0, un = 0;
            for (size_t r = 0; r < byn.size(); ++r) {
                const int64_t node = byn[r].second;
                RPlanRow& R = plan.rows[(size_t)P.row_off + r];
                R.base = ia[node * ncomp];
                R.off = off;
                R.len = (uint32_t)(ia[node * ncomp + 1] - ia[node * ncomp]);
                R.unit0 = un;
                R.nunits = byn[r].first;
                R.pad[0] = R.pad[1] = 0;
                off += R.len * (uint32_t)ncomp;
                un += R.nunits;
                for (int64_t q = nptr[(size_t)node]; q < nptr[(size_t)node + 1]; ++q) {
                    const int32_t e = uelt[(size_t)q];
                    if (stp[(size_t)e] != (int32_t)p) {
                        stp[(size_t)e] = (int32_t)p;
                        slot[(size_t)e] = (int32_t)nv;
                        plan.visits[(size_t)P.visit_off + nv] = e;
                        ++nv;
                    }
                }
            }
            // unit words in "round" order: the k-th unit of every row that has one, k = 0, 1, ...: the lanes of a warp work
            // on different rows (their adds into the row block do not meet) and every lane has work whatever the row lengths
            uint32_t w = 0;
            for (uint32_t k = 0; !byn.empty() && k < byn[0].first; ++k) {
                for (size_t r = 0; r < byn.size() && byn[r].first > k; ++r) {
                    const int64_t q = nptr[(size_t)byn[r].second] + k;
                    plan.units[(size_t)P.unit_off + w++] =
                        (uint32_t)slot[(size_t)uelt[(size_t)q]] | ((uint32_t)uloc[(size_t)q] << 12) | ((uint32_t)r << 17);
                }
            }
        }
    });
    lap("emit");
    return true;
}
