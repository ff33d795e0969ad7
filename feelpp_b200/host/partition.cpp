// partition.cpp -- element partition of the mesh for the multi-GPU path, the way the reference does it.
//
// Feel++ partitions with METIS on the dual graph of the mesh (feelpartition/partitionermetis_impl.hpp:118-253): one graph
// vertex per element, an edge between two elements that share a face (elt.neighbor(ms), :176-186,210-222), vertex weight
// = points per element (:161), one constraint; METIS_PartGraphRecursive for nparts <= 8 (:227-239), METIS_PartGraphKway
// above (:243-253), default options.  METIS 5.1.0 is vendored by the reference (feelpp/contrib/metis, idx_t = int64,
// real_t = double, entry points renamed Feel_METIS_*); the top-level Makefile compiles those C sources where they lie into
// feelpp_b200/lib/libfeelmetis.a (a built artefact, never copied into the repo) and links it here.  Without that archive
// (FEELHOST_NO_METIS) feelhost_partition_metis is not exported and the harness falls back to coordinate bisection.
#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "mesh.hpp"

namespace feelpp_b200 {

// dual graph by faces: xadj[nElt+1], adj (<= 4 per element), neighbours in local face order (the order of elt.neighbor(ms))
void dualGraph(int64_t nPts, int64_t nElt, const int32_t* conn, std::vector<int64_t>& xadj, std::vector<int64_t>& adj)
{
    static const int F2P[4][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}};
    struct Rest { int32_t b, c; int64_t ef; };
    std::vector<int64_t> first((size_t)nPts + 1, 0);
    auto sorted3 = [&](int64_t e, int f, int32_t* v) {
        for (int k = 0; k < 3; ++k) {
            v[k] = conn[4 * e + F2P[f][k]];
            if (v[k] < 0 || v[k] >= nPts) throw std::invalid_argument("dualGraph: point id out of range");
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
    std::vector<int64_t> nb((size_t)nElt * 4, -1);      // neighbour through local face f, -1 on the boundary
    for (int64_t p = 0; p < nPts; ++p) {
        Rest* b = rest.data() + first[(size_t)p];
        Rest* e = rest.data() + first[(size_t)p + 1];
        std::sort(b, e, [](Rest const& x, Rest const& y) { return x.b != y.b ? x.b < y.b : (x.c != y.c ? x.c < y.c : x.ef < y.ef); });
        for (Rest* i = b; i < e;) {
            Rest* k = i + 1;
            while (k < e && k->b == i->b && k->c == i->c) ++k;
            if (k - i == 2) {
                nb[(size_t)i->ef] = (i + 1)->ef / 4;
                nb[(size_t)(i + 1)->ef] = i->ef / 4;
            } else if (k - i > 2)
                throw std::invalid_argument("dualGraph: a face belongs to more than two elements");
            i = k;
        }
    }
    xadj.assign((size_t)nElt + 1, 0);
    adj.clear();
    adj.reserve((size_t)nElt * 4);
    for (int64_t e = 0; e < nElt; ++e) {
        for (int f = 0; f < 4; ++f)
            if (nb[(size_t)e * 4 + f] >= 0) adj.push_back(nb[(size_t)e * 4 + f]);
        xadj[(size_t)e + 1] = (int64_t)adj.size();
    }
}

} // namespace feelpp_b200

#ifndef FEELHOST_NO_METIS
extern "C" {
// prototypes of the vendored library (feelpp/contrib/metis/include/feelmetis.h:195-206; idx_t = int64_t, real_t = double)
int Feel_METIS_PartGraphRecursive(int64_t* nvtxs, int64_t* ncon, int64_t* xadj, int64_t* adjncy, int64_t* vwgt, int64_t* vsize,
                                  int64_t* adjwgt, int64_t* nparts, double* tpwgts, double* ubvec, int64_t* options,
                                  int64_t* edgecut, int64_t* part);
int Feel_METIS_PartGraphKway(int64_t* nvtxs, int64_t* ncon, int64_t* xadj, int64_t* adjncy, int64_t* vwgt, int64_t* vsize,
                             int64_t* adjwgt, int64_t* nparts, double* tpwgts, double* ubvec, int64_t* options,
                             int64_t* edgecut, int64_t* part);
int Feel_METIS_SetDefaultOptions(int64_t* options);
}
#endif

static thread_local std::string g_perr;

extern "C" {

const char* feelhost_partition_last_error() { return g_perr.c_str(); }

int feelhost_dual_graph(int64_t nPts, int64_t nElt, const int32_t* conn, int32_t* xadj, int32_t* adj)
{
    try {
        std::vector<int64_t> xa, ad;
        feelpp_b200::dualGraph(nPts, nElt, conn, xa, ad);
        if (ad.size() >= ((size_t)1 << 31)) throw std::invalid_argument("dual graph: more than 2^31 edges");
        for (size_t k = 0; k < xa.size(); ++k) xadj[k] = (int32_t)xa[k];
        for (size_t k = 0; k < ad.size(); ++k) adj[k] = (int32_t)ad[k];
        return 0;
    } catch (std::exception const& e) {
        g_perr = e.what();
        return 1;
    }
}

#ifndef FEELHOST_NO_METIS
// seed < 0: METIS' default options (NULL, as the reference passes them)
int feelhost_partition_metis(int64_t nPts, int64_t nElt, const int32_t* conn, int nparts, int seed, int32_t* part)
{
    try {
        if (nparts < 1) throw std::invalid_argument("partition: nparts < 1");
        if (nparts == 1 || nElt == 0) {
            std::fill(part, part + nElt, 0);
            return 0;
        }
        std::vector<int64_t> xadj, adj;
        feelpp_b200::dualGraph(nPts, nElt, conn, xadj, adj);
        if (adj.empty()) adj.push_back(0);               // (the reference keeps a non-empty array for a disconnected graph)
        std::vector<int64_t> vwgt((size_t)nElt, 4);      // elt.numPoints (partitionermetis_impl.hpp:161)
        std::vector<int64_t> p((size_t)nElt, 0);
        int64_t n = nElt, ncon = 1, np = nparts, edgecut = 0;
        int64_t options[40];                             // METIS_NOPTIONS
        int64_t* opt = nullptr;
        if (seed >= 0) {
            Feel_METIS_SetDefaultOptions(options);
            options[8] = seed;                           // METIS_OPTION_SEED
            opt = options;
        }
        const int rc = nparts <= 8
            ? Feel_METIS_PartGraphRecursive(&n, &ncon, xadj.data(), adj.data(), vwgt.data(), nullptr, nullptr, &np, nullptr,
                                            nullptr, opt, &edgecut, p.data())
            : Feel_METIS_PartGraphKway(&n, &ncon, xadj.data(), adj.data(), vwgt.data(), nullptr, nullptr, &np, nullptr,
                                       nullptr, opt, &edgecut, p.data());
        if (rc != 1) throw std::runtime_error("METIS returned error " + std::to_string(rc));   // METIS_OK = 1
        for (int64_t e = 0; e < nElt; ++e) part[e] = (int32_t)p[(size_t)e];
        return 0;
    } catch (std::exception const& e) {
        g_perr = e.what();
        return 1;
    }
}
#endif

} // extern "C"
