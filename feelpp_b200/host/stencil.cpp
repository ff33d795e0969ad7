#include "stencil.hpp"

#include <algorithm>
#include <thread>

namespace feelpp_b200 {

void computeGraph(DofTable const& dof, int64_t nElt, GraphCSR& graph, int nthreads)
{
    const int n = dof.n, nd = dof.nd, ncomp = dof.ncomp;
    const int64_t nNodes = dof.nDof / ncomp;
    if (nthreads <= 0) nthreads = (int)std::max(1u, std::thread::hardware_concurrency());
    // node -> elements (counting sort); node id = component-0 dof / ncomp
    std::vector<int64_t> ptr(nNodes + 1, 0);
    for (int64_t e = 0; e < nElt; ++e)
        for (int i = 0; i < n; ++i) ptr[dof.l2g[e * nd + i] / ncomp + 1]++;
    for (int64_t k = 0; k < nNodes; ++k) ptr[k + 1] += ptr[k];
    std::vector<int32_t> elts(ptr[nNodes]);
    {
        std::vector<int64_t> cur(ptr.begin(), ptr.end() - 1);
        for (int64_t e = 0; e < nElt; ++e)
            for (int i = 0; i < n; ++i) elts[cur[dof.l2g[e * nd + i] / ncomp]++] = (int32_t)e;
    }
    // pass 1: node-level row lengths; pass 2: fill
    std::vector<int64_t> nia(nNodes + 1, 0);
    std::vector<std::vector<int32_t>> rowsOfBlock(nthreads);
    auto gather = [&](int64_t node, std::vector<int32_t>& cols) {
        cols.clear();
        for (int64_t k = ptr[node]; k < ptr[node + 1]; ++k) {
            const int32_t* l = &dof.l2g[(int64_t)elts[k] * nd];
            for (int j = 0; j < n; ++j) cols.push_back(l[j] / ncomp);
        }
        cols.push_back((int32_t)node); // diagonal (graphcsr.cpp:633-659)
        std::sort(cols.begin(), cols.end());
        cols.erase(std::unique(cols.begin(), cols.end()), cols.end());
    };
    auto parallel = [&](auto&& fn) {
        std::vector<std::thread> th;
        for (int t = 1; t < nthreads; ++t) th.emplace_back(fn, t);
        fn(0);
        for (auto& x : th) x.join();
    };
    parallel([&](int t) {
        std::vector<int32_t> cols;
        for (int64_t node = nNodes * t / nthreads; node < nNodes * (t + 1) / nthreads; ++node) {
            gather(node, cols);
            nia[node + 1] = (int64_t)cols.size();
        }
    });
    for (int64_t k = 0; k < nNodes; ++k) nia[k + 1] += nia[k];
    graph.nRows = dof.nDof;
    graph.ia.assign(dof.nDof + 1, 0);
    for (int64_t node = 0; node < nNodes; ++node) {
        const int64_t len = (nia[node + 1] - nia[node]) * ncomp;
        for (int c = 0; c < ncomp; ++c) graph.ia[node * ncomp + c + 1] = len;
    }
    for (int64_t r = 0; r < dof.nDof; ++r) graph.ia[r + 1] += graph.ia[r];
    graph.ja.resize(graph.ia[dof.nDof]);
    parallel([&](int t) {
        std::vector<int32_t> cols;
        for (int64_t node = nNodes * t / nthreads; node < nNodes * (t + 1) / nthreads; ++node) {
            gather(node, cols);
            for (int c = 0; c < ncomp; ++c) {
                int64_t* out = &graph.ja[graph.ia[node * ncomp + c]];
                for (size_t k = 0; k < cols.size(); ++k)
                    for (int c2 = 0; c2 < ncomp; ++c2) *out++ = (int64_t)cols[k] * ncomp + c2;
            }
        }
    });
}

} // namespace feelpp_b200
