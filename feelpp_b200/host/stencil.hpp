// stencil.hpp -- sparsity graph of a (test, trial) pair in CSR form.
//
// Produces what Stencil::computeGraph (Pattern::COUPLED, feeldiscr/stencil.hpp:1560-1634) followed by
// addMissingZeroEntriesDiagonal and GraphCSR::close (feelalg/graphcsr.cpp:633-659, 994-1019) hand to
// MatSeqAIJSetPreallocationCSR (feelalg/matrixpetsc.cpp:392-413): ia[nRows+1], ja[nnz] with ascending columns.
// The reference inserts into one std::set per row; here the node-to-element adjacency is inverted once and every
// row is produced by a gather + sort + unique, in parallel over row blocks.
#pragma once
#include <cstdint>
#include <vector>
#include "doftable.hpp"

namespace feelpp_b200 {

struct GraphCSR {
    int64_t nRows = 0;
    std::vector<int64_t> ia, ja;
    int64_t nnz() const { return ia.empty() ? 0 : ia.back(); }
};

// test == trial space (the only case on this path); nthreads <= 0: hardware concurrency
void computeGraph(DofTable const& dof, int64_t nElt, GraphCSR& graph, int nthreads = 0);

} // namespace feelpp_b200
