// rowplan.h -- the owner-computes assembly plan ("row plan").
//
// Idea.  The reference adds each element's local matrix into the CSR with MatSetValues(ADD_VALUES)
// (feelalg/matrixpetsc.cpp:917-936): 100 read-modify-writes per P2 element, ~2.6 per nonzero.  On the GPU that is
// 800 M fp64 atomics per assembly of config 2 and a read-modify-write of the whole value array.  The row plan turns the
// scatter into a gather so that every CSR value is produced by exactly one thread and stored exactly once:
//
//   * unit   = (row node r, element e containing r): the row of e's local matrix that belongs to r.  The element's
//              vertices are re-labelled ("rotated") so that r's dof is always the canonical local dof of its kind
//              (vertex 0 / first dof of edge (0,1) / face (0,1,2)); the Lagrange nodal set is invariant under vertex
//              permutations, so the row is computed by fully static code from the standard reference tables.
//   * bundle = a few rows of one kind whose units fill the 32 lanes of a warp.  Phase 1: lane = unit, computes the n
//              entries of its row into a per-warp staging tile [entry][lane].  Phase 2: lane = a balanced list of output
//              nonzeros; for each it sums its sources from the tile in a fixed order.  Phase 3: coalesced store of the
//              rows' value segments.
//   * CTA    = a spatially compact set of rows; the geometry of every element they touch ("visit") is computed once
//              per CTA into shared memory (barycentric-gradient Gram matrix S = J grad(l_k).grad(l_l), which is
//              equivariant under the vertex re-labelling).
//
// The sum order of every nonzero is fixed by the plan, so the result is bit-reproducible run to run (atomics are not).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

// ---- device layout (v2) ------------------------------------------------------------------------------------------
// Every bundle is one contiguous, 16-byte aligned blob so that all its loads can be issued at once:
//   [units   : nchunks x 8 quads ]  u32 [chunk][lane]
//   [rowinfo : nrows quads       ]  {int64 base (CSR offset of the row's first value), int32 len (node level), int32 pad}
//   [sources : niter4 x 16 quads ]  u16, LANE-major: lane l owns entries [l*4*niter4, (l+1)*4*niter4), read 4 at a time
//   [outmap  : ceil(nout/8) quads]  u16 per output nonzero of the concatenated rows
// and a 16-byte header in a separate array (the next bundle's header is prefetched while the current one runs).
struct RowPlanHeader {           // 16 bytes
    uint32_t q0;                 // blob offset in quads (16 B)
    uint16_t niter4;             // phase-2 iterations / 4
    uint16_t nout;               // output nonzeros (node level)
    uint8_t nrows, nchunks, kind, flags;
    uint32_t pad;
};
#define ROWPLAN_FLAG_WARPDIAG 1  // single row: its diagonal is a full-warp reduction of the canonical entry

struct RowPlanCta {              // 16 bytes
    uint32_t visit_off, nvisit;  // visit element ids: i32 [visit_off + v]
    uint32_t bundle_off, nbundle;
};

// unit word: bits 0-9 visit | bits 10-27 six 3-bit indices of S01,S02,S03,S12,S13,S23 in the visit record | bit 30 active
#define ROWPLAN_UNIT_ACTIVE 0x40000000u
// source entry: staging slot + ROWPLAN_ZOFF (0 = the zero cell) | end-of-output flag (bit 15)
#define ROWPLAN_SRC_END 0x8000u
#define ROWPLAN_ZOFF 32
// outmap entry: bits 0-9 index into the warp's result buffer (0 = zero cell: structural zero, 1 = warp-reduced diagonal,
// 32 + task*32 + lane = task result) | bits 11-15 row slot
#define ROWPLAN_OMAP_ZERO 0u
#define ROWPLAN_OMAP_DIAG 1u

#define ROWPLAN_VMAX 640         // visits per CTA (< 1024)
#define ROWPLAN_WARPS 8          // warps per CTA
#define ROWPLAN_MAXCHUNKS 4      // a row may have up to 128 elements around its node
#define ROWPLAN_MAXROUNDS 23     // results per lane per bundle (32 + 23*32 + 31 < 1024)
#define ROWPLAN_BIGROW 12        // rows with more units are alone in their bundle (warp-reduced diagonal)

struct RowPlanHost {
    int order = 0, n = 0, ncomp = 1;
    int maxchunks = 1, maxrounds = 1;
    std::vector<RowPlanCta> ctas;
    std::vector<RowPlanHeader> hdrs;
    std::vector<int32_t> visits;
    std::vector<uint32_t> blob;  // 4 words per quad
    // statistics
    int64_t nunits = 0, ncontrib = 0, nsrc_padded = 0, nvisits = 0;
    std::string error;
};

struct RowPlan {                 // device mirror
    int order = 0, n = 0, ncomp = 1, maxchunks = 1, maxrounds = 1;
    int64_t ncta = 0, nbundle = 0;
    RowPlanCta* d_ctas = nullptr;
    RowPlanHeader* d_hdrs = nullptr;
    int32_t* d_visits = nullptr;
    uint32_t* d_blob = nullptr;
    int64_t bytes = 0;
    int64_t nunits = 0, ncontrib = 0, nsrc_padded = 0, nvisits = 0;
};

// canonical local dof of each kind and the vertex re-labelling tables (rowplan_build.cpp)
struct RowPlanRot {
    int kind;                    // 0 vertex, 1 edge dof, 2 face dof
    uint8_t perm[4];             // rotated-local vertex k = original-local vertex perm[k]
    uint8_t rmap[20];            // rotated-local dof j' -> original-local dof
};
void rowplan_rotation_tables(int order, RowPlanRot* rot /*[n]*/, int canon[3]);

// Builds the plan on the host.  pos = node-level element->CSR position map [nElt][n][n] (from the device),
// ia = CSR row pointers (dof level), rows are dof ids; nthreads <= 0: hardware concurrency.  Returns false (plan.error)
// when the mesh does not fit the plan's limits -- the caller then keeps the scatter path.
bool rowplan_build_host(int order, int ncomp, int64_t nElt, const int32_t* l2g, int nd, int64_t nPts,
                        const double* xyz, const int32_t* conn, int64_t nRows, const int64_t* ia,
                        const uint16_t* pos, int nthreads, RowPlanHost& plan);
