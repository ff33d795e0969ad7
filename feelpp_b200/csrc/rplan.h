// rplan.h -- the row-thread assembly plan ("rplan"): owner-computes assembly for ANY supported form.
//
// The reference adds every element's local matrix into the CSR with MatSetValues(ADD_VALUES)
// (feelalg/matrixpetsc.cpp:917-936).  The entity plan (eplan.h) turns that into register accumulation for forms whose
// local matrix is a fixed linear combination of a few per-element numbers (constant-coefficient P1/P2 mass/stiffness).
// Variable coefficients (expr("..."), idv(k)), elasticity and P3 have no such structure: their rows have to be computed
// by quadrature per (element, row).  The row-thread plan keeps the owner-computes property for them:
//
//   patch = spatially compact set of node rows (consecutive along a Hilbert curve) whose CSR values fit the CTA's
//           shared-memory row block.
//   visit = an element touching a row of the patch.  Phase A (thread = visit): vertex coordinates -> K, J, B
//           (feelpoly/geomap.hpp:2320-2355,1971-1991) and prod[q]*k(x_q) = w_q J k(x_q) at every quadrature point
//           (feelpoly/pointsetquadrature.hpp:481-490) -> shared memory.  A point-dependent coefficient (idv / expr at xReal)
//           is evaluated once per (element, point) by a small kernel in front of the assembly kernel.
//   unit  = (row r, element e containing r's node): the row of e's local matrix that belongs to r
//           (feelvf/bilinearformcontext.hpp:585-589).  Phase B: one thread per unit (per (unit, column j) for the 3x3 blocks
//           of product spaces) forms the unit's row by the quadrature sum (feelpoly/pointsetquadrature.hpp:403-416) and adds
//           it into row r's segment of the row block through the element->CSR position map.  Units are dealt in round
//           order (the k-th unit of every row), so all lanes are busy whatever the row lengths (a vertex row has ~24 units, a
//           face row 2) and the units of a round belong to different rows; units
//           of one row that run at the same time meet on the columns they share, hence the adds are shared-memory atomics
//           (a compare-and-swap loop: the hardware has no fp64 shared atomic) -- inside the SM, never in L2.  Adding round
//           by round with a barrier in between (plain read-modify-writes, fixed order) was measured: 39 % of the kernel went
//           into the barriers.  Every CSR value leaves the SM exactly once; the order of the <= ~30 additions into a value
//           is not fixed, as with the reference's own MPI runs (values within 1e-12, not bit-identical run to run).
//   Phase C: the row block is streamed to the CSR value array (coalesced, one contiguous piece per node row).
//
// The row-thread name stays from the first version (thread = row), which left the SM waiting for its longest row.
// The plan is small (one 32-bit word per unit, one element id per visit, 32 bytes per node row) and is built in one pass
// over the node -> element adjacency, so its setup cost is a fraction of the entity plan's.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

struct RPlanPatch {              // 32 bytes
    uint32_t visit_off, nvisit;  // element ids: int32 visits[visit_off + v]
    uint32_t row_off, nrows;     // node rows: RPlanRow rows[row_off + r], sorted by descending unit count
    uint32_t unit_off, nunits;   // unit words: uint32 units[unit_off + k]
    uint32_t nslots;             // row-block slots used
    uint32_t pad;
};
struct RPlanRow {                // 32 bytes
    int64_t base;                // CSR offset of the node's component-0 row
    uint32_t off;                // offset of that row in the patch's row block
    uint32_t len;                // dof-level row length (the ncomp rows of a node have the same length and follow each other)
    uint32_t unit0, nunits;      // units of this row: units[patch.unit_off + unit0 ...]
    uint32_t pad[2];
};
static_assert(sizeof(RPlanPatch) == 32 && sizeof(RPlanRow) == 32, "read as two uint4");
// unit word: visit index inside the patch (bits 0-11) | local dof i of the row's node in that element (bits 12-16) | row
// index inside the patch (bits 17-31).  Units are stored in round order: the k-th unit of every row that has one, k = 0, 1, ...

struct RPlanHost {
    int order = 0, n = 0, ncomp = 1;
    int maxslots = 0, maxvisits = 0, maxrows = 0;
    std::vector<RPlanPatch> patches;
    std::vector<int32_t> visits;
    std::vector<RPlanRow> rows;
    std::vector<uint32_t> units;
    int64_t nunits = 0, nvisits = 0;
    int maxpatch_visits = 0, maxpatch_rows = 0, maxpatch_slots = 0;
    std::string error;
};

struct RPlan {                   // device mirror
    int order = 0, n = 0, ncomp = 1;
    int maxslots = 0, maxvisits = 0, maxrows = 0;
    int64_t npatch = 0;
    RPlanPatch* d_patches = nullptr;
    int32_t* d_visits = nullptr;
    RPlanRow* d_rows = nullptr;
    uint32_t* d_units = nullptr;
    double* d_kq = nullptr;      // [nElt][nq] coefficient at the quadrature points, filled by the coefficient pass of an assembly
    size_t kq_bytes = 0;
    int64_t bytes = 0, nunits = 0, nvisits = 0;
    int maxpatch_visits = 0, maxpatch_rows = 0, maxpatch_slots = 0;
    double build_seconds = 0;
};

// Builds the plan on the host.  l2g = [nElt][nd] (component-major local numbering, node = l2g[e][i] / ncomp), ia = CSR row
// pointers (dof level) of the nRows local rows.  maxslots / maxvisits / maxrows bound a patch (row-block doubles, visit
// records, node rows).  Returns false (plan.error) when a single row does not fit.
bool rplan_build_host(int order, int ncomp, int64_t nElt, const int32_t* l2g, int nd, int64_t nPts, const double* xyz,
                      const int32_t* conn, int64_t nRows, const int64_t* ia, int maxslots, int maxvisits, int maxrows,
                      int nthreads, RPlanHost& plan);
