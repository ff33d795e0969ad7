// eplan.h -- the entity-task assembly plan ("eplan"): owner-computes assembly without a shared-memory transpose.
//
// The reference adds every element's local matrix into the CSR with MatSetValues(ADD_VALUES)
// (feelalg/matrixpetsc.cpp:917-936).  For Lagrange dofs attached to simplices (vertex / edge / face) the nonzero (r,c)
// receives a contribution from exactly the elements that contain both entities, i.e. from star(sigma) with
// sigma = entity(r) U entity(c).  Grouping the contributions by
//
//     task = (row r, simplex sigma containing entity(r))
//
// gives every task a STATIC set of output nonzeros (all c with entity(r) U entity(c) = sigma) and a list of elements
// (star(sigma)) that all contribute to all of them.  One thread runs one task: it walks its elements, accumulates in
// registers and stores each finished CSR value exactly once.  Per element it only reads the few entries
// S_kl = c J grad(lambda_k).grad(lambda_l) it needs (stiffness) or J (mass) from the CTA's shared-memory geometry
// cache; the element is re-labelled per task so that entity(r) and sigma occupy its first local vertices (the Lagrange
// nodal set is invariant under vertex permutations), which makes the arithmetic static.
//
//   CTA   = spatially compact set of rows.  Phase A: coordinates of the unique vertices of all elements touching the
//           rows -> shared memory; S (10 doubles) per element visit -> shared memory.  Phase B: task chunks (32 tasks
//           of one type, lane = task).  Phase C: the CTA's row block is streamed from shared memory to the CSR value
//           array with coalesced stores.  No atomics, no memset; the summation order is fixed by the plan.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#ifndef EPLAN_VS
#define EPLAN_VS 512             // visit slots per CTA (the last one is the all-zero record)
#endif
#define EPLAN_VERTS 256          // unique vertices per CTA (u8 local vertex ids)
#ifndef EPLAN_WARPS
#define EPLAN_WARPS 18              // 2 CTAs x 18 warps per SM at 56 registers: measured 1 % faster than 16 warps at 64
#endif
#ifndef EPLAN_MINB
#define EPLAN_MINB 2               // CTAs per SM the kernel is compiled for
#endif
#define EPLAN_ROWS 384            // rows per CTA: the row table reuses the 6 KB vertex cache after phase A
#ifndef EPLAN_OMAX
#define EPLAN_OMAX 7936          // 62 KB row block per CTA: two CTAs per SM next to the 40 KB geometry cache
#endif
#define EPLAN_HDRS 128            // chunk headers kept in shared memory (the rest, if any, is read from global memory)
#define EPLAN_IDX_SHIFT 12          // unit word: bits 0-8 visit slot, 4-bit S indices from bit 12 (index k at 12 + 4k): with the S
                                   // array on a 64 KB boundary of shared memory, (w >> 4k) & 0xF000 | base is the address of load k
#define EPLAN_PADRES 16            // row-block slots the packing pass holds back for phase pads (see eplan_build.cpp)
#define EPLAN_MAXREMOTE 16         // owners a rank's ghost rows can belong to (one node: up to 7; room for more)
#define EPLAN_REMOTE_FLAG (int64_t(1) << 62)   // row-table base of a ghost row that lands in an owner's buffer
#ifndef EPLAN_NSUB
#define EPLAN_NSUB 1               // short chunks run under one header: measured slower on B200 (coarser hand-out, deeper word stream); kept at 1
#endif
#define EPLAN_MAXTYPES 6
#define EPLAN_MAXLD 5
#define EPLAN_MAXOUT 3

// per (order, type): what a task loads per iteration and what it outputs, in re-labelled local numbering
struct EPlanType {
    int nrho, nsig;              // vertices of the row's entity / of sigma
    int nld;                     // S entries loaded per iteration
    int ld[EPLAN_MAXLD][2];      // (k,l), k <= l, re-labelled local vertices
    int nout;
    int outdof[EPLAN_MAXOUT];    // re-labelled local dof of every output column
    int rowdof;                  // re-labelled local dof of the row
    int phaseA;                  // 1: single-element tasks whose outputs are stored by phase A (sigma = the whole tet)
};
struct EPlanSpec {
    int order, n, ntypes;
    EPlanType type[EPLAN_MAXTYPES];
};
const EPlanSpec* eplan_spec(int order);   // nullptr if the order has no entity plan

// index of S_kl in the 10-entry visit record: diagonals 0..3, then (0,1),(0,2),(0,3),(1,2),(1,3),(2,3)
inline int eplan_sidx(int k, int l)
{
    if (k == l) return k;
    if (k > l) { int t = k; k = l; l = t; }
    return 4 + (k == 0 ? l - 1 : (k == 1 ? l + 1 : 5));
}

struct EPlanCta {                // 32 bytes
    uint32_t vert_off;           // unique vertex ids: i32 [vert_off + i]
    uint32_t visit_off;          // visit words (P1: 1 per visit, P2: 4): 4 x u8 local vertex index [, 6 x u16 row-block slots of the
                                 // outputs (edge le, opposite edge), le = original local edge of the row]
    uint32_t chunk_off;          // chunk headers
    uint32_t row_off;            // row table
    uint16_t nverts, nvisit;
    uint16_t nchunk, nrows;
    uint16_t nout;               // out-buffer slots used (incl. slot 0 = trash); bit 15: zero the row block first
    uint16_t blob_nq;            // length of this CTA's chunk data in quads (16 B), for the L2 prefetch
    uint32_t blob_q0;            // its first quad
};
static_assert(sizeof(EPlanCta) == 32, "EPlanCta is read as two uint4");
static_assert(EPLAN_ROWS * 16 <= 3 * EPLAN_VERTS * 8 && EPLAN_ROWS <= EPLAN_WARPS * 32, "row table reuses the vertex cache");
struct EPlanChunk {              // 8 bytes
    uint32_t q0;                 // offset of the chunk data in quads (16 B): words u32 [nsub*niter][32], then slots u16 [nsub*nout][32]
    uint8_t type, niter;
    uint16_t pad;                // nsub: sub-chunks (32 tasks each) run back to back
};
struct EPlanRow {                // 16 bytes
    int64_t base;                // CSR offset of the row's first value (dof level, component 0 row)
    uint32_t off;                // first slot in the CTA's out buffer
    uint32_t len;                // node-level row length
};

// allocator whose resize() leaves new elements uninitialised: the 2.6 GB blob of a large plan is filled by the builder's
// threads (first touch in parallel) instead of being zeroed by one of them first
template <class T>
struct EPlanNoInit : std::allocator<T> {
    template <class U> struct rebind { using other = EPlanNoInit<U>; };
    EPlanNoInit() = default;
    template <class U> EPlanNoInit(const EPlanNoInit<U>&) {}
    template <class U> void construct(U* p) noexcept { ::new (static_cast<void*>(p)) U; }
    template <class U, class... A> void construct(U* p, A&&... a) { ::new (static_cast<void*>(p)) U(std::forward<A>(a)...); }
};

struct EPlanHost {
    int order = 0, n = 0, ncomp = 1, omax = 0;
    std::vector<EPlanCta> ctas;
    std::vector<int32_t> verts;
    std::vector<uint32_t> visits;
    std::vector<EPlanChunk> chunks;
    std::vector<uint32_t, EPlanNoInit<uint32_t>> blob;
    std::vector<EPlanRow> rows;
    // statistics
    int64_t ntasks = 0, niters = 0, niters_padded = 0, nvisits = 0, nouts = 0;
    int64_t nomerge_from = INT64_MAX;   // input: rows whose values start at or after this slot stay one segment each (ghost rows: their segments may be redirected to different owners)
    double lds_wavefronts = 0;   // modelled shared-memory wavefronts of the S loads (stiffness) over the whole plan
    double lds_ideal = 0;        // the same without bank conflicts
    double a1_passes = 0, a1_passes_before = 0;   // modelled passes of one coordinate read of phase A1 (per half-warp and corner), after / before the vertex slots were levelled
    std::string error;
};

struct EPlan {                   // device mirror
    int order = 0, n = 0, ncomp = 1, omax = 0;
    int64_t ncta = 0;
    EPlanCta* d_ctas = nullptr;
    int32_t* d_verts = nullptr;
    uint32_t* d_visits = nullptr;
    EPlanChunk* d_chunks = nullptr;
    uint32_t* d_blob = nullptr;
    EPlanRow* d_rows = nullptr;
    // ghost rows written straight into their owners' memory (feelgpu_eplan_set_remote): the row-table entries of the ghost
    // rows carry (neighbour, offset) instead of a CSR offset: base = EPLAN_REMOTE_FLAG | neighbour << 56 | offset
    int nremote = 0;
    bool remote_fresh = true;    // no assembly has written the current landing buffers yet: the first one stores, later ones may add
    int remote_epoch = 0;
    double* remote_ptr[2][EPLAN_MAXREMOTE];
    int64_t nrowent = 0;         // entries of d_rows
    EPlanRow* d_rows_local = nullptr;   // the row table as built (local CSR offsets), kept to patch from
    int* d_sched = nullptr;      // [2] patch counter and exit counter of the persistent kernel (zero between launches)
    int64_t bytes = 0;
    int64_t ntasks = 0, niters = 0, niters_padded = 0, nvisits = 0, nouts = 0;
    double lds_wavefronts = 0, lds_ideal = 0;
};

// Builds the plan on the host.  pos = element -> CSR position map [nElt][n][n] (dof-level position of the component-0
// column inside the component-0 row), ia = CSR row pointers (dof level).  Returns false (plan.error) when the mesh does
// not fit the plan's limits -- the caller then keeps another path.
bool eplan_build_host(int order, int ncomp, int omax, int64_t nElt, const int32_t* l2g, int nd, int64_t nPts,
                      const double* xyz, const int32_t* conn, int64_t nRows, const int64_t* ia, const uint16_t* pos,
                      int nthreads, EPlanHost& plan);
