/* feelgpu.h -- C ABI of the B200 assembly engine.
 *
 * Drop-in boundary for the reference's element-range assembly path
 *     form2(_test=Xh,_trial=Xh) [+]= integrate(_range=elements(mesh), _expr=...)
 *     form1(_test=Xh)           [+]= integrate(_range=elements(mesh), _expr=...)
 * (feelpp/feel/feelvf/form.hpp:92-147, integrate.hpp:64-98).  A Feel++-side adapter intercepts
 * Integrator::assemble(form, MESH_ELEMENTS, same_mesh) (feelvf/integrator.hpp:975-1238) and calls these
 * entry points instead of running the per-element update/integrate/assemble loop (integrator.hpp:1093-1207).
 * See INTEGRATION.md for the adapter stub.
 *
 * Conventions: every function returns 0 on success or a FEELGPU_ERR_* code (no exceptions, no abort; the
 * reference aborts through CHKERRABORT, feelalg/matrixpetsc.cpp:935); feelgpu_last_error() gives a message.
 * Caller arrays passed to *_create are copied to the device during the call and may be freed afterwards.
 * All work calls are asynchronous on the CUDA stream given (void* = cudaStream_t, NULL = the context's stream).
 * Assemblies through one pattern must be ordered with respect to each other (one stream, or events): the pattern owns
 * the coefficient table and the patch counter of its assembly kernel.
 * There is no CPU fallback: without a CUDA device every call fails with FEELGPU_ERR_CUDA.
 */
#ifndef FEELGPU_H
#define FEELGPU_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define FEELGPU_VERSION 100

enum {
    FEELGPU_OK = 0,
    FEELGPU_ERR_ARG = 1,          /* bad argument                                                        */
    FEELGPU_ERR_CUDA = 2,         /* CUDA runtime error (no device, launch failure, out of memory)       */
    FEELGPU_ERR_NCCL = 3,         /* NCCL error                                                          */
    FEELGPU_ERR_NOT_IN_PATTERN = 4, /* a (row,col) of an element is absent from the CSR graph: the analogue of
                                     PETSc's MAT_NEW_NONZERO_ALLOCATION_ERR (feelalg/matrixpetsc.cpp:2688)  */
    FEELGPU_ERR_UNSUPPORTED = 5,  /* form / order / quadrature outside the supported set                 */
    FEELGPU_ERR_HOST = 6          /* host-side failure while building device data (out of host memory)   */
};

/* bilinear forms recognised on the path (feelvf/operators.hpp:134-153, operations.hpp:585-646, inner.hpp:374-420) */
enum {
    FEELGPU_FORM2_MASS = 0,       /* k * idt(u)*id(v)            / k*inner(idt(u),id(v)) on vector spaces         */
    FEELGPU_FORM2_STIFFNESS = 1,  /* k * gradt(u)*trans(grad(v)) / k*inner(gradt(u),grad(v))                      */
    FEELGPU_FORM2_ELASTICITY = 2  /* inner(lambda*trace(sym(gradt(u)))*Id + 2*mu*sym(gradt(u)), grad(v)), 3 comps */
};
/* linear forms (feelvf/linearformcontext.hpp:294-309) */
enum {
    FEELGPU_FORM1_SOURCE = 0      /* f*id(v) (scalar) / inner(f, id(v)) with a constant vector f (vector spaces)  */
};
/* coefficient kinds: cst(c), idv(k) for a Lagrange field, expr("...") (feelvf/detail/ginacexvf.hpp:750-848)     */
enum { FEELGPU_COEFF_CONST = 0, FEELGPU_COEFF_FIELD = 1, FEELGPU_COEFF_EXPR = 2 };

/* postfix program for closed-form coefficients, evaluated at xReal = G*N(xi_q) (feelpoly/geomap.hpp:1964-1969) */
enum {
    FEELGPU_OP_X = 0, FEELGPU_OP_Y = 1, FEELGPU_OP_Z = 2, FEELGPU_OP_CONST = 3,
    FEELGPU_OP_ADD = 4, FEELGPU_OP_SUB = 5, FEELGPU_OP_MUL = 6, FEELGPU_OP_DIV = 7, FEELGPU_OP_NEG = 8,
    FEELGPU_OP_SIN = 9, FEELGPU_OP_COS = 10, FEELGPU_OP_EXP = 11, FEELGPU_OP_LOG = 12, FEELGPU_OP_SQRT = 13,
    FEELGPU_OP_POW = 14
};
#define FEELGPU_EXPR_MAX_LEN 64

typedef struct feelgpu_context_s* feelgpu_context_t;
typedef struct feelgpu_mesh_s* feelgpu_mesh_t;
typedef struct feelgpu_space_s* feelgpu_space_t;
typedef struct feelgpu_pattern_s* feelgpu_pattern_t;
typedef struct feelgpu_faces_s* feelgpu_faces_t;
typedef struct feelgpu_graph_s* feelgpu_graph_t;

typedef struct {
    int kind;                     /* FEELGPU_COEFF_*                                                            */
    double value[3];              /* CONST: value[0]; form1 on vector spaces: the constant vector               */
    feelgpu_space_t field_space;  /* FIELD: scalar Lagrange space of the field (same mesh)                      */
    const double* field_values;   /* FIELD: DEVICE pointer, [nDofLocalWithGhost(field_space)]                    */
    int expr_len;                 /* EXPR: program length (<= FEELGPU_EXPR_MAX_LEN)                             */
    const int32_t* expr_ops;      /* EXPR: host pointer, opcodes                                                */
    const double* expr_consts;    /* EXPR: host pointer, constant per op slot (read for FEELGPU_OP_CONST)       */
    double lambda, mu;            /* ELASTICITY: Lame coefficients                                              */
} feelgpu_coeff;

/* reference tables the caller tabulated once on the host -- IMGeneral points/weights
 * (feelpoly/pointsetquadrature.hpp:142-153) and PreCompute phi/grad (feelpoly/polynomialset.hpp:924-987) */
typedef struct {
    int nq;                       /* quadrature points                                                          */
    const double* w;              /* [nq]                                                                       */
    const double* xi;             /* [nq][3] reference points                                                   */
    int n;                        /* scalar local dofs of the test=trial basis                                  */
    const double* phi;            /* [n][nq]                                                                    */
    const double* dphi;           /* [n][nq][3]                                                                 */
    int nf;                       /* FIELD coefficient: local dofs of the field basis (0 otherwise)             */
    const double* phif;           /* [nf][nq]                                                                   */
} feelgpu_tables;

typedef struct {
    int kind;                     /* FEELGPU_FORM2_*                                                            */
    feelgpu_coeff coeff;
    feelgpu_tables tables;
    int algorithm;                /* FEELGPU_ALGO_*                                                             */
} feelgpu_form2_desc;

typedef struct {
    int kind;                     /* FEELGPU_FORM1_*                                                            */
    feelgpu_coeff coeff;
    feelgpu_tables tables;
} feelgpu_form1_desc;

/* assembly algorithms for form2 */
enum {
    FEELGPU_ALGO_AUTO = 0,        /* the first plan of the pattern that covers the form: entity plan, row-thread plan, row plan; else scatter */
    FEELGPU_ALGO_SCATTER = 1,     /* element loop + red.global.add.f64 through the element->CSR-slot map         */
    FEELGPU_ALGO_ROWPLAN = 2,     /* owner-computes by (row, element) units: warp tile + planned gather          */
    FEELGPU_ALGO_EPLAN = 3,       /* owner-computes by (row, simplex) tasks: register accumulation over star(sigma)
                                     (constant-coefficient P1/P2 mass and stiffness, scalar and vector)               */
    FEELGPU_ALGO_RPLAN = 4        /* owner-computes by row threads: the thread that owns a CSR row sums the quadrature rows
                                     of the elements around it (cst / idv / expr coefficients, P1-P3, elasticity)      */
};
/* flags of feelgpu_pattern_create (0 = SLOTMAP + RPLAN + EPLAN for P1/P2 spaces).  FEELGPU_PATTERN_ASYNC: the entity plan,
 * whose construction takes seconds on meshes of millions of elements, is built by a background host thread; pattern_create
 * returns when the row-thread plan is ready, FEELGPU_ALGO_AUTO assembles through it until the entity plan is installed
 * (feelgpu_pattern_plans_ready tells / waits), FEELGPU_ALGO_EPLAN waits for it.  The mesh and the spaces must outlive the
 * pattern (as always).                                                                                                 */
enum { FEELGPU_PATTERN_SLOTMAP = 1, FEELGPU_PATTERN_ROWPLAN = 2, FEELGPU_PATTERN_EPLAN = 4, FEELGPU_PATTERN_RPLAN = 8,
       FEELGPU_PATTERN_ASYNC = 16 };

/* ---- context ------------------------------------------------------------------------------------------------ */
int feelgpu_context_create(int device, feelgpu_context_t* ctx);
int feelgpu_context_destroy(feelgpu_context_t ctx);
const char* feelgpu_last_error(feelgpu_context_t ctx);   /* ctx may be NULL: error of the failed context_create */
int feelgpu_sync(feelgpu_context_t ctx);

/* ---- mesh: element connectivity + vertex coordinates (replaces elt.point(i).id(), elt.G(): feelmesh/geond.hpp:626) */
int feelgpu_mesh_create(feelgpu_context_t ctx, int64_t nPts, const double* xyz /*[nPts][3]*/, int64_t nElt,
                        const int32_t* conn /*[nElt][4]*/, feelgpu_mesh_t* mesh);
/* new vertex coordinates (moving meshes, the per-step input of an end-to-end assembly): stream-ordered; page-locked
 * caller memory is read by the copy engine directly and must stay valid until the stream reaches the copy, pageable
 * memory is staged (the call then waits for the stream first) */
int feelgpu_mesh_update_points(feelgpu_mesh_t mesh, const double* xyz /*host [nPts][3]*/, void* stream);
int feelgpu_mesh_destroy(feelgpu_mesh_t mesh);

/* ---- space: flattened DofTable (replaces DofTable::localToGlobalIndices, feeldiscr/doftable.hpp:732-750) ------ */
int feelgpu_space_create(feelgpu_context_t ctx, feelgpu_mesh_t mesh, int order, int ncomp,
                         const int32_t* l2g /*[nElt][ncomp*n] process-local dof ids, local index = n*c+i*/,
                         int64_t nDofLocalWithGhost, int64_t nDofOwned,
                         const int64_t* proc2cluster /*[nDofLocalWithGhost] or NULL = identity (datamap.hpp:467)*/,
                         feelgpu_space_t* space);
int feelgpu_space_destroy(feelgpu_space_t space);

/* ---- pattern: the host-built CSR graph (GraphCSR::ia()/ja(), feelalg/graphcsr.hpp:287-295) ------------------- */
/* Builds on the device the element->CSR-slot map (and, with FEELGPU_PATTERN_ROWPLAN, the owner-computes plan).   */
/* Rows are process-local dof ids (owned first, then ghosts: doftable.hpp:1975-2030); columns are global-cluster. */
int feelgpu_pattern_create(feelgpu_context_t ctx, feelgpu_space_t test, feelgpu_space_t trial, int64_t nRows,
                           const int64_t* ia /*[nRows+1]*/, const int64_t* ja /*[nnz], ascending per row*/,
                           int flags, feelgpu_pattern_t* pattern);
int feelgpu_pattern_nnz(feelgpu_pattern_t pattern, int64_t* nnz);
/* background plan builds: *ready = 1 when none is in flight (wait != 0: block until then); *build_seconds = host time the
 * entity plan took.  Either output may be NULL.                                                                          */
int feelgpu_pattern_plans_ready(feelgpu_pattern_t pattern, int wait, int* ready, double* build_seconds);
int feelgpu_pattern_destroy(feelgpu_pattern_t pattern);

/* ---- work calls (replace the loop integrator.hpp:1093-1207 + MatSetValues/VecSetValues ADD_VALUES) ----------- */
/* accumulate = 0: operator=  (zero, then assemble: bilinearform.hpp:1368-1371); 1: operator+=                   */
int feelgpu_form2_assemble(feelgpu_pattern_t pattern, const feelgpu_form2_desc* desc, int accumulate,
                           double* a_values /*DEVICE [nnz], slot order == ja order*/, void* stream);
int feelgpu_form1_assemble(feelgpu_space_t space, const feelgpu_form1_desc* desc, int accumulate,
                           double* b /*DEVICE [nDofLocalWithGhost]*/, void* stream);
/* end-to-end variants with HOST output buffers (what the adapter calls when the PETSc Mat/Vec lives on the host):
 * assemble on the device, copy the values to a_host / b_host, synchronise.                                       */
int feelgpu_form2_assemble_host(feelgpu_pattern_t pattern, const feelgpu_form2_desc* desc, int accumulate,
                                double* a_host /*[nnz]*/);
int feelgpu_form1_assemble_host(feelgpu_space_t space, const feelgpu_form1_desc* desc, int accumulate,
                                double* b_host);
/* Allocation-free variant of feelgpu_form2_assemble for the forms whose kernels use scratch memory: MASS / STIFFNESS of a
 * scalar space with a point-dependent coefficient (FEELGPU_COEFF_FIELD, FEELGPU_COEFF_EXPR) through the row-thread plan
 * (the pattern must carry it: FEELGPU_PATTERN_RPLAN, the default).  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The coefficient values k(x_q) [nElt][nq], the parameter block and the reference tables live in ONE static
 * device array of the library (16.05 MiB, held by every GPU that loads the library until exit) instead of cudaMalloc'ed
 * buffers; the call issues no cudaMalloc / cudaFree and calls no allocating library routine.  Limit: nElt * tables.nq <=
 * FEELGPU_NOALLOC_MAX_KQ, otherwise FEELGPU_ERR_UNSUPPORTED with a message (use feelgpu_form2_assemble).  Patterns with an
 * exchange attached (multi-GPU) are refused.  Inputs, outputs and semantics are those of feelgpu_form2_assemble
 * (desc->algorithm is ignored); results agree with it within the tolerance of the row-thread plan (1e-12 relative to the
 * row's absolute sum; the order of the shared-memory adds is not fixed).  All calls share the static array: calls that
 * overlap on the GPU (different streams without ordering between them, or several host threads) race on it -- order them
 * on one stream or with events.  Nothing in the array has to start at zero: each call overwrites what its kernels read.  */
#define FEELGPU_NOALLOC_MAX_KQ 2097152   /* doubles of coefficient scratch: elements x quadrature points */
int feelgpu_form2_assemble_noalloc(feelgpu_pattern_t pattern, const feelgpu_form2_desc* desc, int accumulate,
                                   double* a_values /*DEVICE [nnz], slot order == ja order*/, void* stream);

/* ---- on(): Dirichlet elimination on the assembled matrix and right-hand side -----------------------------------------
 * form2 += on(_range=boundaryfaces(mesh), _rhs=l, _element=u, _expr=g) collects the boundary dofs and their values on the
 * host (feelvf/on.hpp:395-449) and calls MatrixPetsc::zeroRows (feelalg/matrixpetsc.cpp:1762-1860), i.e. PETSc's
 * MatZeroRows / MatZeroRowsColumns.  rows: process-local dof ids (host), values: the Dirichlet value of every row (host);
 * a_values and b are DEVICE arrays (b may be NULL).  `strategy` is the reference's ContextOn bit set, passed through
 * unchanged (feelalg/enums.hpp:56-66, string names feelalg/enums.cpp:28-35):
 *   ELIMINATION   0x1  rows zeroed, a_rr = d, b_r = d*x_r; REQUIRED, as MatrixPetsc::zeroRows does nothing without it
 *                      (matrixpetsc.cpp:1784); PENALISATION (0x0) stays on the host and is FEELGPU_ERR_UNSUPPORTED here
 *   KEEP_DIAGONAL 0x2  d = the assembled a_rr instead of value_on_diagonal
 *   SYMMETRIC     0x4  also the columns: b_i -= a_ir*x_r, a_ir = 0             (one space, identity cluster numbering)
 * rows/values are read during the call (any host memory, may be freed afterwards).  A row listed more than once keeps
 * its first value, as the reference's collection loop does (feelvf/on.hpp: std::find(dofs...) -> continue).         */
enum { FEELGPU_ON_PENALISATION = 0, FEELGPU_ON_ELIMINATION = 1, FEELGPU_ON_KEEP_DIAGONAL = 2, FEELGPU_ON_SYMMETRIC = 4 };
int feelgpu_form2_zero_rows(feelgpu_pattern_t pattern, int64_t n, const int32_t* rows, const double* values, int strategy,
                            double value_on_diagonal, double* a_values, double* b, void* stream);

/* ---- sparsity graph on the device -------------------------------------------------------------------------------
 * Replaces Stencil::computeGraph + GraphCSR::close for test space == trial space, pattern COUPLED
 * (feeldiscr/stencil.hpp:1500-1798, feelalg/graphcsr.cpp:662-1022): rows = dofs of the space (sequential numbering),
 * columns ascending as the reference's std::set keeps them, the diagonal always present (graphcsr.cpp:633-659).
 * Bit-exact with the host graph.  feelgpu_graph_get copies ia[nDof+1] / ja[nnz] (int64) to HOST arrays: what the adapter
 * hands to MatSeqAIJSetPreallocationCSR and back to feelgpu_pattern_create.  Either pointer may be NULL.            */
int feelgpu_graph_create(feelgpu_space_t space, feelgpu_graph_t* graph, int64_t* nnz);
int feelgpu_graph_get(feelgpu_graph_t graph, int64_t* ia, int64_t* ja);
int feelgpu_graph_destroy(feelgpu_graph_t graph);
/* The pattern of such a graph without the round trip of its columns through the host (feelgpu_pattern_create with the
 * arrays of feelgpu_graph_get gives the same pattern): test == trial == the graph's space, nRows = its dofs.  Only ia comes to
 * the host (the plan builders read it).  The graph may be destroyed once the call returns.                              */
int feelgpu_pattern_create_from_graph(feelgpu_graph_t graph, int flags, feelgpu_pattern_t* pattern);

/* ---- device hand-off: y = A x on the assembled CSR (MatrixPetsc::multVector -> MatMult, feelalg/matrixpetsc.cpp:998-1062;
 * with it MatrixPetsc::energy :2132-2170 is a dot product).  Sequential matrices (one space, identity cluster numbering);
 * x [nDof], y [nRows] DEVICE.  The matrix never leaves the GPU between assembly, on() and the products of a Krylov loop. */
int feelgpu_matvec(feelgpu_pattern_t pattern, const double* a_values, const double* x, double* y, void* stream);

/* ---- boundary-face integrals: integrate(_range=boundaryfaces(mesh) | markedfaces(mesh, ...), _expr=...) ------------
 * Replaces Integrator::assemble(..., mpl::int_<MESH_FACES>, ...) for faces with one adjacent element
 * (feelvf/integrator.hpp:3242-3704): Neumann terms g*id(v) / inner(g, id(v)) and Robin terms k*idt(u)*id(v).
 * A range is a list of (element id, local face 0..3) pairs: face.element(0).id(), face.pos_first().
 * Tables: the caller maps the triangle rule onto the 4 faces of the reference tet as the reference does
 * (feelpoly/pointsetquadrature.hpp:696-760): tables.nq = 4*Qf, block f = local face f; w already holds the reference-face
 * area factor; xi are points of the reference tet; phi = basis at those points; dphi unused.  Per face the kernel uses
 *     prod[q] = w[q] * J * |B n_f|          (pointsetquadrature.hpp:481-490, feelpoly/geomap.hpp:2021-2034)
 * Coefficients: CONST and EXPR.  accumulate = 0 zeroes the whole target first ('=' on a form).                        */
int feelgpu_faces_create(feelgpu_mesh_t mesh, int64_t nFaces, const int32_t* element, const int32_t* local_face,
                         feelgpu_faces_t* faces);
int feelgpu_faces_destroy(feelgpu_faces_t faces);
int feelgpu_form1_assemble_faces(feelgpu_space_t space, feelgpu_faces_t faces, const feelgpu_form1_desc* desc,
                                 int accumulate, double* b /*DEVICE*/, void* stream);
int feelgpu_form2_assemble_faces(feelgpu_pattern_t pattern, feelgpu_faces_t faces, const feelgpu_form2_desc* desc /*MASS*/,
                                 int accumulate, double* a_values /*DEVICE*/, void* stream);

/* ---- multi-GPU: one context (one MPI rank of the reference) per GPU of a node ------------------------------------------
 * The reference ships stashed off-process rows inside MatAssemblyBegin/End (feelalg/matrixpetsc.cpp:726-760; MatSetValues
 * on ghost rows :2875-2880) after GraphCSR::close has told every owner which (row, columns) its neighbours hold
 * (feelalg/graphcsr.cpp:800-951).  Here:
 *   feelgpu_comm_unique_id / feelgpu_comm_init   an NCCL communicator over the ranks (rank 0 creates the id, the caller
 *       ships the 128 bytes by its own means -- Feel++'s WorldComm / MPI_Bcast -- every rank calls comm_init: collective)
 *   feelgpu_exchange_create (collective, once per pattern)   the graph exchange.  Rows of the pattern are the process-local
 *       dofs in the reference's order: active dofs first, ghost dofs after, each group in first-touch order
 *       (feeldiscr/doftable.hpp:1975-1992) -- the ghost rows of one owner need NOT be contiguous.
 *       ghost_owner[k] = rank owning local row nDofOwned + k (DataMap::procOnGlobalCluster), first_dof_cluster =
 *       firstDofGlobalCluster() of this rank (feelalg/datamap.hpp:305); the cluster id of a ghost row comes from the
 *       space's proc2cluster map.  Builds the owner-side slot lists and, unless FEELGPU_EXCHANGE_NO_PEER_MEMORY, the
 *       peer-memory transport: every owner exports a landing buffer (CUDA IPC over NVLink/NVSwitch) and the assembly kernel
 *       of the entity plan stores a rank's ghost rows straight into their owners' memory while it runs.
 *   feelgpu_form2_close   MatrixPetsc::close (feelvf/bilinearformbase.hpp:353-357): stream-ordered; afterwards the owned
 *       rows hold the contributions of every rank.  Peer transport: one NCCL all-reduce of a word (completion barrier)
 *       + one owner-side add kernel; otherwise pack, ncclSend/ncclRecv (one group), add.  The library tracks which
 *       assemblies since the last close wrote where ('=' and '+=' keep their meaning across ranks).
 *   feelgpu_form1_close   VectorPetscMPI::close: VecAssemblyBegin/End (ghost entries added to their owners) and, with
 *       ghost_update != 0, VecGhostUpdate (ghost copies refreshed from the owners) (feelalg/vectorpetsc.cpp:1667-1693).
 * Work calls of all ranks must be issued in the same order (they contain collectives).                                 */
#define FEELGPU_UNIQUE_ID_BYTES 128
enum { FEELGPU_EXCHANGE_NO_PEER_MEMORY = 1 };
typedef struct feelgpu_exchange_s* feelgpu_exchange_t;
int feelgpu_comm_unique_id(unsigned char id[FEELGPU_UNIQUE_ID_BYTES]);
int feelgpu_comm_init(feelgpu_context_t ctx, const unsigned char id[FEELGPU_UNIQUE_ID_BYTES], int rank, int nranks);
int feelgpu_comm_destroy(feelgpu_context_t ctx);
int feelgpu_exchange_create(feelgpu_pattern_t pattern, const int32_t* ghost_owner /*[nRows - nDofOwned]*/,
                            int64_t first_dof_cluster, int flags, feelgpu_exchange_t* exchange);
int feelgpu_exchange_destroy(feelgpu_exchange_t exchange);
/* info: [0] owners of my ghost rows, [1] ghost rows, [2] ghost-row values sent per close, [3] ranks holding ghosts of my
 * rows, [4] rows received, [5] values received, [6] 1 = peer-memory transport, [7] closes so far (epoch)              */
int feelgpu_exchange_info(feelgpu_exchange_t exchange, int64_t info[8]);
int feelgpu_form2_close(feelgpu_pattern_t pattern, double* a_values /*DEVICE [nnz]*/, void* stream);
int feelgpu_form1_close(feelgpu_exchange_t exchange, double* b /*DEVICE [nDofLocalWithGhost]*/, int ghost_update, void* stream);

/* owner-side helper kept for callers that run their own transport: a[slots[k]] += vals[k], k < n (DEVICE arrays)      */
int feelgpu_scatter_add(feelgpu_context_t ctx, int64_t n, const int64_t* slots, const double* vals, double* a,
                        void* stream);

/* ---- device memory helpers for callers without a CUDA runtime of their own ---------------------------------- */
int feelgpu_malloc(feelgpu_context_t ctx, int64_t bytes, void** dptr);
int feelgpu_free(feelgpu_context_t ctx, void* dptr);
int feelgpu_memcpy_h2d(feelgpu_context_t ctx, void* dst, const void* src, int64_t bytes);
/* zero a device array, asynchronously on `stream` (MatrixPetsc::zero -> MatZeroEntries, feelalg/matrixpetsc.cpp:616-626) */
int feelgpu_memset_zero(feelgpu_context_t ctx, void* dptr, int64_t bytes, void* stream);
int feelgpu_memcpy_d2h(feelgpu_context_t ctx, void* dst, const void* src, int64_t bytes);
int feelgpu_host_alloc_pinned(feelgpu_context_t ctx, int64_t bytes, void** hptr);
int feelgpu_host_free_pinned(feelgpu_context_t ctx, void* hptr);

/* ---- introspection / measurement ----------------------------------------------------------------------------- */
/* number of kernel launches issued by this context since creation (bench.py's gpu_launches)                      */
int feelgpu_launch_count(feelgpu_context_t ctx, int64_t* count);
/* device milliseconds of the last feelgpu_form2_assemble on its stream (CUDA events), per kernel class:
 * ms[0] zero-fill, ms[1] main assembly kernel, ms[2] other                                                       */
int feelgpu_last_timing(feelgpu_context_t ctx, double ms[3]);
int feelgpu_set_timing(feelgpu_context_t ctx, int enable);
/* statistics of the owner-computes plan of a pattern: stats[0] = 1 if a plan exists, [1] CTAs, [2] bundles, [3] units
 * (row, element) pairs, [4] contributions summed in phase 2, [5] source entries incl. padding, [6] element visits
 * (geometry evaluations), [7] plan bytes on the device, [8] max chunks, [9] max rounds                                */
int feelgpu_pattern_plan_stats(feelgpu_pattern_t pattern, int64_t stats[10]);
/* statistics of the entity plan: stats[0] = 1 if it exists, [1] CTAs, [2] tasks, [3] task iterations (= loads of a unit
 * word), [4] iterations incl. padding, [5] element visits, [6] output values, [7] plan bytes on the device,
 * [8] modelled shared-memory wavefronts of the S loads, [9] the same without bank conflicts                          */
int feelgpu_pattern_eplan_stats(feelgpu_pattern_t pattern, int64_t stats[10]);
/* statistics of the row-thread plan: stats[0] = 1 if it exists, [1] patches (CTAs), [2] units (row, element), [3] element
 * visits, [4] plan bytes on the device, [5] largest row block (doubles), [6] most visits in a patch, [7] most rows in a
 * patch, [8] host build time in milliseconds                                                                          */
int feelgpu_pattern_rplan_stats(feelgpu_pattern_t pattern, int64_t stats[10]);


/* ---- host-side view of the entity plan (no device needed): what feelgpu_pattern_create builds for P1/P2 spaces ------
 * Used by the CPU tests to interpret the plan on the host and by tools that want to inspect it.
 * pos = element -> CSR position map [nElt][n][n] (position of column l2g[e][j] inside row l2g[e][i], component 0).
 * sizes: [0] CTAs, [1] vertex ids, [2] visit words, [3] chunks, [4] blob words (u32), [5] rows, [6] out-buffer slots,
 *        [7] visit slots per CTA                                                                                       */
int feelgpu_eplan_host_build(int order, int ncomp, int64_t nElt, const int32_t* l2g, int64_t nPts, const double* xyz,
                             const int32_t* conn, int64_t nRows, const int64_t* ia, const uint16_t* pos,
                             void** plan, int64_t sizes[8]);
int feelgpu_eplan_host_copy(void* plan, void* ctas /*32 B each*/, int32_t* verts, uint32_t* visits,
                            void* chunks /*8 B each*/, uint32_t* blob, void* rows /*16 B each*/, double stats[8]);
const char* feelgpu_eplan_host_error(void* plan);
int feelgpu_eplan_host_free(void* plan);
/* coefficients the kernel applies to the per-task sums, from the caller's reference tables:
 * coef[type][out][load] (6 x 3 x 5), mass[type][out] (6 x 3); form: 0 mass, 1 stiffness                              */
int feelgpu_eplan_host_coefficients(int order, const feelgpu_tables* tables, int form, double* coef, double* mass);

/* ---- host-side view of the row-thread plan (no device needed): what feelgpu_pattern_create builds with
 * FEELGPU_PATTERN_RPLAN; used by the CPU tests to interpret the plan on the host.  sizes: [0] patches, [1] visits,
 * [2] node rows, [3] units, [4] largest row block, [5] most visits in a patch, [6] most rows in a patch.
 * patches: {u32 visit_off, nvisit, row_off, nrows, unit_off, nunits, nslots, pad}; rows: {i64 base, u32 off, len, unit0,
 * nunits, pad, pad}; unit word: visit index in the patch | local dof << 12 | row in the patch << 17, stored in round order
 * (the k-th unit of every row that has one, k = 0, 1, ...)                                                              */
int feelgpu_rplan_host_build(int order, int ncomp, int64_t nElt, const int32_t* l2g, int64_t nPts, const double* xyz,
                             const int32_t* conn, int64_t nRows, const int64_t* ia, int maxslots, int maxvisits, int maxrows,
                             void** plan, int64_t sizes[8]);
int feelgpu_rplan_host_copy(void* plan, void* patches, int32_t* visits, void* rows, uint32_t* units);
const char* feelgpu_rplan_host_error(void* plan);
int feelgpu_rplan_host_free(void* plan);

#ifdef __cplusplus
}
#endif
#endif
