// feelgpu_internal.h -- device-side data layout shared by the C-ABI implementation and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/feelgpu.h"

#define FEELGPU_MAX_NLOC 20   // P3 scalar local dofs
#define FEELGPU_MAX_NQ 64

struct feelgpu_context_s {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string last_error;
    int64_t launches = 0;
    bool timing = false;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    double last_ms[3] = {0, 0, 0};
    int sm_count = 148;
    int* d_flag = nullptr;       // device error flag (slot-map build)
    struct TableEntry { std::vector<char> host; void* dev; };
    std::vector<TableEntry> table_cache;   // parameter blocks + reference tables already on the device
    // multi-GPU (exchange.cu): the NCCL communicator of feelgpu_comm_init (ncclComm_t kept opaque here)
    void* comm = nullptr;
    int rank = 0, nranks = 1;
    int* d_barrier = nullptr;    // the word all-reduced by close() as a completion barrier
};

struct feelgpu_mesh_s {
    feelgpu_context_t ctx;
    int64_t nPts = 0, nElt = 0;
    double* d_xyz = nullptr;     // [nPts][4]: x,y,z,0 -- one 32-byte sector per vertex
    int4* d_conn = nullptr;      // [nElt]
    double* d_raw = nullptr;     // [nPts][3]: the caller's layout as it arrives; k_pad_points writes d_xyz from it
    double* h_stage = nullptr;   // pinned staging for update_points from pageable caller memory
    cudaEvent_t stage_event = nullptr;   // recorded after every coordinate upload (h_stage / d_raw are reused across streams)
    bool stage_busy = false;
    std::vector<double> h_xyz;   // host copies: the plan builder orders rows along a space-filling curve
    std::vector<int32_t> h_conn;
};

struct feelgpu_space_s {
    feelgpu_context_t ctx;
    feelgpu_mesh_t mesh;
    int order = 1, ncomp = 1, n = 4, nd = 4;
    int64_t nDofWithGhost = 0, nDofOwned = 0;
    int32_t* d_l2g = nullptr;    // [nElt][nd]
    int64_t* d_p2c = nullptr;    // process -> cluster ids, or null (identity)
    std::vector<int32_t> h_l2g;  // host copy for the plan builder
    std::vector<int64_t> h_p2c;  // host copy (ghost rows name their owner-side row by cluster id)
};

struct RowPlan;  // kernels_rowplan.cu
struct EPlan;    // kernels_eplan.cu
struct RPlan;    // kernels_rplan.cu

struct feelgpu_pattern_s {
    feelgpu_context_t ctx;
    feelgpu_space_t test, trial;
    int64_t nRows = 0, nnz = 0;
    int64_t h_ia_owned_end = 0;  // ia[nDofOwned]: first value slot of the ghost-row tail
    int64_t* d_ia = nullptr;     // [nRows+1]
    int32_t* d_ja32 = nullptr;   // [nnz] column ids (kept for the Dirichlet elimination); null if an id needs more than 31 bits
    // element -> CSR position map, node level: pos[e][i][j] = position of column cluster(l2g_trial[e][j]) (component 0)
    // inside row l2g_test[e][i] (component 0).  Full slot = ia[row + c1] + pos + c2 (components are interleaved:
    // global = ncomp*node + c, feeldiscr/doftable.hpp:1289-1299).
    uint16_t* d_pos = nullptr;   // [nElt][n][n]
    double* d_stage = nullptr;   // [nnz] device staging of feelgpu_form2_assemble_host, allocated on first use
    RowPlan* plan = nullptr;
    std::string plan_note;       // why no owner-computes plan exists (if none)
    EPlan* eplan = nullptr;
    std::string eplan_note;
    struct EPlanAsync* eplan_async = nullptr;   // a build in flight (FEELGPU_PATTERN_ASYNC) or waiting to be installed
    double eplan_build_seconds = 0;
    RPlan* rplan = nullptr;      // row-thread plan: owner-computes path of the forms the entity plan does not cover
    std::string rplan_note;
    // multi-GPU: host copies of what the graph exchange ships (interface-sized) and the exchange itself
    std::vector<int64_t> h_ia;        // [nRows+1]
    std::vector<int64_t> h_ja_tail;   // columns of the ghost rows: ja[ia[nDofOwned] .. nnz)
    struct feelgpu_exchange_s* exchange = nullptr;
};

// parameter block copied to the device per assemble call
struct feelgpu_faces_s {
    feelgpu_mesh_t mesh = nullptr;
    int64_t n = 0;
    int32_t* d_elt = nullptr;    // [n] element of the face
    int32_t* d_lface = nullptr;  // [n] local face id 0..3
};

struct feelgpu_graph_s {         // the COUPLED graph of (space, space) on the device (kernels_graph.cu)
    feelgpu_space_t space = nullptr;
    int64_t nRows = 0, nnz = 0;
    int64_t* d_ia = nullptr;   // [nRows + 1]
    int64_t* d_ja = nullptr;   // [nnz]
};

struct FormParams {
    int kind, coeff_kind, n, nq, nf, ncomp, expr_len, pad;
    double cval[3];
    double lambda, mu;
    double w[FEELGPU_MAX_NQ];
    double xi[FEELGPU_MAX_NQ * 3];
    int32_t expr_ops[FEELGPU_EXPR_MAX_LEN];
    double expr_consts[FEELGPU_EXPR_MAX_LEN];
};

#define FEELGPU_CHECK_CUDA(ctx, call)                                                                   \
    do {                                                                                                \
        cudaError_t err__ = (call);                                                                     \
        if (err__ != cudaSuccess) {                                                                     \
            (ctx)->last_error = std::string(#call) + ": " + cudaGetErrorString(err__);                  \
            return FEELGPU_ERR_CUDA;                                                                    \
        }                                                                                               \
    } while (0)

// host threads a plan builder may use: FEELGPU_BUILD_THREADS, else the hardware threads shared among the ranks that build
// at the same time on this node (after feelgpu_comm_init: min(ranks of the communicator, devices of the node))
int feelgpu_builder_threads(feelgpu_context_t ctx);
// ---- kernel launchers (kernels_scatter.cu) ----
int feelgpu_launch_build_pos(feelgpu_pattern_t pat, const int64_t* d_ja, cudaStream_t s);
int feelgpu_launch_narrow_columns(feelgpu_pattern_t pat, const int64_t* d_ja, cudaStream_t s);
int feelgpu_launch_form2_scatter(feelgpu_pattern_t pat, const FormParams& hp, const feelgpu_form2_desc* desc,
                                 double* a, cudaStream_t s);
int feelgpu_launch_form1(feelgpu_space_t sp, const FormParams& hp, const feelgpu_form1_desc* desc, double* b,
                         cudaStream_t s);
int feelgpu_launch_form1_faces(feelgpu_space_t sp, feelgpu_faces_t fc, const FormParams& hp, const feelgpu_form1_desc* desc,
                               double* b, cudaStream_t s);
int feelgpu_launch_form2_faces(feelgpu_pattern_t pat, feelgpu_faces_t fc, const FormParams& hp,
                               const feelgpu_form2_desc* desc, double* a, cudaStream_t s);
// ---- kernels_on.cu ----
int feelgpu_launch_matvec(feelgpu_pattern_t pat, const double* a, const double* x, double* y, cudaStream_t s);
int feelgpu_launch_zero_rows(feelgpu_pattern_t pat, int n, const int32_t* h_rows, const double* h_x, int strategy,
                             double value_on_diagonal, double* a, double* b, cudaStream_t s);
// ---- kernels_rowplan.cu ----
int feelgpu_rowplan_build(feelgpu_pattern_t pat, const int64_t* h_ia, const int64_t* h_ja);
void feelgpu_rowplan_destroy(feelgpu_pattern_t pat);
int feelgpu_launch_form2_rowplan(feelgpu_pattern_t pat, const FormParams& hp, const feelgpu_form2_desc* desc,
                                 int accumulate, double* a, cudaStream_t s);
bool feelgpu_rowplan_supports(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc);
// ---- kernels_rplan.cu ----
int feelgpu_rplan_build(feelgpu_pattern_t pat, const int64_t* h_ia);
void feelgpu_rplan_destroy(feelgpu_pattern_t pat);
int feelgpu_launch_form2_rplan(feelgpu_pattern_t pat, const FormParams& hp, const feelgpu_form2_desc* desc,
                               int accumulate, double* a, cudaStream_t s);
bool feelgpu_rplan_supports(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc);
int feelgpu_rplan_stats(feelgpu_pattern_t pat, int64_t stats[10]);
// allocation-free variant (feelgpu_form2_assemble_noalloc): scratch in a static device array of kernels_rplan.cu
bool feelgpu_rplan_noalloc_supports(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc);
int feelgpu_launch_form2_rplan_noalloc(feelgpu_pattern_t pat, const FormParams& hp, const feelgpu_form2_desc* desc,
                                       int accumulate, double* a, cudaStream_t s);
// ---- kernels_eplan.cu ----
int feelgpu_eplan_build(feelgpu_pattern_t pat, const int64_t* h_ia, bool async);
int feelgpu_eplan_poll(feelgpu_pattern_t pat, bool wait);
void feelgpu_eplan_destroy(feelgpu_pattern_t pat);
int feelgpu_launch_form2_eplan(feelgpu_pattern_t pat, const FormParams& hp, const feelgpu_form2_desc* desc,
                               int accumulate, double* a, cudaStream_t s);
bool feelgpu_eplan_supports(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc);
int feelgpu_eplan_stats(feelgpu_pattern_t pat, int64_t stats[10]);
// ghost rows of the plan go to landing buffers in their owners' memory: per ghost dof row (local row nDofOwned + g) the
// neighbour index and the offset (in doubles) inside that neighbour's landing buffer; ptr[epoch][neighbour] = the buffers.
// n = 0 restores local stores.
int feelgpu_eplan_set_remote(feelgpu_pattern_t pat, int n, const int32_t* h_ghost_nb, const int64_t* h_ghost_off,
                             double* const* ptr_epoch0, double* const* ptr_epoch1);
void feelgpu_eplan_remote_epoch(feelgpu_pattern_t pat, int epoch);   // which set the next launches write; the set counts as empty
void feelgpu_eplan_remote_fresh(feelgpu_pattern_t pat);              // the current set counts as empty again ('=' on the form)
// ---- exchange.cu ----
void feelgpu_exchange_note_assembly(feelgpu_pattern_t pat, bool through_eplan, int accumulate, double* a, cudaStream_t s);
void feelgpu_exchange_detach(feelgpu_pattern_t pat);
