// feelgpu_api.cu -- implementation of the C ABI declared in include/feelgpu.h.
#include "feelgpu_internal.h"
#include "rowplan.h"
#include "eplan.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <new>

static std::string g_create_error;

#define CTX_CUDA(ctx, call) FEELGPU_CHECK_CUDA(ctx, call)
#define ARG_FAIL(ctx, msg)             \
    do {                               \
        (ctx)->last_error = (msg);     \
        return FEELGPU_ERR_ARG;        \
    } while (0)

// No exception crosses the ABI (include/feelgpu.h): the create calls build std::vector-based host plans, so they run
// inside this guard.  `cleanup` releases what the body had allocated when the exception was thrown.
template <class Body, class Cleanup>
static int abi_guard(feelgpu_context_t ctx, const char* what, Body body, Cleanup cleanup)
{
    try {
        return body();
    } catch (const std::exception& e) {
        cleanup();
        ctx->last_error = std::string(what) + ": " + e.what();
    } catch (...) {
        cleanup();
        ctx->last_error = std::string(what) + ": unknown host exception";
    }
    return FEELGPU_ERR_HOST;
}

extern "C" {

int feelgpu_context_create(int device, feelgpu_context_t* out)
{
    if (!out) return FEELGPU_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    cudaError_t err = cudaGetDeviceCount(&ndev);
    if (err != cudaSuccess || ndev == 0) {
        g_create_error = std::string("no CUDA device: ") + (err != cudaSuccess ? cudaGetErrorString(err) : "device count is 0");
        return FEELGPU_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) {
        g_create_error = "device index out of range";
        return FEELGPU_ERR_ARG;
    }
    feelgpu_context_t ctx = new (std::nothrow) feelgpu_context_s;
    if (!ctx) return FEELGPU_ERR_ARG;
    ctx->device = device;
    if ((err = cudaSetDevice(device)) != cudaSuccess || (err = cudaStreamCreate(&ctx->stream)) != cudaSuccess ||
        (err = cudaMalloc(&ctx->d_flag, sizeof(int))) != cudaSuccess) {
        g_create_error = cudaGetErrorString(err);
        if (ctx->stream) cudaStreamDestroy(ctx->stream);
        delete ctx;
        return FEELGPU_ERR_CUDA;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) ctx->sm_count = prop.multiProcessorCount;
    for (int k = 0; k < 4; ++k) cudaEventCreate(&ctx->ev[k]);
    *out = ctx;
    return FEELGPU_OK;
}

int feelgpu_context_destroy(feelgpu_context_t ctx)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    feelgpu_comm_destroy(ctx);
    for (auto& e : ctx->table_cache) cudaFree(e.dev);
    for (int k = 0; k < 4; ++k)
        if (ctx->ev[k]) cudaEventDestroy(ctx->ev[k]);
    if (ctx->d_flag) cudaFree(ctx->d_flag);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return FEELGPU_OK;
}

const char* feelgpu_last_error(feelgpu_context_t ctx) { return ctx ? ctx->last_error.c_str() : g_create_error.c_str(); }

int feelgpu_sync(feelgpu_context_t ctx)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    CTX_CUDA(ctx, cudaDeviceSynchronize());
    return FEELGPU_OK;
}

// ---- mesh ------------------------------------------------------------------------------------------------------
// [nPts][3] (the caller's layout, elt.G() columns) -> [nPts][4], one 32-byte sector per vertex
static __global__ void k_pad_points(int64_t nPts, const double* __restrict__ raw, double* __restrict__ xyz)
{
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < nPts; p += (int64_t)gridDim.x * blockDim.x) {
        const double x = raw[3 * p], y = raw[3 * p + 1], z = raw[3 * p + 2];
        reinterpret_cast<double4*>(xyz)[p] = make_double4(x, y, z, 0.0);
    }
}

// Coordinates travel in the caller's 24-byte layout and are padded on the device.  Page-locked caller memory is read
// by the copy engine directly; pageable memory goes through the mesh's pinned staging buffer (one memcpy).
static int upload_points(feelgpu_mesh_t m, const double* xyz, cudaStream_t s)
{
    feelgpu_context_t ctx = m->ctx;
    const size_t bytes = sizeof(double) * 3 * (size_t)m->nPts;
    cudaPointerAttributes at;
    const bool pinned = cudaPointerGetAttributes(&at, xyz) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();   // (older drivers report unregistered host memory as an error)
    const double* src = xyz;
    if (!pinned) {
        // the staging buffer is reused: the previous upload through it (on whatever stream) must have left it
        if (m->stage_busy) CTX_CUDA(ctx, cudaEventSynchronize(m->stage_event));
        std::memcpy(m->h_stage, xyz, bytes);
        src = m->h_stage;
    }
    // d_raw is shared as well: order this upload after the previous one's padding kernel when the stream differs
    if (m->stage_busy) CTX_CUDA(ctx, cudaStreamWaitEvent(s, m->stage_event, 0));
    CTX_CUDA(ctx, cudaMemcpyAsync(m->d_raw, src, bytes, cudaMemcpyHostToDevice, s));
    const int grid = (int)std::min<int64_t>((m->nPts + 255) / 256, (int64_t)std::max(ctx->sm_count, 1) * 8);
    k_pad_points<<<grid, 256, 0, s>>>(m->nPts, m->d_raw, m->d_xyz);
    ctx->launches++;
    CTX_CUDA(ctx, cudaGetLastError());
    CTX_CUDA(ctx, cudaEventRecord(m->stage_event, s));
    m->stage_busy = true;
    return FEELGPU_OK;
}

int feelgpu_mesh_create(feelgpu_context_t ctx, int64_t nPts, const double* xyz, int64_t nElt, const int32_t* conn,
                        feelgpu_mesh_t* out)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    if (!out || !xyz || !conn || nPts <= 0 || nElt < 0) ARG_FAIL(ctx, "mesh_create: bad argument");
    if (nPts >= (int64_t)1 << 31) ARG_FAIL(ctx, "mesh_create: more than 2^31 points");
    for (int64_t k = 0; k < 4 * nElt; ++k)
        if (conn[k] < 0 || conn[k] >= nPts) ARG_FAIL(ctx, "mesh_create: connectivity entry outside [0,nPts)");
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    feelgpu_mesh_t m = nullptr;
    return abi_guard(ctx, "mesh_create", [&]() -> int {
    m = new feelgpu_mesh_s;
    m->ctx = ctx;
    m->nPts = nPts;
    m->nElt = nElt;
    cudaError_t err;
    if ((err = cudaMalloc(&m->d_xyz, sizeof(double) * 4 * nPts)) != cudaSuccess ||
        (err = cudaMalloc(&m->d_conn, sizeof(int4) * std::max<int64_t>(nElt, 1))) != cudaSuccess ||
        (err = cudaMalloc(&m->d_raw, sizeof(double) * 3 * nPts)) != cudaSuccess ||
        (err = cudaMallocHost(&m->h_stage, sizeof(double) * 3 * nPts)) != cudaSuccess ||
        (err = cudaEventCreateWithFlags(&m->stage_event, cudaEventDisableTiming)) != cudaSuccess) {
        ctx->last_error = std::string("mesh_create: ") + cudaGetErrorString(err);
        feelgpu_mesh_destroy(m);
        return FEELGPU_ERR_CUDA;
    }
    m->h_xyz.assign(xyz, xyz + 3 * nPts);
    m->h_conn.assign(conn, conn + 4 * nElt);
    int rc = upload_points(m, xyz, ctx->stream);
    if (rc) { feelgpu_mesh_destroy(m); return rc; }
    err = cudaMemcpyAsync(m->d_conn, conn, sizeof(int32_t) * 4 * nElt, cudaMemcpyHostToDevice, ctx->stream);
    if (err == cudaSuccess) err = cudaStreamSynchronize(ctx->stream);
    if (err != cudaSuccess) {
        ctx->last_error = std::string("mesh_create: ") + cudaGetErrorString(err);
        feelgpu_mesh_destroy(m);
        return FEELGPU_ERR_CUDA;
    }
    *out = m;
    return FEELGPU_OK;
    }, [&] { if (m) feelgpu_mesh_destroy(m); });
}

int feelgpu_mesh_update_points(feelgpu_mesh_t m, const double* xyz, void* stream)
{
    if (!m) return FEELGPU_ERR_ARG;
    if (!xyz) ARG_FAIL(m->ctx, "mesh_update_points: null xyz");
    cudaStream_t s = stream ? (cudaStream_t)stream : m->ctx->stream;
    CTX_CUDA(m->ctx, cudaSetDevice(m->ctx->device));
    return upload_points(m, xyz, s);
}

int feelgpu_mesh_destroy(feelgpu_mesh_t m)
{
    if (!m) return FEELGPU_ERR_ARG;
    if (m->d_xyz) cudaFree(m->d_xyz);
    if (m->d_raw) cudaFree(m->d_raw);
    if (m->d_conn) cudaFree(m->d_conn);
    if (m->h_stage) cudaFreeHost(m->h_stage);
    if (m->stage_event) cudaEventDestroy(m->stage_event);
    delete m;
    return FEELGPU_OK;
}

// ---- space -----------------------------------------------------------------------------------------------------
int feelgpu_space_create(feelgpu_context_t ctx, feelgpu_mesh_t mesh, int order, int ncomp, const int32_t* l2g,
                         int64_t nWithGhost, int64_t nOwned, const int64_t* p2c, feelgpu_space_t* out)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    if (!mesh || !l2g || !out) ARG_FAIL(ctx, "space_create: bad argument");
    if (order < 1 || order > 3) { ctx->last_error = "space_create: Lagrange order must be 1, 2 or 3"; return FEELGPU_ERR_UNSUPPORTED; }
    if (ncomp != 1 && ncomp != 3) { ctx->last_error = "space_create: ncomp must be 1 or 3"; return FEELGPU_ERR_UNSUPPORTED; }
    if (nOwned < 0 || nOwned > nWithGhost || nWithGhost >= (int64_t)1 << 31) ARG_FAIL(ctx, "space_create: bad dof counts");
    const int n = (order + 1) * (order + 2) * (order + 3) / 6, nd = n * ncomp;
    for (int64_t e = 0; e < mesh->nElt; ++e)
        for (int i = 0; i < n; ++i) {
            int32_t g0 = l2g[e * nd + i];
            if (g0 < 0 || g0 + ncomp - 1 >= nWithGhost) ARG_FAIL(ctx, "space_create: dof id outside [0,nDofLocalWithGhost)");
            // product spaces: local = n*c + i, global = base + c (feeldiscr/doftable.hpp:1289-1299)
            for (int c = 1; c < ncomp; ++c)
                if (l2g[e * nd + n * c + i] != g0 + c) {
                    ctx->last_error = "space_create: vector dofs must be interleaved (global = ncomp*node + c)";
                    return FEELGPU_ERR_UNSUPPORTED;
                }
        }
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    feelgpu_space_t sp = nullptr;
    return abi_guard(ctx, "space_create", [&]() -> int {
    sp = new feelgpu_space_s;
    sp->ctx = ctx;
    sp->mesh = mesh;
    sp->order = order;
    sp->ncomp = ncomp;
    sp->n = n;
    sp->nd = nd;
    sp->nDofWithGhost = nWithGhost;
    sp->nDofOwned = nOwned;
    sp->h_l2g.assign(l2g, l2g + mesh->nElt * nd);
    if (p2c) sp->h_p2c.assign(p2c, p2c + nWithGhost);
    cudaError_t err = cudaMalloc(&sp->d_l2g, sizeof(int32_t) * std::max<int64_t>(mesh->nElt * nd, 1));
    if (err == cudaSuccess) err = cudaMemcpy(sp->d_l2g, l2g, sizeof(int32_t) * mesh->nElt * nd, cudaMemcpyHostToDevice);
    if (err == cudaSuccess && p2c) {
        err = cudaMalloc(&sp->d_p2c, sizeof(int64_t) * std::max<int64_t>(nWithGhost, 1));
        if (err == cudaSuccess) err = cudaMemcpy(sp->d_p2c, p2c, sizeof(int64_t) * nWithGhost, cudaMemcpyHostToDevice);
    }
    if (err != cudaSuccess) {
        ctx->last_error = std::string("space_create: ") + cudaGetErrorString(err);
        feelgpu_space_destroy(sp);
        return FEELGPU_ERR_CUDA;
    }
    *out = sp;
    return FEELGPU_OK;
    }, [&] { if (sp) feelgpu_space_destroy(sp); });
}

int feelgpu_space_destroy(feelgpu_space_t sp)
{
    if (!sp) return FEELGPU_ERR_ARG;
    if (sp->d_l2g) cudaFree(sp->d_l2g);
    if (sp->d_p2c) cudaFree(sp->d_p2c);
    delete sp;
    return FEELGPU_OK;
}

// ---- pattern ---------------------------------------------------------------------------------------------------
// ja: host columns, or null when the columns are already on the device (d_ja_graph, owned by a feelgpu_graph_t)
static int pattern_create_impl(feelgpu_context_t ctx, feelgpu_space_t test, feelgpu_space_t trial, int64_t nRows,
                               const int64_t* ia, const int64_t* ja, const int64_t* d_ja_graph, int flags, feelgpu_pattern_t* out)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    if (!test || !trial || !ia || (!ja && !d_ja_graph) || !out || nRows < 0) ARG_FAIL(ctx, "pattern_create: bad argument");
    if (test->mesh != trial->mesh) ARG_FAIL(ctx, "pattern_create: test and trial spaces live on different meshes");
    if (test->n != trial->n || test->ncomp != trial->ncomp) {
        ctx->last_error = "pattern_create: test and trial spaces must have the same order and components";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    if (nRows > test->nDofWithGhost) ARG_FAIL(ctx, "pattern_create: more rows than local test dofs");
    if (ia[0] != 0) ARG_FAIL(ctx, "pattern_create: ia[0] != 0");
    for (int64_t r = 0; r < nRows; ++r)
        if (ia[r + 1] < ia[r]) ARG_FAIL(ctx, "pattern_create: ia not monotone");
    const bool async = (flags & FEELGPU_PATTERN_ASYNC) != 0;
    flags &= ~FEELGPU_PATTERN_ASYNC;
    if (flags == 0) flags = FEELGPU_PATTERN_SLOTMAP | FEELGPU_PATTERN_RPLAN | (test->order <= 2 ? FEELGPU_PATTERN_EPLAN : 0);
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    feelgpu_pattern_t pat = nullptr;
    int64_t* d_ja = nullptr;
    return abi_guard(ctx, "pattern_create", [&]() -> int {
    pat = new feelgpu_pattern_s;
    pat->ctx = ctx;
    pat->test = test;
    pat->trial = trial;
    pat->nRows = nRows;
    pat->nnz = ia[nRows];
    pat->h_ia_owned_end = ia[std::min<int64_t>(test->nDofOwned, nRows)];
    pat->h_ia.assign(ia, ia + nRows + 1);
    if (test->nDofOwned < nRows) {
        if (ja) pat->h_ja_tail.assign(ja + pat->h_ia_owned_end, ja + ia[nRows]);
        else {
            pat->h_ja_tail.resize((size_t)(ia[nRows] - pat->h_ia_owned_end));
            if (cudaMemcpy(pat->h_ja_tail.data(), d_ja_graph + pat->h_ia_owned_end, sizeof(int64_t) * pat->h_ja_tail.size(),
                           cudaMemcpyDeviceToHost) != cudaSuccess) {
                ctx->last_error = "pattern_create: copy of the ghost rows' columns failed";
                feelgpu_pattern_destroy(pat);
                pat = nullptr;
                return FEELGPU_ERR_CUDA;
            }
        }
    }
    const int n = test->n;
    const int64_t nElt = test->mesh->nElt;
    cudaError_t err = cudaMalloc(&pat->d_ia, sizeof(int64_t) * (nRows + 1));
    if (err == cudaSuccess) err = cudaMemcpy(pat->d_ia, ia, sizeof(int64_t) * (nRows + 1), cudaMemcpyHostToDevice);
    if (!d_ja_graph) {
        if (err == cudaSuccess) err = cudaMalloc(&d_ja, sizeof(int64_t) * std::max<int64_t>(pat->nnz, 1));
        if (err == cudaSuccess) err = cudaMemcpy(d_ja, ja, sizeof(int64_t) * pat->nnz, cudaMemcpyHostToDevice);
    }
    const int64_t* d_cols = d_ja_graph ? d_ja_graph : d_ja;
    if (err == cudaSuccess) err = cudaMalloc(&pat->d_pos, sizeof(uint16_t) * std::max<int64_t>(nElt * n * n, 1));
    if (err != cudaSuccess) {
        ctx->last_error = std::string("pattern_create: ") + cudaGetErrorString(err);
        if (d_ja) cudaFree(d_ja);
        feelgpu_pattern_destroy(pat);
        return FEELGPU_ERR_CUDA;
    }
    int rc = feelgpu_launch_build_pos(pat, d_cols, ctx->stream);
    // 32-bit copy of the columns for feelgpu_form2_zero_rows / matvec / the exchange, narrowed on the device
    if (rc == FEELGPU_OK) rc = feelgpu_launch_narrow_columns(pat, d_cols, ctx->stream);
    if (d_ja) cudaFree(d_ja); // the 64-bit column array is only needed to build the maps
    d_ja = nullptr;
    if (rc != FEELGPU_OK) {
        feelgpu_pattern_destroy(pat);
        pat = nullptr;
        return rc;
    }
    // (the background build of the entity plan starts first: it overlaps the row-thread plan's build)
    if (rc == FEELGPU_OK && (flags & FEELGPU_PATTERN_EPLAN)) rc = feelgpu_eplan_build(pat, ia, async && (flags & FEELGPU_PATTERN_RPLAN));
    if (rc == FEELGPU_OK && (flags & FEELGPU_PATTERN_RPLAN)) rc = feelgpu_rplan_build(pat, ia);
    // the row plan is also the fallback of a mesh the entity plan could not take
    if (rc == FEELGPU_OK && ((flags & FEELGPU_PATTERN_ROWPLAN) ||
                             ((flags & FEELGPU_PATTERN_EPLAN) && !pat->eplan && !pat->eplan_async && !pat->rplan)))
    {
        std::vector<int64_t> hja;                      // (the row plan's builder reads the columns on the host)
        if (!ja) {
            hja.resize((size_t)std::max<int64_t>(pat->nnz, 1));
            if (cudaMemcpy(hja.data(), d_ja_graph, sizeof(int64_t) * pat->nnz, cudaMemcpyDeviceToHost) != cudaSuccess) {
                ctx->last_error = "pattern_create: copy of the columns to the host failed";
                rc = FEELGPU_ERR_CUDA;
            }
        }
        if (rc == FEELGPU_OK) rc = feelgpu_rowplan_build(pat, ia, ja ? ja : hja.data());
    }
    if (rc != FEELGPU_OK) {
        feelgpu_pattern_destroy(pat);
        pat = nullptr;
        return rc;
    }
    *out = pat;
    return FEELGPU_OK;
    }, [&] {
        if (d_ja) cudaFree(d_ja);
        if (pat) feelgpu_pattern_destroy(pat);
    });
}

int feelgpu_pattern_create(feelgpu_context_t ctx, feelgpu_space_t test, feelgpu_space_t trial, int64_t nRows,
                           const int64_t* ia, const int64_t* ja, int flags, feelgpu_pattern_t* out)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    if (!ja) ARG_FAIL(ctx, "pattern_create: bad argument");
    return pattern_create_impl(ctx, test, trial, nRows, ia, ja, nullptr, flags, out);
}

// The pattern of a graph that was built on the device (feelgpu_graph_create): the columns stay where they are, only the row
// pointers come to the host (the plan builders read them).  The graph may be destroyed afterwards.
int feelgpu_pattern_create_from_graph(feelgpu_graph_t graph, int flags, feelgpu_pattern_t* out)
{
    if (!graph || !graph->space) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = graph->space->ctx;
    if (!out) ARG_FAIL(ctx, "pattern_create_from_graph: bad argument");
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    std::vector<int64_t> ia;
    int rc = abi_guard(ctx, "pattern_create_from_graph", [&]() -> int {
        ia.resize((size_t)graph->nRows + 1);
        CTX_CUDA(ctx, cudaMemcpy(ia.data(), graph->d_ia, sizeof(int64_t) * ia.size(), cudaMemcpyDeviceToHost));
        return FEELGPU_OK;
    }, [] {});
    if (rc) return rc;
    return pattern_create_impl(ctx, graph->space, graph->space, graph->nRows, ia.data(), nullptr, graph->d_ja, flags, out);
}

int feelgpu_pattern_nnz(feelgpu_pattern_t pat, int64_t* nnz)
{
    if (!pat || !nnz) return FEELGPU_ERR_ARG;
    *nnz = pat->nnz;
    return FEELGPU_OK;
}

int feelgpu_pattern_plan_stats(feelgpu_pattern_t pat, int64_t stats[10])
{
    if (!pat || !stats) return FEELGPU_ERR_ARG;
    for (int k = 0; k < 10; ++k) stats[k] = 0;
    const RowPlan* p = pat->plan;
    if (!p) return FEELGPU_OK;
    stats[0] = 1;
    stats[1] = p->ncta;
    stats[2] = p->nbundle;
    stats[3] = p->nunits;
    stats[4] = p->ncontrib;
    stats[5] = p->nsrc_padded;
    stats[6] = p->nvisits;
    stats[7] = p->bytes;
    stats[8] = p->maxchunks;
    stats[9] = p->maxrounds;
    return FEELGPU_OK;
}

int feelgpu_pattern_eplan_stats(feelgpu_pattern_t pat, int64_t stats[10])
{
    if (!pat || !stats) return FEELGPU_ERR_ARG;
    int rc = feelgpu_eplan_poll(pat, false);
    if (rc) return rc;
    return feelgpu_eplan_stats(pat, stats);
}

int feelgpu_pattern_plans_ready(feelgpu_pattern_t pat, int wait, int* ready, double* build_seconds)
{
    if (!pat) return FEELGPU_ERR_ARG;
    int rc = feelgpu_eplan_poll(pat, wait != 0);
    if (ready) *ready = pat->eplan_async ? 0 : 1;
    if (build_seconds) *build_seconds = pat->eplan_build_seconds;
    return rc;
}

int feelgpu_pattern_rplan_stats(feelgpu_pattern_t pat, int64_t stats[10])
{
    if (!pat || !stats) return FEELGPU_ERR_ARG;
    return feelgpu_rplan_stats(pat, stats);
}

int feelgpu_pattern_destroy(feelgpu_pattern_t pat)
{
    if (!pat) return FEELGPU_ERR_ARG;
    feelgpu_exchange_detach(pat);
    feelgpu_rowplan_destroy(pat);
    feelgpu_eplan_destroy(pat);
    feelgpu_rplan_destroy(pat);
    if (pat->d_ia) cudaFree(pat->d_ia);
    if (pat->d_pos) cudaFree(pat->d_pos);
    if (pat->d_stage) cudaFree(pat->d_stage);
    if (pat->d_ja32) cudaFree(pat->d_ja32);
    delete pat;
    return FEELGPU_OK;
}

extern "C++" int feelgpu_builder_threads(feelgpu_context_t ctx)
{
    if (const char* e = std::getenv("FEELGPU_BUILD_THREADS"))
        if (std::atoi(e) > 0) return std::atoi(e);
    const int hw = (int)std::max(1u, std::thread::hardware_concurrency());
    int share = 1;
    if (ctx && ctx->nranks > 1) {
        int ndev = 1;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) ndev = 1;
        share = std::min(ctx->nranks, ndev);
    }
    // (at least 8 where the machine has them: mild oversubscription keeps the cores busy when the ranks finish at different times)
    return std::max(hw / share, std::min(hw, 8));
}

// ---- work calls ------------------------------------------------------------------------------------------------
static int fill_params(feelgpu_context_t ctx, int kind, const feelgpu_coeff& c, const feelgpu_tables& t,
                       feelgpu_space_t sp, bool form2, FormParams* hp)
{
    std::memset(hp, 0, sizeof(FormParams));
    if (!t.w || !t.xi || !t.phi || t.nq < 1 || t.nq > FEELGPU_MAX_NQ) ARG_FAIL(ctx, "assemble: bad tables (nq must be in [1,64])");
    if (t.n != sp->n) ARG_FAIL(ctx, "assemble: tables.n does not match the space");
    if (form2 && kind != FEELGPU_FORM2_MASS && !t.dphi) ARG_FAIL(ctx, "assemble: tables.dphi is required");
    hp->kind = kind;
    hp->coeff_kind = c.kind;
    hp->n = t.n;
    hp->nq = t.nq;
    hp->ncomp = sp->ncomp;
    hp->cval[0] = c.value[0];
    hp->cval[1] = c.value[1];
    hp->cval[2] = c.value[2];
    hp->lambda = c.lambda;
    hp->mu = c.mu;
    std::memcpy(hp->w, t.w, sizeof(double) * t.nq);
    std::memcpy(hp->xi, t.xi, sizeof(double) * 3 * t.nq);
    if (form2 && kind == FEELGPU_FORM2_ELASTICITY) {
        if (sp->ncomp != 3) { ctx->last_error = "elasticity needs a 3-component space"; return FEELGPU_ERR_UNSUPPORTED; }
        hp->coeff_kind = FEELGPU_COEFF_CONST;
        return FEELGPU_OK;
    }
    switch (c.kind) {
    case FEELGPU_COEFF_CONST: break;
    case FEELGPU_COEFF_FIELD:
        if (!c.field_space || !c.field_values || c.field_space->ncomp != 1 || c.field_space->mesh != sp->mesh)
            ARG_FAIL(ctx, "assemble: FIELD coefficient needs a scalar space on the same mesh and device values");
        if (t.nf != c.field_space->n || !t.phif) ARG_FAIL(ctx, "assemble: tables.phif/nf do not match the field space");
        hp->nf = t.nf;
        break;
    case FEELGPU_COEFF_EXPR:
        if (c.expr_len < 1 || c.expr_len > FEELGPU_EXPR_MAX_LEN || !c.expr_ops || !c.expr_consts)
            ARG_FAIL(ctx, "assemble: bad EXPR program");
        {
            int sp_ = 0, maxsp = 0; // validate the postfix program's stack discipline
            for (int k = 0; k < c.expr_len; ++k) {
                int op = c.expr_ops[k];
                if (op >= FEELGPU_OP_X && op <= FEELGPU_OP_CONST) sp_++;
                else if ((op >= FEELGPU_OP_ADD && op <= FEELGPU_OP_DIV) || op == FEELGPU_OP_POW) sp_--;
                else if (op < 0 || op > FEELGPU_OP_POW) ARG_FAIL(ctx, "assemble: unknown EXPR opcode");
                if (sp_ < 1) ARG_FAIL(ctx, "assemble: EXPR stack underflow");
                maxsp = std::max(maxsp, sp_);
            }
            if (sp_ != 1 || maxsp > 16) ARG_FAIL(ctx, "assemble: EXPR must leave one value and use <= 16 stack slots");
        }
        hp->expr_len = c.expr_len;
        std::memcpy(hp->expr_ops, c.expr_ops, sizeof(int32_t) * c.expr_len);
        std::memcpy(hp->expr_consts, c.expr_consts, sizeof(double) * c.expr_len);
        break;
    default: ARG_FAIL(ctx, "assemble: unknown coefficient kind");
    }
    if (sp->ncomp > 1 && c.kind != FEELGPU_COEFF_CONST) {
        ctx->last_error = "assemble: vector spaces support constant coefficients only";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    return FEELGPU_OK;
}

int feelgpu_form2_assemble(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc, int accumulate, double* a,
                           void* stream)
{
    if (!pat) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = pat->ctx;
    if (!desc || !a) ARG_FAIL(ctx, "form2_assemble: null argument");
    if (desc->kind < FEELGPU_FORM2_MASS || desc->kind > FEELGPU_FORM2_ELASTICITY) {
        ctx->last_error = "form2_assemble: unknown form kind";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    FormParams hp;
    int rc = fill_params(ctx, desc->kind, desc->coeff, desc->tables, pat->test, true, &hp);
    if (rc) return rc;
    // a plan built in the background is installed as soon as it is ready; asking for it by name waits for it
    if (pat->eplan_async && (rc = feelgpu_eplan_poll(pat, desc->algorithm == FEELGPU_ALGO_EPLAN))) return rc;
    int algo = desc->algorithm;
    if (algo == FEELGPU_ALGO_AUTO)
        algo = feelgpu_eplan_supports(pat, desc)     ? FEELGPU_ALGO_EPLAN
               : feelgpu_rplan_supports(pat, desc)   ? FEELGPU_ALGO_RPLAN
               : feelgpu_rowplan_supports(pat, desc) ? FEELGPU_ALGO_ROWPLAN
                                                     : FEELGPU_ALGO_SCATTER;
    if (algo < FEELGPU_ALGO_SCATTER || algo > FEELGPU_ALGO_RPLAN) ARG_FAIL(ctx, "form2_assemble: unknown algorithm");
    if ((algo == FEELGPU_ALGO_ROWPLAN && !feelgpu_rowplan_supports(pat, desc)) ||
        (algo == FEELGPU_ALGO_RPLAN && !feelgpu_rplan_supports(pat, desc)) ||
        (algo == FEELGPU_ALGO_EPLAN && !feelgpu_eplan_supports(pat, desc))) {
        ctx->last_error = "form2_assemble: the owner-computes plan does not cover this form/pattern";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    feelgpu_exchange_note_assembly(pat, algo == FEELGPU_ALGO_EPLAN, accumulate, a, s);   // multi-GPU: where the ghost rows go
    if (ctx->timing) cudaEventRecord(ctx->ev[0], s);
    if (algo == FEELGPU_ALGO_EPLAN) {
        if (ctx->timing) cudaEventRecord(ctx->ev[1], s);
        rc = feelgpu_launch_form2_eplan(pat, hp, desc, accumulate, a, s);
    } else if (algo == FEELGPU_ALGO_RPLAN) {
        if (ctx->timing) cudaEventRecord(ctx->ev[1], s);
        rc = feelgpu_launch_form2_rplan(pat, hp, desc, accumulate, a, s);
    } else if (algo == FEELGPU_ALGO_ROWPLAN) {
        if (ctx->timing) cudaEventRecord(ctx->ev[1], s);
        rc = feelgpu_launch_form2_rowplan(pat, hp, desc, accumulate, a, s);
    } else {
        if (!accumulate) CTX_CUDA(ctx, cudaMemsetAsync(a, 0, sizeof(double) * pat->nnz, s));
        if (ctx->timing) cudaEventRecord(ctx->ev[1], s);
        rc = feelgpu_launch_form2_scatter(pat, hp, desc, a, s);
    }
    if (ctx->timing) cudaEventRecord(ctx->ev[2], s);
    return rc;
}

// Allocation-free variant (include/feelgpu.h): no cudaMalloc / cudaFree / allocating library call between entry and return;
// the table cache, the background plan installer and the exchange bookkeeping of feelgpu_form2_assemble are not touched.
int feelgpu_form2_assemble_noalloc(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc, int accumulate, double* a,
                                   void* stream)
{
    if (!pat) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = pat->ctx;
    if (!desc || !a) ARG_FAIL(ctx, "form2_assemble_noalloc: null argument");
    if (pat->exchange) {
        ctx->last_error = "form2_assemble_noalloc: patterns with an exchange attached are not supported";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    FormParams hp;
    int rc = fill_params(ctx, desc->kind, desc->coeff, desc->tables, pat->test, true, &hp);
    if (rc) return rc;
    if (!feelgpu_rplan_noalloc_supports(pat, desc)) {
        ctx->last_error = "form2_assemble_noalloc: needs a MASS / STIFFNESS form with a FIELD or EXPR coefficient on a scalar "
                          "space whose pattern carries the row-thread plan";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    return feelgpu_launch_form2_rplan_noalloc(pat, hp, desc, accumulate, a, s);
}

int feelgpu_form1_assemble(feelgpu_space_t sp, const feelgpu_form1_desc* desc, int accumulate, double* b, void* stream)
{
    if (!sp) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = sp->ctx;
    if (!desc || !b) ARG_FAIL(ctx, "form1_assemble: null argument");
    if (desc->kind != FEELGPU_FORM1_SOURCE) {
        ctx->last_error = "form1_assemble: unknown form kind";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    FormParams hp;
    int rc = fill_params(ctx, desc->kind, desc->coeff, desc->tables, sp, false, &hp);
    if (rc) return rc;
    if (!accumulate) CTX_CUDA(ctx, cudaMemsetAsync(b, 0, sizeof(double) * sp->nDofWithGhost, s));
    return feelgpu_launch_form1(sp, hp, desc, b, s);
}

int feelgpu_form2_assemble_host(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc, int accumulate, double* a_host)
{
    if (!pat) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = pat->ctx;
    if (!a_host) ARG_FAIL(ctx, "form2_assemble_host: null output");
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    // the device staging array lives with the pattern: allocating 8*nnz bytes per call costs more than the assembly
    if (!pat->d_stage) CTX_CUDA(ctx, cudaMalloc(&pat->d_stage, sizeof(double) * std::max<int64_t>(pat->nnz, 1)));
    double* d = pat->d_stage;
    int rc = FEELGPU_OK;
    cudaError_t err = cudaSuccess;
    if (accumulate) err = cudaMemcpyAsync(d, a_host, sizeof(double) * pat->nnz, cudaMemcpyHostToDevice, ctx->stream);
    if (err == cudaSuccess) rc = feelgpu_form2_assemble(pat, desc, accumulate, d, ctx->stream);
    if (err == cudaSuccess && rc == FEELGPU_OK)
        err = cudaMemcpyAsync(a_host, d, sizeof(double) * pat->nnz, cudaMemcpyDeviceToHost, ctx->stream);
    if (err == cudaSuccess) err = cudaStreamSynchronize(ctx->stream);
    if (err != cudaSuccess) {
        ctx->last_error = std::string("form2_assemble_host: ") + cudaGetErrorString(err);
        return FEELGPU_ERR_CUDA;
    }
    return rc;
}

int feelgpu_form1_assemble_host(feelgpu_space_t sp, const feelgpu_form1_desc* desc, int accumulate, double* b_host)
{
    if (!sp) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = sp->ctx;
    if (!b_host) ARG_FAIL(ctx, "form1_assemble_host: null output");
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    double* d = nullptr;
    CTX_CUDA(ctx, cudaMalloc(&d, sizeof(double) * std::max<int64_t>(sp->nDofWithGhost, 1)));
    int rc = FEELGPU_OK;
    cudaError_t err = cudaSuccess;
    if (accumulate) err = cudaMemcpyAsync(d, b_host, sizeof(double) * sp->nDofWithGhost, cudaMemcpyHostToDevice, ctx->stream);
    if (err == cudaSuccess) rc = feelgpu_form1_assemble(sp, desc, accumulate, d, ctx->stream);
    if (err == cudaSuccess && rc == FEELGPU_OK)
        err = cudaMemcpyAsync(b_host, d, sizeof(double) * sp->nDofWithGhost, cudaMemcpyDeviceToHost, ctx->stream);
    if (err == cudaSuccess) err = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    if (err != cudaSuccess) {
        ctx->last_error = std::string("form1_assemble_host: ") + cudaGetErrorString(err);
        return FEELGPU_ERR_CUDA;
    }
    return rc;
}

int feelgpu_form2_zero_rows(feelgpu_pattern_t pat, int64_t n, const int32_t* rows, const double* values, int strategy,
                            double value_on_diagonal, double* a_values, double* b, void* stream)
{
    if (!pat) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = pat->ctx;
    if (n < 0 || n > pat->nRows || (n && (!rows || !values)) || !a_values) ARG_FAIL(ctx, "zero_rows: bad argument");
    if (!(strategy & FEELGPU_ON_ELIMINATION) || (strategy & ~(FEELGPU_ON_ELIMINATION | FEELGPU_ON_KEEP_DIAGONAL | FEELGPU_ON_SYMMETRIC))) {
        ctx->last_error = "zero_rows: the strategy must contain ContextOn::ELIMINATION (penalisation stays on the host)";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    return feelgpu_launch_zero_rows(pat, (int)n, rows, values, strategy, value_on_diagonal, a_values, b, s);
}

// ---- memory helpers -----------------------------------------------------------------------------------------
int feelgpu_matvec(feelgpu_pattern_t pat, const double* a_values, const double* x, double* y, void* stream)
{
    if (!pat) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = pat->ctx;
    if (!a_values || !x || !y) ARG_FAIL(ctx, "matvec: null argument");
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    return feelgpu_launch_matvec(pat, a_values, x, y, stream ? (cudaStream_t)stream : ctx->stream);
}

int feelgpu_faces_create(feelgpu_mesh_t mesh, int64_t nFaces, const int32_t* element, const int32_t* local_face,
                         feelgpu_faces_t* out)
{
    if (!mesh) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = mesh->ctx;
    if (!out || nFaces < 0 || (nFaces > 0 && (!element || !local_face))) ARG_FAIL(ctx, "faces_create: bad argument");
    for (int64_t k = 0; k < nFaces; ++k)
        if (element[k] < 0 || element[k] >= mesh->nElt || local_face[k] < 0 || local_face[k] > 3)
            ARG_FAIL(ctx, "faces_create: (element, local face) out of range");
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    feelgpu_faces_t fc = new (std::nothrow) feelgpu_faces_s;
    if (!fc) { ctx->last_error = "faces_create: out of host memory"; return FEELGPU_ERR_HOST; }
    fc->mesh = mesh;
    fc->n = nFaces;
    const size_t bytes = sizeof(int32_t) * (size_t)std::max<int64_t>(nFaces, 1);
    cudaError_t err = cudaMalloc(&fc->d_elt, bytes);
    if (err == cudaSuccess) err = cudaMalloc(&fc->d_lface, bytes);
    if (err == cudaSuccess && nFaces > 0) err = cudaMemcpy(fc->d_elt, element, sizeof(int32_t) * nFaces, cudaMemcpyHostToDevice);
    if (err == cudaSuccess && nFaces > 0) err = cudaMemcpy(fc->d_lface, local_face, sizeof(int32_t) * nFaces, cudaMemcpyHostToDevice);
    if (err != cudaSuccess) {
        ctx->last_error = std::string("faces_create: ") + cudaGetErrorString(err);
        feelgpu_faces_destroy(fc);
        return FEELGPU_ERR_CUDA;
    }
    *out = fc;
    return FEELGPU_OK;
}

int feelgpu_faces_destroy(feelgpu_faces_t fc)
{
    if (!fc) return FEELGPU_ERR_ARG;
    cudaFree(fc->d_elt);
    cudaFree(fc->d_lface);
    delete fc;
    return FEELGPU_OK;
}

static int check_face_tables(feelgpu_context_t ctx, const feelgpu_coeff& c, const feelgpu_tables& t)
{
    if (t.nq % 4 != 0) ARG_FAIL(ctx, "assemble_faces: tables.nq must be 4 x (points per face)");
    if (c.kind == FEELGPU_COEFF_FIELD) {
        ctx->last_error = "assemble_faces: FIELD coefficients are not covered on faces";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    return FEELGPU_OK;
}

int feelgpu_form1_assemble_faces(feelgpu_space_t sp, feelgpu_faces_t fc, const feelgpu_form1_desc* desc, int accumulate,
                                 double* b, void* stream)
{
    if (!sp) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = sp->ctx;
    if (!fc || !desc || !b) ARG_FAIL(ctx, "form1_assemble_faces: null argument");
    if (fc->mesh != sp->mesh) ARG_FAIL(ctx, "form1_assemble_faces: the faces belong to another mesh");
    if (desc->kind != FEELGPU_FORM1_SOURCE) {
        ctx->last_error = "form1_assemble_faces: unknown form kind";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    FormParams hp;
    int rc = fill_params(ctx, desc->kind, desc->coeff, desc->tables, sp, false, &hp);
    if (rc) return rc;
    if ((rc = check_face_tables(ctx, desc->coeff, desc->tables))) return rc;
    if (!accumulate) CTX_CUDA(ctx, cudaMemsetAsync(b, 0, sizeof(double) * sp->nDofWithGhost, s));
    return fc->n ? feelgpu_launch_form1_faces(sp, fc, hp, desc, b, s) : FEELGPU_OK;
}

int feelgpu_form2_assemble_faces(feelgpu_pattern_t pat, feelgpu_faces_t fc, const feelgpu_form2_desc* desc, int accumulate,
                                 double* a, void* stream)
{
    if (!pat) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = pat->ctx;
    if (!fc || !desc || !a) ARG_FAIL(ctx, "form2_assemble_faces: null argument");
    if (fc->mesh != pat->test->mesh) ARG_FAIL(ctx, "form2_assemble_faces: the faces belong to another mesh");
    if (desc->kind != FEELGPU_FORM2_MASS) {
        ctx->last_error = "form2_assemble_faces: only k*idt(u)*id(v) is covered on faces";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    if (!pat->d_pos) {
        ctx->last_error = "form2_assemble_faces: the pattern was created without the element->slot map";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    FormParams hp;
    int rc = fill_params(ctx, desc->kind, desc->coeff, desc->tables, pat->test, true, &hp);
    if (rc) return rc;
    if ((rc = check_face_tables(ctx, desc->coeff, desc->tables))) return rc;
    if (!accumulate) CTX_CUDA(ctx, cudaMemsetAsync(a, 0, sizeof(double) * pat->nnz, s));
    return fc->n ? feelgpu_launch_form2_faces(pat, fc, hp, desc, a, s) : FEELGPU_OK;
}

int feelgpu_malloc(feelgpu_context_t ctx, int64_t bytes, void** dptr)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    if (!dptr || bytes < 0) ARG_FAIL(ctx, "malloc: bad argument");
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    CTX_CUDA(ctx, cudaMalloc(dptr, (size_t)std::max<int64_t>(bytes, 1)));
    return FEELGPU_OK;
}
int feelgpu_free(feelgpu_context_t ctx, void* dptr)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    CTX_CUDA(ctx, cudaFree(dptr));
    return FEELGPU_OK;
}
int feelgpu_memcpy_h2d(feelgpu_context_t ctx, void* dst, const void* src, int64_t bytes)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    CTX_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyHostToDevice, ctx->stream));
    CTX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return FEELGPU_OK;
}
int feelgpu_memset_zero(feelgpu_context_t ctx, void* dptr, int64_t bytes, void* stream)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    if (!dptr || bytes < 0) ARG_FAIL(ctx, "memset_zero: null pointer or negative size");
    CTX_CUDA(ctx, cudaSetDevice(ctx->device));
    CTX_CUDA(ctx, cudaMemsetAsync(dptr, 0, (size_t)bytes, stream ? (cudaStream_t)stream : ctx->stream));
    return FEELGPU_OK;
}
int feelgpu_memcpy_d2h(feelgpu_context_t ctx, void* dst, const void* src, int64_t bytes)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    CTX_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CTX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return FEELGPU_OK;
}
int feelgpu_host_alloc_pinned(feelgpu_context_t ctx, int64_t bytes, void** hptr)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    if (!hptr || bytes < 0) ARG_FAIL(ctx, "host_alloc_pinned: bad argument");
    CTX_CUDA(ctx, cudaMallocHost(hptr, (size_t)std::max<int64_t>(bytes, 1)));
    return FEELGPU_OK;
}
int feelgpu_host_free_pinned(feelgpu_context_t ctx, void* hptr)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    CTX_CUDA(ctx, cudaFreeHost(hptr));
    return FEELGPU_OK;
}

// ---- introspection -------------------------------------------------------------------------------------------
int feelgpu_launch_count(feelgpu_context_t ctx, int64_t* count)
{
    if (!ctx || !count) return FEELGPU_ERR_ARG;
    *count = ctx->launches;
    return FEELGPU_OK;
}
int feelgpu_set_timing(feelgpu_context_t ctx, int enable)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    ctx->timing = enable != 0;
    return FEELGPU_OK;
}
int feelgpu_last_timing(feelgpu_context_t ctx, double ms[3])
{
    if (!ctx || !ms) return FEELGPU_ERR_ARG;
    if (!ctx->timing) ARG_FAIL(ctx, "last_timing: timing is disabled (feelgpu_set_timing)");
    CTX_CUDA(ctx, cudaEventSynchronize(ctx->ev[2]));
    float t01 = 0, t12 = 0;
    CTX_CUDA(ctx, cudaEventElapsedTime(&t01, ctx->ev[0], ctx->ev[1]));
    CTX_CUDA(ctx, cudaEventElapsedTime(&t12, ctx->ev[1], ctx->ev[2]));
    ms[0] = t01;
    ms[1] = t12;
    ms[2] = 0;
    return FEELGPU_OK;
}

} // extern "C"
