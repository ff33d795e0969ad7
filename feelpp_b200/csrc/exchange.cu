// exchange.cu -- the multi-GPU side of the C ABI: communicator, ghost-row exchange plan, close().
//
// Replaces, for one node of GPUs (one process / MPI rank per GPU):
//   * GraphCSR::close, the graph exchange that tells every owner which (row, columns) its neighbours hold as ghost rows
//     (feelalg/graphcsr.cpp:800-951)                                              -> feelgpu_exchange_create (once)
//   * MatrixPetsc::close -> MatAssemblyBegin/End, the stash of off-process rows shipped to their owners and added
//     (feelalg/matrixpetsc.cpp:726-760, MatSetValues on ghost rows :2875-2880)     -> feelgpu_form2_close
//   * VectorPetscMPI::close -> VecAssemblyBegin/End + VecGhostUpdate (feelalg/vectorpetsc.cpp:1683-1693,1667-1676)
//                                                                                 -> feelgpu_form1_close
// Ghost rows come in the reference's order: local numbering = active dofs first, ghosts after, each group in first-touch
// order (feeldiscr/doftable.hpp:1975-1992), so the ghost rows of one owner are not contiguous.  Everything here works
// from a per-ghost-row owner table (DataMap::procOnGlobalCluster) and the process->cluster map.
//
// Two transports for the matrix values:
//   peer memory  every owner exports ONE landing buffer per epoch (CUDA IPC, NVLink/NVSwitch); the entity-plan kernel
//                stores a rank's ghost rows straight into their owners' buffers while it runs (the row table of the plan
//                carries (owner, offset) for ghost rows); close() = one NCCL all-reduce of a word (kernel completion on
//                every rank = remote stores visible) + ONE owner-side add kernel over all neighbours.  Two epochs of
//                buffers alternate so that a rank may start the next assembly while a slower owner still reads.
//   send/recv    forms the entity plan does not cover write the ghost rows into the tail of the local value array;
//                close() packs them per owner, ncclSend/ncclRecv (one group), and the same add kernel runs on the
//                receive buffer.
// Both may occur between two close() calls; the library tracks which rows went where.
#include "feelgpu_internal.h"
#include "eplan.h"

#include <nccl.h>
#include <dlfcn.h>

#include <algorithm>
#include <cstring>
#include <numeric>

// NCCL is bound at run time, not through DT_NEEDED: a process that also carries PyTorch (the test/bench harness) has
// torch's own libnccl.so.2 under the same soname, and whichever copy is mapped first serves both -- the system copy is older
// than the one torch was built against, so mapping it early breaks `import torch`.  feelgpu_comm_* therefore take the copy
// that is already mapped (RTLD_NOLOAD) and only otherwise load the system library (a Feel++ process: no torch).
namespace {
struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    bool ok = false;
    std::string error;
};
NcclApi& nccl_api()
{
    static NcclApi api = [] {
        NcclApi a;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { a.error = std::string("cannot load libnccl.so.2: ") + dlerror(); return a; }
        bool all = true;
        auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p) { all = false; a.error = std::string("libnccl: missing ") + name; } return p; };
        a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
        a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
        a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
        a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
        a.AllGather = reinterpret_cast<decltype(a.AllGather)>(sym("ncclAllGather"));
        a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(sym("ncclAllReduce"));
        a.Send = reinterpret_cast<decltype(a.Send)>(sym("ncclSend"));
        a.Recv = reinterpret_cast<decltype(a.Recv)>(sym("ncclRecv"));
        a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
        a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
        a.GetVersion = reinterpret_cast<decltype(a.GetVersion)>(sym("ncclGetVersion"));
        a.ok = all;
        return a;
    }();
    return api;
}
} // namespace
#define ncclGetUniqueId nccl_api().GetUniqueId
#define ncclCommInitRank nccl_api().CommInitRank
#define ncclCommDestroy nccl_api().CommDestroy
#define ncclGetErrorString nccl_api().GetErrorString
#define ncclAllGather nccl_api().AllGather
#define ncclAllReduce nccl_api().AllReduce
#define ncclSend nccl_api().Send
#define ncclRecv nccl_api().Recv
#define ncclGroupStart nccl_api().GroupStart
#define ncclGroupEnd nccl_api().GroupEnd

struct feelgpu_exchange_s {
    feelgpu_context_t ctx = nullptr;
    feelgpu_pattern_t pat = nullptr;
    int nranks = 1, rank = 0;
    int64_t nOwned = 0, nGhost = 0, tail0 = 0;          // dof rows; first value slot of the ghost rows
    // neighbours = ranks that own ghost rows of mine (send side) and ranks that hold ghost copies of my rows (recv side)
    std::vector<int> send_rank, recv_rank;               // ascending
    std::vector<int64_t> send_row0, send_val0;           // [nsend+1] offsets into the packed row / value lists
    std::vector<int64_t> recv_row0, recv_val0;           // [nrecv+1]
    int32_t* d_send_rows = nullptr;                      // ghost rows (local ids), grouped by owner, local order inside
    int64_t* d_send_idx = nullptr;                       // value slot of every packed ghost-row value
    int32_t* d_recv_rows = nullptr;                      // my owned row of every incoming ghost row
    int64_t* d_recv_slots = nullptr;                     // my value slot of every incoming ghost-row value
    double* d_sendbuf = nullptr;                         // [send values] (also rows for vectors: values >= rows)
    double* d_recvbuf = nullptr;                         // [recv values]
    // peer-memory transport
    bool peer = false;
    double* d_land[2] = {nullptr, nullptr};              // my landing buffers (epoch parity), [recv values]
    std::vector<void*> opened;                           // IPC mappings of the owners' buffers (to close)
    int epoch = 0;
    bool tail_used = false, direct_used = false;
    int64_t launches = 0;
};

namespace {

#define NCCL_TRY(ctx, call)                                                                             \
    do {                                                                                                \
        ncclResult_t r__ = (call);                                                                      \
        if (r__ != ncclSuccess) {                                                                       \
            (ctx)->last_error = std::string(#call) + ": " + ncclGetErrorString(r__);                    \
            return FEELGPU_ERR_NCCL;                                                                    \
        }                                                                                               \
    } while (0)

inline ncclComm_t comm_of(feelgpu_context_t ctx) { return static_cast<ncclComm_t>(ctx->comm); }

// slot of every incoming (row, column): the search MatSetValues would do at every assembly, once
__global__ void k_exchange_slots(int64_t n, const int32_t* __restrict__ vrow, const int32_t* __restrict__ vcol,
                                 const int64_t* __restrict__ ia, const int32_t* __restrict__ ja, int64_t* __restrict__ slots,
                                 int* __restrict__ flag)
{
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = vrow[k];
        const int32_t c = vcol[k];
        int64_t lo = ia[r], hi = ia[r + 1];
        const int64_t end = hi;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (ja[mid] < c) lo = mid + 1;
            else hi = mid;
        }
        if (lo >= end || ja[lo] != c) {
            atomicExch(flag, 1);
            slots[k] = 0;
        } else
            slots[k] = lo;
    }
}

__global__ void k_pack(int64_t n, const int64_t* __restrict__ idx, const double* __restrict__ a, double* __restrict__ out)
{
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) out[k] = a[idx[k]];
}
__global__ void k_pack_rows(int64_t n, const int32_t* __restrict__ rows, const double* __restrict__ b, double* __restrict__ out)
{
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) out[k] = b[rows[k]];
}
// owner side: a[slot[k]] += v[k] for the values of ALL neighbours at once (two neighbours may hit one slot: atomics)
__global__ void k_add_slots(int64_t n, const int64_t* __restrict__ slots, const double* __restrict__ v, double* __restrict__ a)
{
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(&a[slots[k]], v[k]);
}
__global__ void k_add_rows(int64_t n, const int32_t* __restrict__ rows, const double* __restrict__ v, double* __restrict__ b)
{
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(&b[rows[k]], v[k]);
}
__global__ void k_set_rows(int64_t n, const int32_t* __restrict__ rows, const double* __restrict__ v, double* __restrict__ b)
{
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) b[rows[k]] = v[k];
}

inline int grid_for(feelgpu_context_t ctx, int64_t n)
{
    return (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16));
}

template <typename T>
cudaError_t to_device(T** d, const std::vector<T>& h)
{
    cudaError_t err = cudaMalloc(d, std::max<size_t>(h.size(), 1) * sizeof(T));
    if (err != cudaSuccess || h.empty()) return err;
    return cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
}

// grouped neighbour exchange of int64 / double payloads living on the device
template <typename T>
ncclResult_t neighbour_exchange(feelgpu_exchange_t ex, const T* sendbuf, const std::vector<int64_t>& s0, T* recvbuf,
                                const std::vector<int64_t>& r0, bool reverse, cudaStream_t s)
{
    // reverse: the owners send, the ghost holders receive (VecGhostUpdate)
    const ncclDataType_t dt = sizeof(T) == 8 ? ncclDouble : ncclInt32;   // 8-byte payloads travel as doubles (bit copies)
    ncclComm_t comm = comm_of(ex->ctx);
    ncclResult_t r = ncclGroupStart();
    if (r != ncclSuccess) return r;
    const std::vector<int>& sr = reverse ? ex->recv_rank : ex->send_rank;
    const std::vector<int>& rr = reverse ? ex->send_rank : ex->recv_rank;
    for (size_t k = 0; k < sr.size() && r == ncclSuccess; ++k)
        if (s0[k + 1] > s0[k]) r = ncclSend(sendbuf + s0[k], (size_t)(s0[k + 1] - s0[k]), dt, sr[k], comm, s);
    for (size_t k = 0; k < rr.size() && r == ncclSuccess; ++k)
        if (r0[k + 1] > r0[k]) r = ncclRecv(recvbuf + r0[k], (size_t)(r0[k + 1] - r0[k]), dt, rr[k], comm, s);
    ncclResult_t e = ncclGroupEnd();
    return r != ncclSuccess ? r : e;
}

} // namespace

extern "C" {

// ---- communicator -------------------------------------------------------------------------------------------------------
int feelgpu_comm_unique_id(unsigned char id[FEELGPU_UNIQUE_ID_BYTES])
{
    if (!id) return FEELGPU_ERR_ARG;
    static_assert(sizeof(ncclUniqueId) <= FEELGPU_UNIQUE_ID_BYTES, "ncclUniqueId size");
    if (!nccl_api().ok) return FEELGPU_ERR_NCCL;
    ncclUniqueId u;
    if (ncclGetUniqueId(&u) != ncclSuccess) return FEELGPU_ERR_NCCL;
    std::memset(id, 0, FEELGPU_UNIQUE_ID_BYTES);
    std::memcpy(id, &u, sizeof(u));
    return FEELGPU_OK;
}

int feelgpu_comm_init(feelgpu_context_t ctx, const unsigned char id[FEELGPU_UNIQUE_ID_BYTES], int rank, int nranks)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    if (!id || nranks < 1 || rank < 0 || rank >= nranks) {
        ctx->last_error = "comm_init: bad argument";
        return FEELGPU_ERR_ARG;
    }
    if (ctx->comm) {
        ctx->last_error = "comm_init: the context already has a communicator";
        return FEELGPU_ERR_ARG;
    }
    if (!nccl_api().ok) {
        ctx->last_error = "comm_init: " + nccl_api().error;
        return FEELGPU_ERR_NCCL;
    }
    FEELGPU_CHECK_CUDA(ctx, cudaSetDevice(ctx->device));
    ncclUniqueId u;
    std::memcpy(&u, id, sizeof(u));
    ncclComm_t comm = nullptr;
    NCCL_TRY(ctx, ncclCommInitRank(&comm, nranks, u, rank));
    ctx->comm = comm;
    ctx->rank = rank;
    ctx->nranks = nranks;
    FEELGPU_CHECK_CUDA(ctx, cudaMalloc(&ctx->d_barrier, sizeof(int)));
    FEELGPU_CHECK_CUDA(ctx, cudaMemset(ctx->d_barrier, 0, sizeof(int)));
    return FEELGPU_OK;
}

int feelgpu_comm_destroy(feelgpu_context_t ctx)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    if (ctx->comm) {
        cudaSetDevice(ctx->device);
        cudaDeviceSynchronize();
        ncclCommDestroy(comm_of(ctx));
        ctx->comm = nullptr;
    }
    if (ctx->d_barrier) cudaFree(ctx->d_barrier);
    ctx->d_barrier = nullptr;
    ctx->nranks = 1;
    ctx->rank = 0;
    return FEELGPU_OK;
}

// ---- exchange plan ------------------------------------------------------------------------------------------------------
int feelgpu_exchange_destroy(feelgpu_exchange_t ex)
{
    if (!ex) return FEELGPU_ERR_ARG;
    cudaSetDevice(ex->ctx->device);
    cudaDeviceSynchronize();
    if (ex->pat && ex->pat->exchange == ex) {
        if (ex->pat->eplan) feelgpu_eplan_set_remote(ex->pat, 0, nullptr, nullptr, nullptr, nullptr);
        ex->pat->exchange = nullptr;
    }
    for (void* p : ex->opened) cudaIpcCloseMemHandle(p);
    cudaFree(ex->d_send_rows);
    cudaFree(ex->d_send_idx);
    cudaFree(ex->d_recv_rows);
    cudaFree(ex->d_recv_slots);
    cudaFree(ex->d_sendbuf);
    cudaFree(ex->d_recvbuf);
    cudaFree(ex->d_land[0]);
    cudaFree(ex->d_land[1]);
    delete ex;
    return FEELGPU_OK;
}

int feelgpu_exchange_create(feelgpu_pattern_t pat, const int32_t* ghost_owner, int64_t first_dof_cluster, int flags,
                            feelgpu_exchange_t* out)
{
    if (!pat) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = pat->ctx;
    if (!out) { ctx->last_error = "exchange_create: null output"; return FEELGPU_ERR_ARG; }
    *out = nullptr;
    if (!ctx->comm) { ctx->last_error = "exchange_create: no communicator (feelgpu_comm_init)"; return FEELGPU_ERR_ARG; }
    if (pat->exchange) { ctx->last_error = "exchange_create: the pattern already has an exchange"; return FEELGPU_ERR_ARG; }
    if (int prc = feelgpu_eplan_poll(pat, true)) return prc;     // the peer-memory transport patches the entity plan
    feelgpu_space_t sp = pat->test;
    const int P = ctx->nranks, me = ctx->rank;
    const int64_t nO = sp->nDofOwned, nG = pat->nRows - nO;
    if (nG < 0) { ctx->last_error = "exchange_create: the pattern has fewer rows than owned dofs"; return FEELGPU_ERR_ARG; }
    if (nG > 0 && (!ghost_owner || sp->h_p2c.empty())) {
        ctx->last_error = "exchange_create: ghost rows need their owners and the process->cluster map of the space";
        return FEELGPU_ERR_ARG;
    }
    for (int64_t g = 0; g < nG; ++g)
        if (ghost_owner[g] < 0 || ghost_owner[g] >= P || ghost_owner[g] == me) {
            ctx->last_error = "exchange_create: ghost_owner must name another rank of the communicator";
            return FEELGPU_ERR_ARG;
        }
    if (!pat->d_ja32) { ctx->last_error = "exchange_create: the pattern holds column ids beyond 32 bits"; return FEELGPU_ERR_UNSUPPORTED; }
    FEELGPU_CHECK_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    feelgpu_exchange_t ex = new feelgpu_exchange_s;
    ex->ctx = ctx;
    ex->pat = pat;
    ex->nranks = P;
    ex->rank = me;
    ex->nOwned = nO;
    ex->nGhost = nG;
    ex->tail0 = pat->h_ia[nO];
    const std::vector<int64_t>& ia = pat->h_ia;
    int rc = FEELGPU_OK;
    int64_t *d_cnt = nullptr, *d_all = nullptr, *d_msg = nullptr, *d_in = nullptr;
    int32_t *d_vrow = nullptr, *d_vcol = nullptr;
    unsigned char *d_h = nullptr, *d_hall = nullptr;
    auto cleanup = [&]() {
        cudaFree(d_cnt); cudaFree(d_all); cudaFree(d_msg); cudaFree(d_in); cudaFree(d_vrow); cudaFree(d_vcol);
        cudaFree(d_h); cudaFree(d_hall);
    };
#define EX_CUDA(call)                                                                     \
    do {                                                                                  \
        cudaError_t e__ = (call);                                                         \
        if (e__ != cudaSuccess) {                                                         \
            ctx->last_error = std::string("exchange_create: ") + cudaGetErrorString(e__); \
            cleanup();                                                                    \
            feelgpu_exchange_destroy(ex);                                                 \
            return FEELGPU_ERR_CUDA;                                                      \
        }                                                                                 \
    } while (0)
#define EX_NCCL(call)                                                                     \
    do {                                                                                  \
        ncclResult_t r__ = (call);                                                        \
        if (r__ != ncclSuccess) {                                                         \
            ctx->last_error = std::string("exchange_create: ") + ncclGetErrorString(r__); \
            cleanup();                                                                    \
            feelgpu_exchange_destroy(ex);                                                 \
            return FEELGPU_ERR_NCCL;                                                      \
        }                                                                                 \
    } while (0)
    try {
        // ---- my ghost rows grouped by owner, local order inside a group ----
        std::vector<int64_t> rows_of(P, 0), vals_of(P, 0);
        for (int64_t g = 0; g < nG; ++g) {
            rows_of[ghost_owner[g]]++;
            vals_of[ghost_owner[g]] += ia[nO + g + 1] - ia[nO + g];
        }
        // counts[src][dst] = (rows, values): all-gathered so that every rank knows what it receives and where it lands
        std::vector<int64_t> cnt(2 * (size_t)P), all(2 * (size_t)P * P);
        for (int q = 0; q < P; ++q) { cnt[2 * q] = rows_of[q]; cnt[2 * q + 1] = vals_of[q]; }
        EX_CUDA(cudaMalloc(&d_cnt, cnt.size() * 8));
        EX_CUDA(cudaMalloc(&d_all, all.size() * 8));
        EX_CUDA(cudaMemcpyAsync(d_cnt, cnt.data(), cnt.size() * 8, cudaMemcpyHostToDevice, s));
        EX_NCCL(ncclAllGather(d_cnt, d_all, cnt.size(), ncclInt64, comm_of(ctx), s));
        EX_CUDA(cudaMemcpyAsync(all.data(), d_all, all.size() * 8, cudaMemcpyDeviceToHost, s));
        EX_CUDA(cudaStreamSynchronize(s));
        auto rows_sd = [&](int src, int dst) { return all[2 * ((size_t)src * P + dst)]; };
        auto vals_sd = [&](int src, int dst) { return all[2 * ((size_t)src * P + dst) + 1]; };
        ex->send_row0.push_back(0);
        ex->send_val0.push_back(0);
        for (int q = 0; q < P; ++q)
            if (rows_of[q] > 0) {
                ex->send_rank.push_back(q);
                ex->send_row0.push_back(ex->send_row0.back() + rows_of[q]);
                ex->send_val0.push_back(ex->send_val0.back() + vals_of[q]);
            }
        ex->recv_row0.push_back(0);
        ex->recv_val0.push_back(0);
        for (int p = 0; p < P; ++p)
            if (p != me && rows_sd(p, me) > 0) {
                ex->recv_rank.push_back(p);
                ex->recv_row0.push_back(ex->recv_row0.back() + rows_sd(p, me));
                ex->recv_val0.push_back(ex->recv_val0.back() + vals_sd(p, me));
            }
        const int nsend = (int)ex->send_rank.size(), nrecv = (int)ex->recv_rank.size();
        const int64_t nsr = ex->send_row0.back(), nsv = ex->send_val0.back();
        const int64_t nrr = ex->recv_row0.back(), nrv = ex->recv_val0.back();
        std::vector<int> nb_of(P, -1);
        for (int k = 0; k < nsend; ++k) nb_of[ex->send_rank[k]] = k;
        // packed lists + the message to every owner: [row cluster ids | row lengths | columns] as int64
        std::vector<int32_t> send_rows((size_t)nsr), ghost_nb((size_t)nG);
        std::vector<int64_t> send_idx((size_t)nsv), ghost_off((size_t)nG);
        std::vector<int64_t> msg0(nsend + 1, 0);
        for (int k = 0; k < nsend; ++k)
            msg0[k + 1] = msg0[k] + 2 * (ex->send_row0[k + 1] - ex->send_row0[k]) + (ex->send_val0[k + 1] - ex->send_val0[k]);
        std::vector<int64_t> msg((size_t)msg0[nsend]);
        {
            std::vector<int64_t> rat(ex->send_row0.begin(), ex->send_row0.end() - 1), vat(ex->send_val0.begin(), ex->send_val0.end() - 1);
            for (int64_t g = 0; g < nG; ++g) {
                const int k = nb_of[ghost_owner[g]];
                const int64_t r = nO + g, len = ia[r + 1] - ia[r];
                const int64_t nr = ex->send_row0[k + 1] - ex->send_row0[k];
                const int64_t ir = rat[k] - ex->send_row0[k], iv = vat[k] - ex->send_val0[k];
                int64_t* m = msg.data() + msg0[k];
                m[ir] = sp->h_p2c[(size_t)r];
                m[nr + ir] = len;
                for (int64_t j = 0; j < len; ++j) {
                    m[2 * nr + iv + j] = pat->h_ja_tail[(size_t)(ia[r] - ex->tail0 + j)];
                    send_idx[(size_t)(vat[k] + j)] = ia[r] + j;
                }
                send_rows[(size_t)rat[k]] = (int32_t)r;
                ghost_nb[(size_t)g] = k;
                ghost_off[(size_t)g] = iv;              // offset inside my block of the owner's landing buffer (block base added below)
                rat[k]++;
                vat[k] += len;
            }
        }
        // ---- ship the messages ----
        std::vector<int64_t> in0(nrecv + 1, 0);
        for (int k = 0; k < nrecv; ++k)
            in0[k + 1] = in0[k] + 2 * (ex->recv_row0[k + 1] - ex->recv_row0[k]) + (ex->recv_val0[k + 1] - ex->recv_val0[k]);
        std::vector<int64_t> in((size_t)in0[nrecv]);
        EX_CUDA(cudaMalloc(&d_msg, std::max<size_t>(msg.size(), 1) * 8));
        EX_CUDA(cudaMalloc(&d_in, std::max<size_t>(in.size(), 1) * 8));
        if (!msg.empty()) EX_CUDA(cudaMemcpyAsync(d_msg, msg.data(), msg.size() * 8, cudaMemcpyHostToDevice, s));
        EX_NCCL(neighbour_exchange(ex, reinterpret_cast<const double*>(d_msg), msg0, reinterpret_cast<double*>(d_in), in0, false, s));
        if (!in.empty()) EX_CUDA(cudaMemcpyAsync(in.data(), d_in, in.size() * 8, cudaMemcpyDeviceToHost, s));
        EX_CUDA(cudaStreamSynchronize(s));
        // ---- owner side: my row and slot of every incoming value ----
        std::vector<int32_t> recv_rows((size_t)nrr), vrow((size_t)nrv), vcol((size_t)nrv);
        for (int k = 0; k < nrecv; ++k) {
            const int64_t nr = ex->recv_row0[k + 1] - ex->recv_row0[k];
            const int64_t* m = in.data() + in0[k];
            int64_t v = ex->recv_val0[k];
            for (int64_t i = 0; i < nr; ++i) {
                const int64_t lrow = m[i] - first_dof_cluster, len = m[nr + i];
                if (lrow < 0 || lrow >= nO) {
                    ctx->last_error = "exchange_create: received a ghost row this rank does not own";
                    cleanup();
                    feelgpu_exchange_destroy(ex);
                    return FEELGPU_ERR_ARG;
                }
                recv_rows[(size_t)(ex->recv_row0[k] + i)] = (int32_t)lrow;
                for (int64_t j = 0; j < len; ++j, ++v) {
                    vrow[(size_t)v] = (int32_t)lrow;
                    vcol[(size_t)v] = (int32_t)m[2 * nr + (v - ex->recv_val0[k])];
                }
            }
            if (v != ex->recv_val0[k + 1]) {
                ctx->last_error = "exchange_create: inconsistent ghost-row message";
                cleanup();
                feelgpu_exchange_destroy(ex);
                return FEELGPU_ERR_ARG;
            }
        }
        EX_CUDA(to_device(&ex->d_send_rows, send_rows));
        EX_CUDA(to_device(&ex->d_send_idx, send_idx));
        EX_CUDA(to_device(&ex->d_recv_rows, recv_rows));
        EX_CUDA(to_device(&d_vrow, vrow));
        EX_CUDA(to_device(&d_vcol, vcol));
        EX_CUDA(cudaMalloc(&ex->d_recv_slots, std::max<int64_t>(nrv, 1) * 8));
        EX_CUDA(cudaMalloc(&ex->d_sendbuf, std::max<int64_t>(std::max(nsv, nsr), std::max(nrr, (int64_t)1)) * 8));
        EX_CUDA(cudaMalloc(&ex->d_recvbuf, std::max<int64_t>(std::max(nrv, nrr), std::max(nsr, (int64_t)1)) * 8));
        int flag = 0;
        if (nrv > 0) {
            EX_CUDA(cudaMemsetAsync(ctx->d_flag, 0, sizeof(int), s));
            k_exchange_slots<<<grid_for(ctx, nrv), 256, 0, s>>>(nrv, d_vrow, d_vcol, pat->d_ia, pat->d_ja32, ex->d_recv_slots, ctx->d_flag);
            ctx->launches++;
            EX_CUDA(cudaGetLastError());
            EX_CUDA(cudaMemcpyAsync(&flag, ctx->d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
            EX_CUDA(cudaStreamSynchronize(s));
        }
        // every rank learns whether some owner misses an entry (the calls below are collective)
        {
            int* d_f = ctx->d_barrier;
            EX_CUDA(cudaMemcpyAsync(d_f, &flag, sizeof(int), cudaMemcpyHostToDevice, s));
            EX_NCCL(ncclAllReduce(d_f, d_f, 1, ncclInt32, ncclMax, comm_of(ctx), s));
            EX_CUDA(cudaMemcpyAsync(&flag, d_f, sizeof(int), cudaMemcpyDeviceToHost, s));
            EX_CUDA(cudaStreamSynchronize(s));
            if (flag) {
                ctx->last_error = "exchange_create: a ghost-row entry is not in its owner's graph";
                cleanup();
                feelgpu_exchange_destroy(ex);
                return FEELGPU_ERR_NOT_IN_PATTERN;
            }
        }
        // ---- peer-memory transport: one landing buffer per epoch, exported to everybody ----
        const bool want_peer = !(flags & FEELGPU_EXCHANGE_NO_PEER_MEMORY);
        int ok = want_peer && pat->eplan && nsend <= EPLAN_MAXREMOTE ? 1 : 0;
        cudaIpcMemHandle_t hmine[2];
        std::memset(hmine, 0, sizeof(hmine));
        if (want_peer) {      // (collective from here on: every rank goes through the same calls)
            for (int e = 0; e < 2 && ok; ++e) {
                if (cudaMalloc(&ex->d_land[e], std::max<int64_t>(nrv, 1) * 8) != cudaSuccess ||
                    cudaIpcGetMemHandle(&hmine[e], ex->d_land[e]) != cudaSuccess) {
                    cudaGetLastError();
                    ok = 0;
                }
            }
            std::vector<unsigned char> hall((size_t)P * 128);
            EX_CUDA(cudaMalloc(&d_h, 128));
            EX_CUDA(cudaMalloc(&d_hall, hall.size()));
            EX_CUDA(cudaMemcpyAsync(d_h, hmine, 128, cudaMemcpyHostToDevice, s));
            EX_NCCL(ncclAllGather(d_h, d_hall, 128, ncclUint8, comm_of(ctx), s));
            EX_CUDA(cudaMemcpyAsync(hall.data(), d_hall, hall.size(), cudaMemcpyDeviceToHost, s));
            EX_CUDA(cudaStreamSynchronize(s));
            double* rptr[2][EPLAN_MAXREMOTE];
            for (int e = 0; e < 2; ++e)
                for (int k = 0; k < EPLAN_MAXREMOTE; ++k) rptr[e][k] = nullptr;
            for (int k = 0; k < nsend && ok; ++k) {
                const int q = ex->send_rank[k];
                // my block inside q's landing buffer starts after the blocks of the lower ranks that send to q
                int64_t off = 0;
                for (int p = 0; p < me; ++p)
                    if (p != q) off += vals_sd(p, q);
                for (int e = 0; e < 2 && ok; ++e) {
                    cudaIpcMemHandle_t h;
                    std::memcpy(&h, hall.data() + (size_t)q * 128 + 64 * e, 64);
                    void* p = nullptr;
                    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                        cudaGetLastError();
                        ok = 0;
                        break;
                    }
                    ex->opened.push_back(p);
                    rptr[e][k] = static_cast<double*>(p) + off;
                }
            }
            // all ranks or none
            int* d_f = ctx->d_barrier;
            EX_CUDA(cudaMemcpyAsync(d_f, &ok, sizeof(int), cudaMemcpyHostToDevice, s));
            EX_NCCL(ncclAllReduce(d_f, d_f, 1, ncclInt32, ncclMin, comm_of(ctx), s));
            EX_CUDA(cudaMemcpyAsync(&ok, d_f, sizeof(int), cudaMemcpyDeviceToHost, s));
            EX_CUDA(cudaStreamSynchronize(s));
            const int zero = 0;
            EX_CUDA(cudaMemcpyAsync(d_f, &zero, sizeof(int), cudaMemcpyHostToDevice, s));
            EX_CUDA(cudaStreamSynchronize(s));
            if (ok) {
                rc = feelgpu_eplan_set_remote(pat, nsend, ghost_nb.data(), ghost_off.data(), rptr[0], rptr[1]);
                if (rc) { cleanup(); feelgpu_exchange_destroy(ex); return rc; }
                ex->peer = true;
            } else {
                for (void* p : ex->opened) cudaIpcCloseMemHandle(p);
                ex->opened.clear();
                cudaFree(ex->d_land[0]);
                cudaFree(ex->d_land[1]);
                ex->d_land[0] = ex->d_land[1] = nullptr;
            }
        }
    } catch (const std::exception& e) {
        ctx->last_error = std::string("exchange_create: ") + e.what();
        cleanup();
        feelgpu_exchange_destroy(ex);
        return FEELGPU_ERR_HOST;
    }
#undef EX_CUDA
#undef EX_NCCL
    cleanup();
    pat->exchange = ex;
    *out = ex;
    return FEELGPU_OK;
}

int feelgpu_exchange_info(feelgpu_exchange_t ex, int64_t info[8])
{
    if (!ex || !info) return FEELGPU_ERR_ARG;
    info[0] = (int64_t)ex->send_rank.size();
    info[1] = ex->send_row0.back();
    info[2] = ex->send_val0.back();
    info[3] = (int64_t)ex->recv_rank.size();
    info[4] = ex->recv_row0.back();
    info[5] = ex->recv_val0.back();
    info[6] = ex->peer ? 1 : 0;
    info[7] = ex->epoch;
    return FEELGPU_OK;
}

// ---- close() ------------------------------------------------------------------------------------------------------------
int feelgpu_form2_close(feelgpu_pattern_t pat, double* a, void* stream)
{
    if (!pat) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = pat->ctx;
    feelgpu_exchange_t ex = pat->exchange;
    if (!ex) {
        if (ctx->nranks == 1 || pat->nRows == pat->test->nDofOwned) return FEELGPU_OK;   // sequential matrix: nothing to ship
        ctx->last_error = "form2_close: the pattern has ghost rows but no exchange (feelgpu_exchange_create)";
        return FEELGPU_ERR_ARG;
    }
    if (!a) { ctx->last_error = "form2_close: null value array"; return FEELGPU_ERR_ARG; }
    FEELGPU_CHECK_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    const int64_t nsv = ex->send_val0.back(), nrv = ex->recv_val0.back();
    if (ex->peer && ex->direct_used) {
        // every rank's kernel is complete = its stores into my landing buffer are visible
        NCCL_TRY(ctx, ncclAllReduce(ctx->d_barrier, ctx->d_barrier, 1, ncclInt32, ncclSum, comm_of(ctx), s));
        if (nrv > 0) {
            k_add_slots<<<grid_for(ctx, nrv), 256, 0, s>>>(nrv, ex->d_recv_slots, ex->d_land[ex->epoch & 1], a);
            ctx->launches++;
            FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
        }
        ex->epoch++;
        feelgpu_eplan_remote_epoch(pat, ex->epoch & 1);
        ex->direct_used = false;
    }
    if (ex->tail_used || !ex->peer) {
        if (nsv > 0) {
            k_pack<<<grid_for(ctx, nsv), 256, 0, s>>>(nsv, ex->d_send_idx, a, ex->d_sendbuf);
            ctx->launches++;
            FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
        }
        NCCL_TRY(ctx, neighbour_exchange(ex, ex->d_sendbuf, ex->send_val0, ex->d_recvbuf, ex->recv_val0, false, s));
        if (nrv > 0) {
            k_add_slots<<<grid_for(ctx, nrv), 256, 0, s>>>(nrv, ex->d_recv_slots, ex->d_recvbuf, a);
            ctx->launches++;
            FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
        }
        ex->tail_used = false;
    }
    return FEELGPU_OK;
}

int feelgpu_form1_close(feelgpu_exchange_t ex, double* b, int ghost_update, void* stream)
{
    if (!ex) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = ex->ctx;
    if (!b) { ctx->last_error = "form1_close: null vector"; return FEELGPU_ERR_ARG; }
    FEELGPU_CHECK_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    const int64_t nsr = ex->send_row0.back(), nrr = ex->recv_row0.back();
    // VecAssemblyBegin/End: the ghost entries go to their owners and are added there
    if (nsr > 0) {
        k_pack_rows<<<grid_for(ctx, nsr), 256, 0, s>>>(nsr, ex->d_send_rows, b, ex->d_sendbuf);
        ctx->launches++;
    }
    NCCL_TRY(ctx, neighbour_exchange(ex, ex->d_sendbuf, ex->send_row0, ex->d_recvbuf, ex->recv_row0, false, s));
    if (nrr > 0) {
        k_add_rows<<<grid_for(ctx, nrr), 256, 0, s>>>(nrr, ex->d_recv_rows, ex->d_recvbuf, b);
        ctx->launches++;
    }
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    if (ghost_update) {
        // VecGhostUpdate(INSERT_VALUES, SCATTER_FORWARD): the ghost copies take the owners' final values
        if (nrr > 0) {
            k_pack_rows<<<grid_for(ctx, nrr), 256, 0, s>>>(nrr, ex->d_recv_rows, b, ex->d_sendbuf);
            ctx->launches++;
        }
        NCCL_TRY(ctx, neighbour_exchange(ex, ex->d_sendbuf, ex->recv_row0, ex->d_recvbuf, ex->send_row0, true, s));
        if (nsr > 0) {
            k_set_rows<<<grid_for(ctx, nsr), 256, 0, s>>>(nsr, ex->d_send_rows, ex->d_recvbuf, b);
            ctx->launches++;
        }
        FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    }
    return FEELGPU_OK;
}

} // extern "C"

// called by feelgpu_form2_assemble: where do the ghost rows of this assembly go?
void feelgpu_exchange_note_assembly(feelgpu_pattern_t pat, bool through_eplan, int accumulate, double* a, cudaStream_t s)
{
    feelgpu_exchange_t ex = pat->exchange;
    if (!ex) return;
    if (through_eplan && ex->peer) {
        if (!accumulate) ex->tail_used = false;          // '=' voids what the local tail held
        ex->direct_used = true;
    } else {
        if (ex->peer && !accumulate) {
            ex->direct_used = false;                     // '=' voids the landing buffers: they count as empty again
            feelgpu_eplan_remote_fresh(pat);
        }
        if (accumulate && !ex->tail_used && pat->nnz > ex->tail0)
            cudaMemsetAsync(a + ex->tail0, 0, sizeof(double) * (size_t)(pat->nnz - ex->tail0), s);   // the tail held nothing yet
        ex->tail_used = true;
    }
}

void feelgpu_exchange_detach(feelgpu_pattern_t pat)
{
    if (pat->exchange) feelgpu_exchange_destroy(pat->exchange);
}
