// kernels_graph.cu -- the COUPLED sparsity graph of one space, built on the device.
//
// Replaces Stencil::computeGraph + GraphCSR::close for test space == trial space (feeldiscr/stencil.hpp:1500-1798: every
// row dof of an element is coupled with every column dof of that element, kept in a std::set per row;
// feelalg/graphcsr.cpp:662-1022: rows flattened to ia/ja with ascending columns, the diagonal always present :633-659).
// At 8 M P2 elements the host std::set graph costs far more than the assembly; here:
//   1. one 64-bit key (row node << 32 | column node) per (element, i, j)       -- k_graph_keys
//   2. radix sort + unique of the keys (CUB: library code, setup path)          -- the set semantics, ascending columns
//   3. rows counted, exclusive scan -> ia; low words -> ja                      -- k_graph_count / k_graph_fill
// Product spaces (global dof = ncomp * node + c, feeldiscr/doftable.hpp:1289-1292) are built at node level and expanded:
// every node entry becomes an ncomp x ncomp block, columns stay ascending.
#include <string>
#include "feelgpu_internal.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>

#include <algorithm>
#include <new>

// (struct feelgpu_graph_s: feelgpu_internal.h)

namespace {

__global__ void k_graph_keys(int64_t total, int n, int nd, int ncomp, const int32_t* __restrict__ l2g,
                             uint64_t* __restrict__ keys)
{
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t e = t / (n * n);
        const int r = (int)(t - e * n * n);
        const int i = r / n, j = r - i * n;
        const uint64_t row = (uint64_t)(l2g[e * nd + i] / ncomp), col = (uint64_t)(l2g[e * nd + j] / ncomp);
        keys[t] = (row << 32) | col;
    }
}

__global__ void k_graph_count(int64_t nUnique, const uint64_t* __restrict__ keys, int64_t* __restrict__ cnt)
{
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < nUnique; k += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(reinterpret_cast<unsigned long long*>(&cnt[keys[k] >> 32]), 1ull);
}

// node graph (ian, keys) -> dof graph: row ncomp*node + c holds, for every node column m, the columns ncomp*m + c2
__global__ void k_graph_fill(int64_t nNode, int ncomp, const int64_t* __restrict__ ian, const uint64_t* __restrict__ keys,
                             int64_t* __restrict__ ia, int64_t* __restrict__ ja)
{
    const int lane = threadIdx.x & 31;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t node = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; node < nNode; node += nw) {
        const int64_t p0 = ian[node], len = ian[node + 1] - p0;
        for (int c = 0; c < ncomp; ++c) {
            const int64_t base = (int64_t)ncomp * ncomp * p0 + (int64_t)c * ncomp * len;
            if (lane == 0) ia[node * ncomp + c] = base;
            for (int64_t k = lane; k < len * ncomp; k += 32)
                ja[base + k] = (int64_t)ncomp * (int64_t)(keys[p0 + k / ncomp] & 0xFFFFFFFFull) + k % ncomp;
        }
        if (node == nNode - 1 && lane == 0) ia[nNode * ncomp] = (int64_t)ncomp * ncomp * ian[nNode];
    }
}

} // namespace

#define GRAPH_CUDA(call)                                                                                \
    do {                                                                                                \
        cudaError_t err__ = (call);                                                                     \
        if (err__ != cudaSuccess) {                                                                     \
            ctx->last_error = std::string(#call) + ": " + cudaGetErrorString(err__);                    \
            rc = FEELGPU_ERR_CUDA;                                                                      \
            goto done;                                                                                  \
        }                                                                                               \
    } while (0)

extern "C" int feelgpu_graph_create(feelgpu_space_t sp, feelgpu_graph_t* out, int64_t* nnz)
{
    if (!sp) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = sp->ctx;
    if (!out || !nnz) {
        ctx->last_error = "graph_create: null argument";
        return FEELGPU_ERR_ARG;
    }
    if (sp->d_p2c) {
        ctx->last_error = "graph_create: partition-local spaces get their graph from the caller (ghost rows, cluster columns)";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    int rc = FEELGPU_OK;
    const int n = sp->n, ncomp = sp->ncomp;
    const int64_t nElt = sp->mesh->nElt, nNode = sp->nDofWithGhost / ncomp, total = nElt * n * n;
    cudaStream_t s = ctx->stream;
    uint64_t *d_keys = nullptr, *d_sorted = nullptr, *d_uniq = nullptr;
    int64_t *d_cnt = nullptr, *d_ian = nullptr, *d_nuniq = nullptr;
    void* d_tmp = nullptr;
    size_t tmp_bytes = 0, need = 0;
    int64_t nUnique = 0;
    feelgpu_graph_t g = nullptr;
    int bits = 1;
    while (bits < 32 && ((int64_t)1 << bits) < std::max<int64_t>(nNode, 2)) ++bits;
    const int grid = (int)std::min<int64_t>(std::max<int64_t>((total + 255) / 256, 1), (int64_t)ctx->sm_count * 32);

    GRAPH_CUDA(cudaSetDevice(ctx->device));
    GRAPH_CUDA(cudaMalloc(&d_keys, sizeof(uint64_t) * std::max<int64_t>(total, 1)));
    GRAPH_CUDA(cudaMalloc(&d_sorted, sizeof(uint64_t) * std::max<int64_t>(total, 1)));
    GRAPH_CUDA(cudaMalloc(&d_nuniq, sizeof(int64_t)));
    k_graph_keys<<<grid, 256, 0, s>>>(total, n, sp->nd, ncomp, sp->d_l2g, d_keys);
    ctx->launches++;
    // temporary storage: the largest of the three CUB calls
    GRAPH_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, need, d_keys, d_sorted, total, 0, 32 + bits, s));
    tmp_bytes = need;
    GRAPH_CUDA(cub::DeviceSelect::Unique(nullptr, need, d_sorted, d_keys, d_nuniq, total, s));
    tmp_bytes = std::max(tmp_bytes, need);
    GRAPH_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, d_cnt, d_ian, nNode + 1, s));
    tmp_bytes = std::max(tmp_bytes, need);
    GRAPH_CUDA(cudaMalloc(&d_tmp, std::max<size_t>(tmp_bytes, 16)));
    need = tmp_bytes;
    GRAPH_CUDA(cub::DeviceRadixSort::SortKeys(d_tmp, need, d_keys, d_sorted, total, 0, 32 + bits, s));
    d_uniq = d_keys;   // the unsorted keys are no longer needed
    need = tmp_bytes;
    GRAPH_CUDA(cub::DeviceSelect::Unique(d_tmp, need, d_sorted, d_uniq, d_nuniq, total, s));
    GRAPH_CUDA(cudaMemcpyAsync(&nUnique, d_nuniq, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    GRAPH_CUDA(cudaStreamSynchronize(s));
    GRAPH_CUDA(cudaFree(d_sorted));
    d_sorted = nullptr;
    GRAPH_CUDA(cudaMalloc(&d_cnt, sizeof(int64_t) * (nNode + 1)));
    GRAPH_CUDA(cudaMalloc(&d_ian, sizeof(int64_t) * (nNode + 1)));
    GRAPH_CUDA(cudaMemsetAsync(d_cnt, 0, sizeof(int64_t) * (nNode + 1), s));
    k_graph_count<<<grid, 256, 0, s>>>(nUnique, d_uniq, d_cnt);
    ctx->launches++;
    need = tmp_bytes;
    GRAPH_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp, need, d_cnt, d_ian, nNode + 1, s));
    g = new (std::nothrow) feelgpu_graph_s;
    if (!g) {
        ctx->last_error = "graph_create: out of host memory";
        rc = FEELGPU_ERR_HOST;
        goto done;
    }
    g->space = sp;
    g->nRows = nNode * ncomp;
    g->nnz = nUnique * ncomp * ncomp;
    GRAPH_CUDA(cudaMalloc(&g->d_ia, sizeof(int64_t) * (g->nRows + 1)));
    GRAPH_CUDA(cudaMalloc(&g->d_ja, sizeof(int64_t) * std::max<int64_t>(g->nnz, 1)));
    if (nNode > 0) {
        const int fgrid = (int)std::min<int64_t>((nNode * 32 + 255) / 256, (int64_t)ctx->sm_count * 32);
        k_graph_fill<<<fgrid, 256, 0, s>>>(nNode, ncomp, d_ian, d_uniq, g->d_ia, g->d_ja);
        ctx->launches++;
    } else
        GRAPH_CUDA(cudaMemsetAsync(g->d_ia, 0, sizeof(int64_t), s));
    GRAPH_CUDA(cudaGetLastError());
    GRAPH_CUDA(cudaStreamSynchronize(s));
    *out = g;
    *nnz = g->nnz;
    g = nullptr;
done:
    cudaFree(d_keys);
    cudaFree(d_sorted);
    cudaFree(d_nuniq);
    cudaFree(d_cnt);
    cudaFree(d_ian);
    cudaFree(d_tmp);
    if (g) {
        cudaFree(g->d_ia);
        cudaFree(g->d_ja);
        delete g;
    }
    return rc;
}

extern "C" int feelgpu_graph_get(feelgpu_graph_t g, int64_t* ia, int64_t* ja)
{
    if (!g) return FEELGPU_ERR_ARG;
    feelgpu_context_t ctx = g->space->ctx;
    if (ia && cudaMemcpy(ia, g->d_ia, sizeof(int64_t) * (g->nRows + 1), cudaMemcpyDeviceToHost) != cudaSuccess) {
        ctx->last_error = "graph_get: copy of ia failed";
        return FEELGPU_ERR_CUDA;
    }
    if (ja && g->nnz && cudaMemcpy(ja, g->d_ja, sizeof(int64_t) * g->nnz, cudaMemcpyDeviceToHost) != cudaSuccess) {
        ctx->last_error = "graph_get: copy of ja failed";
        return FEELGPU_ERR_CUDA;
    }
    return FEELGPU_OK;
}

extern "C" int feelgpu_graph_destroy(feelgpu_graph_t g)
{
    if (!g) return FEELGPU_ERR_ARG;
    cudaFree(g->d_ia);
    cudaFree(g->d_ja);
    delete g;
    return FEELGPU_OK;
}
