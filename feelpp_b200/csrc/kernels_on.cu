// kernels_on.cu -- Dirichlet elimination on the assembled CSR: form2 += on(_range=boundaryfaces(mesh), _rhs=l, _element=u,
// _expr=g).  The reference collects the boundary dofs and their values on the host (feelvf/on.hpp:395-449) and calls
// MatrixPetsc::zeroRows (feelalg/matrixpetsc.cpp:1762-1860), i.e. PETSc's MatZeroRows / MatZeroRowsColumns with the
// value vector x and the right-hand side b:
//     ELIMINATION            rows r: a_r* = 0, a_rr = d,  b_r = d x_r
//     SYMMETRIC              additionally columns: b_i -= a_ir x_r, a_ir = 0 for every other row i
//     KEEP_DIAGONAL          d = the old a_rr instead of value_on_diagonal (matrixpetsc.cpp:1817-1829)
// Here the matrix stays on the device: one warp per eliminated row.  The column pass uses the structural symmetry of the
// COUPLED pattern of one space: the transpose of entry (r,c) is found by a binary search in row c.
#include <string>
#include "feelgpu_internal.h"

#include <algorithm>
#include <cstdlib>
#include <vector>

namespace {

__global__ void k_on_mark(int n, const int32_t* __restrict__ rows, int32_t* __restrict__ mark)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) mark[rows[k]] = k + 1;
}

// SYMMETRIC: for every eliminated row r and every entry (r,c) of it: the transpose (c,r) is zeroed and, when c itself is
// not eliminated, b_c -= a_cr x_r
__global__ void k_on_columns(int n, const int32_t* __restrict__ rows, const double* __restrict__ x,
                             const int32_t* __restrict__ mark, const int64_t* __restrict__ ia,
                             const int32_t* __restrict__ ja, double* __restrict__ a, double* __restrict__ b)
{
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= n) return;
    const int32_t r = rows[w];
    const double xr = x[w];
    for (int64_t p = ia[r] + lane; p < ia[r + 1]; p += 32) {
        const int32_t c = ja[p];
        if (c == r) continue;
        int64_t lo = ia[c], hi = ia[c + 1];
        const int64_t end = hi;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (ja[mid] < r) lo = mid + 1;
            else hi = mid;
        }
        if (lo < end && ja[lo] == r) {
            if (b && !mark[c]) atomicAdd(&b[c], -a[lo] * xr);
            a[lo] = 0.0;
        }
    }
}

__global__ void k_on_rows(int n, const int32_t* __restrict__ rows, const double* __restrict__ x,
                          const int64_t* __restrict__ ia, const int32_t* __restrict__ ja, const int64_t* __restrict__ p2c,
                          double* __restrict__ a, double* __restrict__ b, int keep_diagonal, double value_on_diagonal)
{
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= n) return;
    const int32_t r = rows[w];
    const int32_t dcol = p2c ? (int32_t)p2c[r] : r;   // columns are global-cluster ids (feelalg/datamap.hpp:467)
    double d = value_on_diagonal;
    for (int64_t p = ia[r] + lane; p < ia[r + 1]; p += 32) {
        if (ja[p] == dcol) {
            if (keep_diagonal) d = a[p];
            else a[p] = d;
            if (b) b[r] = x[w] * d;      // exactly one lane holds the diagonal (GraphCSR always inserts it, graphcsr.cpp:633-659)
        } else
            a[p] = 0.0;
    }
}

// y = A x on the device CSR (MatrixPetsc::multVector -> MatMult, feelalg/matrixpetsc.cpp:998-1062): one warp per row,
// 32-bit column ids, x through the read-only path (it stays in L2: 8 B per dof against 12 B per nonzero streamed)
// Row-per-warp variant: used for rows longer than MATVEC_CAP and under FEELGPU_MATVEC_ROW_PER_WARP=1 (A/B timing).
__global__ void __launch_bounds__(256)
k_matvec(int64_t nRows, const int64_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
         const double* __restrict__ x, double* __restrict__ y)
{
    const int lane = threadIdx.x & 31;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < nRows; r += nw) {
        const int64_t p0 = ia[r], p1 = ia[r + 1];
        double s = 0.0;
        for (int64_t p = p0 + lane; p < p1; p += 32) s += a[p] * __ldg(&x[ja[p]]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) y[r] = s;
    }
}

// Streaming variant (FE rows are short: 15 / 29 / 48 nonzeros for P1 / P2 / P3): a warp takes up to 32 consecutive rows
// whose nonzeros fit MATVEC_CAP values, streams that contiguous piece of (a, ja) with all lanes -- coalesced, many loads in
// flight -- leaves the products in shared memory, and lane l sums row l's piece.
constexpr int MATVEC_CAP = 768, MATVEC_WARPS = 4;

__global__ void __launch_bounds__(32 * MATVEC_WARPS, 8)
k_matvec_stream(int64_t nRows, const int64_t* __restrict__ ia, const int32_t* __restrict__ ja,
                const double* __restrict__ a, const double* __restrict__ x, double* __restrict__ y)
{
    __shared__ double s_prod[MATVEC_WARPS][MATVEC_CAP];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    double* sp = s_prod[w];
    const int64_t nw = (int64_t)gridDim.x * MATVEC_WARPS;
    for (int64_t r0 = ((int64_t)blockIdx.x * MATVEC_WARPS + w) * 32; r0 < nRows; r0 += nw * 32) {
        const int64_t rend = min(r0 + 32, nRows);
        int64_t rbeg = r0;
        while (rbeg < rend) {
            const int64_t myr = rbeg + lane;
            const int64_t plo = ia[rbeg];
            const int64_t myend = myr < rend ? ia[myr + 1] : -1;
            const bool fits = myr < rend && myend - plo <= MATVEC_CAP;     // monotone in the lane: a prefix of the warp
            const int R = __popc(__ballot_sync(full, fits));
            if (R == 0) {   // one row longer than the staging buffer
                const int64_t p1 = __shfl_sync(full, myend, 0);
                double s = 0.0;
                for (int64_t p = plo + lane; p < p1; p += 32) s += a[p] * __ldg(&x[ja[p]]);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(full, s, o);
                if (lane == 0) y[rbeg] = s;
                rbeg += 1;
                continue;
            }
            const int cnt = (int)(__shfl_sync(full, myend, R - 1) - plo);
            int64_t mybeg = __shfl_up_sync(full, myend, 1);
            if (lane == 0) mybeg = plo;
#pragma unroll 4
            for (int k = lane; k < cnt; k += 32) sp[k] = a[plo + k] * __ldg(&x[ja[plo + k]]);
            __syncwarp();
            if (lane < R) {
                double s = 0.0;
                for (int k = (int)(mybeg - plo); k < (int)(myend - plo); ++k) s += sp[k];
                y[myr] = s;
            }
            __syncwarp();
            rbeg += R;
        }
    }
}

} // namespace

int feelgpu_launch_matvec(feelgpu_pattern_t pat, const double* a, const double* x, double* y, cudaStream_t s)
{
    feelgpu_context_t ctx = pat->ctx;
    if (pat->test->d_p2c || pat->test != pat->trial) {
        ctx->last_error = "matvec: needs one space with identity process->cluster numbering (a sequential matrix)";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    if (!pat->d_ja32) {
        ctx->last_error = "matvec: the pattern holds column ids beyond 32 bits";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    if (pat->nRows == 0) return FEELGPU_OK;
    static const bool row_per_warp = [] {
        const char* e = std::getenv("FEELGPU_MATVEC_ROW_PER_WARP");
        return e && e[0] == '1';
    }();
    if (row_per_warp) {
        const int threads = 256;
        const int grid = (int)std::min<int64_t>((pat->nRows * 32 + threads - 1) / threads, (int64_t)ctx->sm_count * 32);
        k_matvec<<<grid, threads, 0, s>>>(pat->nRows, pat->d_ia, pat->d_ja32, a, x, y);
    } else {
        const int64_t groups = (pat->nRows + 31) / 32;
        const int grid = (int)std::min<int64_t>((groups + MATVEC_WARPS - 1) / MATVEC_WARPS, (int64_t)ctx->sm_count * 28);
        k_matvec_stream<<<grid, 32 * MATVEC_WARPS, 0, s>>>(pat->nRows, pat->d_ia, pat->d_ja32, a, x, y);
    }
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    return FEELGPU_OK;
}

int feelgpu_launch_zero_rows(feelgpu_pattern_t pat, int n, const int32_t* h_rows, const double* h_x, int strategy,
                             double value_on_diagonal, double* a, double* b, cudaStream_t s)
{
    feelgpu_context_t ctx = pat->ctx;
    if (n == 0) return FEELGPU_OK;
    const bool symmetric = (strategy & FEELGPU_ON_SYMMETRIC) != 0;
    if (symmetric && (pat->test->d_p2c || pat->test != pat->trial)) {
        ctx->last_error = "zero_rows: the symmetric elimination needs one space with identity process->cluster numbering";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    if (!pat->d_ja32) {
        ctx->last_error = "zero_rows: the pattern holds column ids beyond 32 bits";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    // The reference de-duplicates the dofs while it collects them (feelvf/on.hpp: std::find(dofs...) -> continue), so a row
    // listed twice keeps its first value.  The same is done here, on private copies: two warps of k_on_columns working on
    // one row would otherwise both subtract a_cr x_r from b_c.  The copies are pageable, so the asynchronous uploads
    // below have staged them when cudaMemcpyAsync returns, whatever kind of memory the caller passed.
    std::vector<int32_t> rows;
    std::vector<double> xs;
    rows.reserve(n);
    xs.reserve(n);
    {
        std::vector<uint8_t> seen((size_t)pat->nRows, 0);
        for (int k = 0; k < n; ++k) {
            const int32_t r = h_rows[k];
            if (r < 0 || r >= pat->nRows) {
                ctx->last_error = "zero_rows: row id outside [0,nRows)";
                return FEELGPU_ERR_ARG;
            }
            if (seen[r]) continue;
            seen[r] = 1;
            rows.push_back(r);
            xs.push_back(h_x[k]);
        }
    }
    n = (int)rows.size();
    int32_t* d_rows = nullptr;
    double* d_x = nullptr;
    int32_t* d_mark = nullptr;
    // every exit path releases the stream-ordered temporaries
    auto fail = [&](cudaError_t err, const char* what) {
        ctx->last_error = std::string("zero_rows: ") + what + ": " + cudaGetErrorString(err);
        if (d_rows) cudaFreeAsync(d_rows, s);
        if (d_x) cudaFreeAsync(d_x, s);
        if (d_mark) cudaFreeAsync(d_mark, s);
        return FEELGPU_ERR_CUDA;
    };
    cudaError_t err;
    if ((err = cudaMallocAsync(&d_rows, sizeof(int32_t) * n, s)) != cudaSuccess) return fail(err, "cudaMallocAsync");
    if ((err = cudaMallocAsync(&d_x, sizeof(double) * n, s)) != cudaSuccess) return fail(err, "cudaMallocAsync");
    if ((err = cudaMemcpyAsync(d_rows, rows.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, s)) != cudaSuccess ||
        (err = cudaMemcpyAsync(d_x, xs.data(), sizeof(double) * n, cudaMemcpyHostToDevice, s)) != cudaSuccess)
        return fail(err, "cudaMemcpyAsync");
    const int threads = 256, wblocks = (int)(((int64_t)n * 32 + threads - 1) / threads);
    if (symmetric) {
        if ((err = cudaMallocAsync(&d_mark, sizeof(int32_t) * std::max<int64_t>(pat->nRows, 1), s)) != cudaSuccess ||
            (err = cudaMemsetAsync(d_mark, 0, sizeof(int32_t) * pat->nRows, s)) != cudaSuccess)
            return fail(err, "mark array");
        k_on_mark<<<(n + threads - 1) / threads, threads, 0, s>>>(n, d_rows, d_mark);
        k_on_columns<<<wblocks, threads, 0, s>>>(n, d_rows, d_x, d_mark, pat->d_ia, pat->d_ja32, a, b);
        ctx->launches += 2;
    }
    k_on_rows<<<wblocks, threads, 0, s>>>(n, d_rows, d_x, pat->d_ia, pat->d_ja32, pat->test->d_p2c, a, b,
                                          (strategy & FEELGPU_ON_KEEP_DIAGONAL) ? 1 : 0, value_on_diagonal);
    ctx->launches++;
    if ((err = cudaGetLastError()) != cudaSuccess) return fail(err, "kernel launch");
    cudaFreeAsync(d_rows, s);
    cudaFreeAsync(d_x, s);
    if (d_mark) cudaFreeAsync(d_mark, s);
    return FEELGPU_OK;
}
