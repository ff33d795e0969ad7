// kernels_scatter.cu -- element loop + scatter-add through the element->CSR position map (sm_100a).
//
// This is the direct device counterpart of the reference's per-element update/integrate/assemble
// (feelvf/integrator.hpp:1093-1207): one warp per element, the affine geomap K,J,B (feelpoly/geomap.hpp:2320-2355,
// 1971-1991) in registers, lanes spread over the (i,j) entries of the local matrix
// (feelvf/bilinearformcontext.hpp:585-589), quadrature contraction per entry
// (feelpoly/pointsetquadrature.hpp:403-416) and red.global.add.f64 into the CSR values, which replaces
// MatSetValues(ADD_VALUES) (feelalg/matrixpetsc.cpp:917-936).  It is the general path (any coefficient, any
// supported order) and the one used for form1; the owner-computes path of kernels_rowplan.cu is the fast one.
#include "feelgpu_internal.h"
#include "device_common.cuh"
#include <algorithm>
#include <cstring>

namespace {

// ---------------------------------------------------------------------------------------------------------------
// element -> CSR position map: pos[e][i][j] = index of column cluster(l2g_trial[e][j]) in row l2g_test[e][i]
// (the search MatSetValues performs per entry per assembly, done once here)
// ---------------------------------------------------------------------------------------------------------------
__global__ void k_build_pos(int64_t total, int n, int nd_test, int nd_trial, const int32_t* __restrict__ l2g_test,
                            const int32_t* __restrict__ l2g_trial, const int64_t* __restrict__ p2c,
                            const int64_t* __restrict__ ia, const int64_t* __restrict__ ja, int64_t nRows,
                            uint16_t* __restrict__ pos, int* __restrict__ flag)
{
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t e = t / (n * n);
        int r = (int)(t - e * n * n);
        int i = r / n, j = r - i * n;
        int64_t row = l2g_test[e * nd_test + i];
        int64_t col = l2g_trial[e * nd_trial + j];
        if (p2c) col = p2c[col];
        if (row < 0 || row >= nRows) {
            atomicExch(flag, 2);
            pos[t] = 0;
            continue;
        }
        int64_t lo = ia[row], hi = ia[row + 1], b = lo;
        while (lo < hi) {
            int64_t mid = (lo + hi) >> 1;
            if (ja[mid] < col) lo = mid + 1;
            else hi = mid;
        }
        if (lo >= ia[row + 1] || ja[lo] != col || lo - b > 65535) {
            atomicExch(flag, (lo - b > 65535) ? 3 : 1);
            pos[t] = 0;
        } else
            pos[t] = (uint16_t)(lo - b);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// form2, one warp per element
// ---------------------------------------------------------------------------------------------------------------
template <int KIND, bool PRECONTRACTED>
__global__ void __launch_bounds__(256)
k_form2_scatter(const FormParams* __restrict__ P, DevTables T, int64_t nElt, const double* __restrict__ xyz,
                const int4* __restrict__ conn, const int32_t* __restrict__ l2g, int nd,
                const int64_t* __restrict__ ia, const uint16_t* __restrict__ pos,
                const int32_t* __restrict__ l2gf, const double* __restrict__ fvals, double* __restrict__ a)
{
    extern __shared__ double smem[];
    const int n = P->n, nq = P->nq, nf = P->nf, ncomp = P->ncomp;
    // shared tables
    double* s_A = smem;                         // PRECONTRACTED: M9 (n*n*9) or Mm (n*n); else dphi (n*nq*3) or phi
    double* s_phif = nullptr;
    double* s_kq = nullptr;
    int off = 0;
    if (PRECONTRACTED) {
        int cnt = (KIND == FEELGPU_FORM2_MASS) ? n * n : n * n * 9;
        const double* src = (KIND == FEELGPU_FORM2_MASS) ? T.Mm : T.M9;
        for (int k = threadIdx.x; k < cnt; k += blockDim.x) s_A[k] = src[k];
        off = cnt;
    } else {
        int cnt = (KIND == FEELGPU_FORM2_MASS) ? n * nq : n * nq * 3;
        const double* src = (KIND == FEELGPU_FORM2_MASS) ? T.phi : T.dphi;
        for (int k = threadIdx.x; k < cnt; k += blockDim.x) s_A[k] = src[k];
        off = cnt;
        s_phif = smem + off;
        for (int k = threadIdx.x; k < nf * nq; k += blockDim.x) s_phif[k] = T.phif[k];
        off += nf * nq;
        s_kq = smem + off + (threadIdx.x >> 5) * nq;   // per-warp coefficient values
    }
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;

    for (int64_t e = warp; e < nElt; e += nwarps) {
        Geo g;
        geo_update(xyz, __ldg(&conn[e]), g);
        // G = J * B^T B  (G_bd = J sum_a B_ab B_ad)
        double G[3][3];
#pragma unroll
        for (int b = 0; b < 3; ++b)
#pragma unroll
            for (int d = 0; d < 3; ++d)
                G[b][d] = g.J * (g.B[0][b] * g.B[0][d] + g.B[1][b] * g.B[1][d] + g.B[2][b] * g.B[2][d]);
        if (!PRECONTRACTED) {
            __syncwarp();
            for (int q = lane; q < nq; q += 32) s_kq[q] = P->w[q] * coeff_at(P, s_phif, g, e, q, l2gf, fvals);
            __syncwarp();
        }
        const int32_t* rows = l2g + e * nd;
        const uint16_t* pe = pos + e * (int64_t)(n * n);
        for (int idx = lane; idx < n * n; idx += 32) {
            const int i = idx / n, j = idx - i * n;
            const int64_t row = rows[i];
            const int64_t p0 = pe[idx];
            if (KIND == FEELGPU_FORM2_MASS) {
                double s;
                if (PRECONTRACTED) s = g.J * P->cval[0] * s_A[idx];
                else {
                    s = 0;
                    for (int q = 0; q < nq; ++q) s += s_kq[q] * (s_A[i * nq + q] * s_A[j * nq + q]);
                    s *= g.J;
                }
                for (int c = 0; c < ncomp; ++c) atomicAdd(&a[ia[row + c] + p0 + c], s);
            } else if (KIND == FEELGPU_FORM2_STIFFNESS) {
                double s = 0;
                if (PRECONTRACTED) {
                    const double* M = &s_A[idx * 9];
#pragma unroll
                    for (int b = 0; b < 3; ++b)
#pragma unroll
                        for (int d = 0; d < 3; ++d) s += G[b][d] * M[b * 3 + d];
                    s *= P->cval[0];
                } else {
                    for (int q = 0; q < nq; ++q) {
                        const double* gi = &s_A[(i * nq + q) * 3];
                        const double* gj = &s_A[(j * nq + q) * 3];
                        double t0 = G[0][0] * gj[0] + G[0][1] * gj[1] + G[0][2] * gj[2];
                        double t1 = G[1][0] * gj[0] + G[1][1] * gj[1] + G[1][2] * gj[2];
                        double t2 = G[2][0] * gj[0] + G[2][1] * gj[1] + G[2][2] * gj[2];
                        s += s_kq[q] * (gi[0] * t0 + gi[1] * t1 + gi[2] * t2);
                    }
                }
                for (int c = 0; c < ncomp; ++c) atomicAdd(&a[ia[row + c] + p0 + c], s);
            } else { // elasticity, constant Lame coefficients: block (c1,c2) = J [lam P_c1c2 + mu P_c2c1 + mu d tr P]
                const double* M = &s_A[idx * 9];
                double BD[3][3], Pm[3][3];
#pragma unroll
                for (int a1 = 0; a1 < 3; ++a1)
#pragma unroll
                    for (int d = 0; d < 3; ++d)
                        BD[a1][d] = g.B[a1][0] * M[0 * 3 + d] + g.B[a1][1] * M[1 * 3 + d] + g.B[a1][2] * M[2 * 3 + d];
#pragma unroll
                for (int a1 = 0; a1 < 3; ++a1)
#pragma unroll
                    for (int c2 = 0; c2 < 3; ++c2)
                        Pm[a1][c2] = BD[a1][0] * g.B[c2][0] + BD[a1][1] * g.B[c2][1] + BD[a1][2] * g.B[c2][2];
                const double tr = Pm[0][0] + Pm[1][1] + Pm[2][2];
#pragma unroll
                for (int c1 = 0; c1 < 3; ++c1) {
                    const int64_t base = ia[row + c1] + p0;
#pragma unroll
                    for (int c2 = 0; c2 < 3; ++c2) {
                        double v = P->lambda * Pm[c1][c2] + P->mu * Pm[c2][c1];
                        if (c1 == c2) v += P->mu * tr;
                        atomicAdd(&a[base + c2], g.J * v);
                    }
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// form1: one thread per (element, node); b[l2g] += sum_q w_q J f_q phi_i(q)   (VecSetValues ADD, vectorpetsc.cpp:177-190)
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_form1(const FormParams* __restrict__ P, DevTables T, int64_t nElt, const double* __restrict__ xyz,
        const int4* __restrict__ conn, const int32_t* __restrict__ l2g, int nd, const int32_t* __restrict__ l2gf,
        const double* __restrict__ fvals, double* __restrict__ b)
{
    extern __shared__ double smem[];
    const int n = P->n, nq = P->nq, nf = P->nf, ncomp = P->ncomp;
    double* s_phi = smem;
    double* s_phif = smem + n * nq;
    for (int k = threadIdx.x; k < n * nq; k += blockDim.x) s_phi[k] = T.phi[k];
    for (int k = threadIdx.x; k < nf * nq; k += blockDim.x) s_phif[k] = T.phif[k];
    __syncthreads();
    const int64_t total = nElt * n;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t e = t / n;
        const int i = (int)(t - e * n);
        Geo g;
        geo_update(xyz, __ldg(&conn[e]), g);
        if (ncomp == 1) {
            double s = 0;
            for (int q = 0; q < nq; ++q)
                s += (P->w[q] * g.J) * (coeff_at(P, s_phif, g, e, q, l2gf, fvals) * s_phi[i * nq + q]);
            atomicAdd(&b[l2g[e * nd + i]], s);
        } else {
            double s = 0;
            for (int q = 0; q < nq; ++q) s += (P->w[q] * g.J) * s_phi[i * nq + q];
            for (int c = 0; c < ncomp; ++c) atomicAdd(&b[l2g[e * nd + n * c + i]], s * P->cval[c]);
        }
    }
}

// form1 with a point-dependent coefficient (expr("...") at xReal, idv(field)): the coefficient is evaluated once per
// (element, quadrature point) -- thread = (element, q) -- and left in shared memory together with prod[q] = w_q J; then
// thread = (element, node) forms the same sum, in the same order, as k_form1.  A block works on tiles of `tile` elements.
__global__ void __launch_bounds__(256)
k_form1_coeff(const FormParams* __restrict__ P, DevTables T, int64_t nElt, int tile, const double* __restrict__ xyz,
              const int4* __restrict__ conn, const int32_t* __restrict__ l2g, int nd, const int32_t* __restrict__ l2gf,
              const double* __restrict__ fvals, double* __restrict__ b)
{
    extern __shared__ double smem[];
    const int n = P->n, nq = P->nq, nf = P->nf;
    double* s_phi = smem;
    double* s_phif = s_phi + n * nq;
    double* s_prod = s_phif + nf * nq;       // [tile][nq]
    double* s_f = s_prod + tile * nq;        // [tile][nq]
    for (int k = threadIdx.x; k < n * nq; k += blockDim.x) s_phi[k] = T.phi[k];
    for (int k = threadIdx.x; k < nf * nq; k += blockDim.x) s_phif[k] = T.phif[k];
    __syncthreads();
    for (int64_t e0 = (int64_t)blockIdx.x * tile; e0 < nElt; e0 += (int64_t)gridDim.x * tile) {
        const int ne = (int)min((int64_t)tile, nElt - e0);
        for (int t = threadIdx.x; t < ne * nq; t += blockDim.x) {
            const int el = t / nq, q = t - el * nq;
            Geo g;
            geo_update(xyz, __ldg(&conn[e0 + el]), g);
            s_prod[t] = P->w[q] * g.J;
            s_f[t] = coeff_at(P, s_phif, g, e0 + el, q, l2gf, fvals);
        }
        __syncthreads();
        for (int t = threadIdx.x; t < ne * n; t += blockDim.x) {
            const int el = t / n, i = t - el * n;
            double s = 0;
            for (int q = 0; q < nq; ++q) s += s_prod[el * nq + q] * (s_f[el * nq + q] * s_phi[i * nq + q]);
            atomicAdd(&b[l2g[(e0 + el) * nd + i]], s);
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------------------
// boundary faces (feelvf/integrator.hpp:3242-3704): the points q of block f = local face of the tables lie on that face of
// the reference tet; prod[q] = w[q] * J * |B n_f| (feelpoly/pointsetquadrature.hpp:481-490), n_f the reference normal:
// (1,1,1)/sqrt(3) for face 0, -e_{f-1} otherwise (feelmesh/refsimplex.hpp:1045-1083)
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double face_measure(const Geo& g, int f)
{
    double N[3];
    if (f == 0) {
        const double r = sqrt(1.0 / 3.0);
#pragma unroll
        for (int a = 0; a < 3; ++a) N[a] = g.B[a][0] * r + g.B[a][1] * r + g.B[a][2] * r;
    } else {
#pragma unroll
        for (int a = 0; a < 3; ++a) N[a] = -(f == 1 ? g.B[a][0] : f == 2 ? g.B[a][1] : g.B[a][2]);
    }
    return g.J * sqrt(N[0] * N[0] + N[1] * N[1] + N[2] * N[2]);
}

// one thread per (face, local dof): b[l2g] += sum_q prod[q] g(x_q) phi_i(q)
__global__ void __launch_bounds__(256)
k_form1_faces(const FormParams* __restrict__ P, DevTables T, int64_t nFaces, const int32_t* __restrict__ felt,
              const int32_t* __restrict__ flocal, const double* __restrict__ xyz, const int4* __restrict__ conn,
              const int32_t* __restrict__ l2g, int nd, double* __restrict__ b)
{
    extern __shared__ double smem[];
    const int n = P->n, nq = P->nq, nqf = nq / 4, ncomp = P->ncomp;
    double* s_phi = smem;
    for (int k = threadIdx.x; k < n * nq; k += blockDim.x) s_phi[k] = T.phi[k];
    __syncthreads();
    const int64_t total = nFaces * n;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = t / n;
        const int i = (int)(t - k * n);
        const int64_t e = felt[k];
        const int f = flocal[k];
        Geo g;
        geo_update(xyz, __ldg(&conn[e]), g);
        const double mf = face_measure(g, f);
        if (ncomp == 1) {
            double s = 0;
            for (int q = f * nqf; q < (f + 1) * nqf; ++q)
                s += (P->w[q] * mf) * (coeff_at(P, nullptr, g, e, q, nullptr, nullptr) * s_phi[i * nq + q]);
            atomicAdd(&b[l2g[e * nd + i]], s);
        } else {
            for (int c = 0; c < ncomp; ++c) {
                double s = 0;
                for (int q = f * nqf; q < (f + 1) * nqf; ++q) s += (P->w[q] * mf) * (P->cval[c] * s_phi[i * nq + q]);
                atomicAdd(&b[l2g[e * nd + n * c + i]], s);
            }
        }
    }
}

// one thread per (face, i, j): a[row_i, col_j] += sum_q prod[q] k(x_q) phi_j(q) phi_i(q)
__global__ void __launch_bounds__(256)
k_form2_faces_mass(const FormParams* __restrict__ P, DevTables T, int64_t nFaces, const int32_t* __restrict__ felt,
                   const int32_t* __restrict__ flocal, const double* __restrict__ xyz, const int4* __restrict__ conn,
                   const int32_t* __restrict__ l2g, int nd, const int64_t* __restrict__ ia,
                   const uint16_t* __restrict__ pos, double* __restrict__ a)
{
    extern __shared__ double smem[];
    const int n = P->n, nq = P->nq, nqf = nq / 4, ncomp = P->ncomp;
    double* s_phi = smem;
    for (int k = threadIdx.x; k < n * nq; k += blockDim.x) s_phi[k] = T.phi[k];
    __syncthreads();
    const int64_t total = nFaces * n * n;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = t / (n * n);
        const int idx = (int)(t - k * n * n);
        const int i = idx / n, j = idx - i * n;
        const int64_t e = felt[k];
        const int f = flocal[k];
        Geo g;
        geo_update(xyz, __ldg(&conn[e]), g);
        const double mf = face_measure(g, f);
        double s = 0;
        for (int q = f * nqf; q < (f + 1) * nqf; ++q)
            s += (P->w[q] * mf) * (coeff_at(P, nullptr, g, e, q, nullptr, nullptr) * (s_phi[j * nq + q] * s_phi[i * nq + q]));
        const int64_t row = l2g[e * nd + i];
        const int64_t p0 = pos[e * (int64_t)(n * n) + idx];
        for (int c = 0; c < ncomp; ++c) atomicAdd(&a[ia[row + c] + p0 + c], s);
    }
}

__global__ void k_scatter_add(int64_t n, const int64_t* __restrict__ slots, const double* __restrict__ vals,
                              double* __restrict__ a)
{
    // every slot appears at most once per neighbour message, but messages of different neighbours may overlap
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(&a[slots[k]], vals[k]);
}

} // namespace

extern "C" int feelgpu_scatter_add(feelgpu_context_t ctx, int64_t n, const int64_t* slots, const double* vals, double* a,
                                   void* stream)
{
    if (!ctx) return FEELGPU_ERR_ARG;
    if (n < 0 || (n > 0 && (!slots || !vals || !a))) {
        ctx->last_error = "scatter_add: bad argument";
        return FEELGPU_ERR_ARG;
    }
    if (n == 0) return FEELGPU_OK;
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
    k_scatter_add<<<grid, 256, 0, s>>>(n, slots, vals, a);
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    return FEELGPU_OK;
}

// 32-bit copy of the column ids (matvec, zero_rows, the exchange's slot search), narrowed on the device from the 64-bit array
// that is there anyway for the position map; flag = 1 when an id does not fit
__global__ void k_narrow_columns(int64_t nnz, const int64_t* __restrict__ ja, int32_t* __restrict__ ja32, int* __restrict__ flag)
{
    bool bad = false;
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < nnz; k += (int64_t)gridDim.x * blockDim.x) {
        const int64_t c = ja[k];
        bad |= c < 0 || c >= ((int64_t)1 << 31);
        ja32[k] = (int32_t)c;
    }
    if (bad) atomicOr(flag, 1);
}

// leaves pat->d_ja32 null (not an error) when the ids do not fit 32 bits
int feelgpu_launch_narrow_columns(feelgpu_pattern_t pat, const int64_t* d_ja, cudaStream_t s)
{
    feelgpu_context_t ctx = pat->ctx;
    FEELGPU_CHECK_CUDA(ctx, cudaMalloc(&pat->d_ja32, sizeof(int32_t) * std::max<int64_t>(pat->nnz, 1)));
    FEELGPU_CHECK_CUDA(ctx, cudaMemsetAsync(ctx->d_flag, 0, sizeof(int), s));
    if (pat->nnz > 0) {
        const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((pat->nnz + 255) / 256, (int64_t)ctx->sm_count * 16));
        k_narrow_columns<<<grid, 256, 0, s>>>(pat->nnz, d_ja, pat->d_ja32, ctx->d_flag);
        ctx->launches++;
        FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    }
    int flag = 0;
    FEELGPU_CHECK_CUDA(ctx, cudaMemcpyAsync(&flag, ctx->d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
    FEELGPU_CHECK_CUDA(ctx, cudaStreamSynchronize(s));
    if (flag) {
        cudaFree(pat->d_ja32);
        pat->d_ja32 = nullptr;
    }
    return FEELGPU_OK;
}

int feelgpu_launch_build_pos(feelgpu_pattern_t pat, const int64_t* d_ja, cudaStream_t s)
{
    feelgpu_context_t ctx = pat->ctx;
    const int n = pat->test->n;
    const int64_t total = pat->test->mesh->nElt * n * n;
    FEELGPU_CHECK_CUDA(ctx, cudaMemsetAsync(ctx->d_flag, 0, sizeof(int), s));
    int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)ctx->sm_count * 32);
    if (grid < 1) grid = 1;
    k_build_pos<<<grid, 256, 0, s>>>(total, n, pat->test->nd, pat->trial->nd, pat->test->d_l2g, pat->trial->d_l2g,
                                     pat->trial->d_p2c, pat->d_ia, d_ja, pat->nRows, pat->d_pos, ctx->d_flag);
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    int flag = 0;
    FEELGPU_CHECK_CUDA(ctx, cudaMemcpyAsync(&flag, ctx->d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
    FEELGPU_CHECK_CUDA(ctx, cudaStreamSynchronize(s));
    if (flag == 1) {
        ctx->last_error = "pattern_create: an element (row,col) pair is not in the CSR graph";
        return FEELGPU_ERR_NOT_IN_PATTERN;
    }
    if (flag == 2) {
        ctx->last_error = "pattern_create: a test dof id is outside [0,nRows)";
        return FEELGPU_ERR_ARG;
    }
    if (flag == 3) {
        ctx->last_error = "pattern_create: a CSR row is longer than 65535 entries";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    return FEELGPU_OK;
}

// uploads the per-call parameter block + tables into ctx->d_scratch; returns device pointers
static int upload_tables(feelgpu_context_t ctx, const FormParams& hp, const feelgpu_tables& t, cudaStream_t s,
                         const FormParams** dP, DevTables* dT)
{
    const int n = t.n, nq = t.nq, nf = t.nf;
    size_t cnt = (size_t)n * nq + (size_t)n * nq * 3 + (size_t)nf * nq + (size_t)n * n * 9 + (size_t)n * n;
    size_t bytes = sizeof(FormParams) + cnt * sizeof(double);
    static thread_local std::vector<char> host;
    host.resize(bytes);
    std::memcpy(host.data(), &hp, sizeof(FormParams));
    double* p = reinterpret_cast<double*>(host.data() + sizeof(FormParams));
    double* phi = p;
    double* dphi = phi + (size_t)n * nq;
    double* phif = dphi + (size_t)n * nq * 3;
    double* M9 = phif + (size_t)nf * nq;
    double* Mm = M9 + (size_t)n * n * 9;
    std::memcpy(phi, t.phi, sizeof(double) * n * nq);
    if (t.dphi) std::memcpy(dphi, t.dphi, sizeof(double) * n * nq * 3);
    else std::memset(dphi, 0, sizeof(double) * n * nq * 3);
    if (nf) std::memcpy(phif, t.phif, sizeof(double) * nf * nq);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            double m = 0;
            for (int q = 0; q < nq; ++q) m += t.w[q] * (phi[i * nq + q] * phi[j * nq + q]);
            Mm[i * n + j] = m;
            for (int b = 0; b < 3; ++b)
                for (int d = 0; d < 3; ++d) {
                    double v = 0;
                    for (int q = 0; q < nq; ++q) v += t.w[q] * (dphi[(i * nq + q) * 3 + b] * dphi[(j * nq + q) * 3 + d]);
                    M9[(i * n + j) * 9 + b * 3 + d] = v;
                }
        }
    // device-side cache keyed by content: repeated assemblies of the same form issue no host->device copy
    (void)s;
    void* dev = nullptr;
    for (auto& ent : ctx->table_cache)
        if (ent.host.size() == bytes && std::memcmp(ent.host.data(), host.data(), bytes) == 0) dev = ent.dev;
    if (!dev) {
        FEELGPU_CHECK_CUDA(ctx, cudaMalloc(&dev, bytes));
        FEELGPU_CHECK_CUDA(ctx, cudaMemcpy(dev, host.data(), bytes, cudaMemcpyHostToDevice));
        if (ctx->table_cache.size() >= 64) {
            // entries may still be referenced by in-flight kernels: drain before recycling
            cudaDeviceSynchronize();
            for (auto& ent : ctx->table_cache) cudaFree(ent.dev);
            ctx->table_cache.clear();
        }
        ctx->table_cache.push_back({host, dev});
    }
    char* d = reinterpret_cast<char*>(dev);
    *dP = reinterpret_cast<const FormParams*>(d);
    const double* dp = reinterpret_cast<const double*>(d + sizeof(FormParams));
    dT->phi = dp;
    dT->dphi = dp + (size_t)n * nq;
    dT->phif = dT->dphi + (size_t)n * nq * 3;
    dT->M9 = dT->phif + (size_t)nf * nq;
    dT->Mm = dT->M9 + (size_t)n * n * 9;
    return FEELGPU_OK;
}

int feelgpu_upload_tables_public(feelgpu_context_t ctx, const FormParams& hp, const feelgpu_tables& t,
                                 cudaStream_t s, const FormParams** dP, const double** phi, const double** dphi,
                                 const double** phif, const double** M9, const double** Mm)
{
    DevTables T;
    int rc = upload_tables(ctx, hp, t, s, dP, &T);
    if (rc) return rc;
    *phi = T.phi; *dphi = T.dphi; *phif = T.phif; *M9 = T.M9; *Mm = T.Mm;
    return FEELGPU_OK;
}

int feelgpu_launch_form2_scatter(feelgpu_pattern_t pat, const FormParams& hp, const feelgpu_form2_desc* desc,
                                 double* a, cudaStream_t s)
{
    feelgpu_context_t ctx = pat->ctx;
    const FormParams* dP;
    DevTables T;
    int rc = upload_tables(ctx, hp, desc->tables, s, &dP, &T);
    if (rc) return rc;
    feelgpu_space_t sp = pat->test;
    feelgpu_mesh_t m = sp->mesh;
    const int n = hp.n, nq = hp.nq, nf = hp.nf;
    const bool pre = hp.coeff_kind == FEELGPU_COEFF_CONST || hp.kind == FEELGPU_FORM2_ELASTICITY;
    const int threads = 256;
    size_t sh;
    if (pre) sh = sizeof(double) * (hp.kind == FEELGPU_FORM2_MASS ? n * n : n * n * 9);
    else sh = sizeof(double) * ((hp.kind == FEELGPU_FORM2_MASS ? n * nq : n * nq * 3) + nf * nq + (threads / 32) * nq);
    const int32_t* l2gf = nullptr;
    const double* fvals = nullptr;
    if (hp.coeff_kind == FEELGPU_COEFF_FIELD) {
        l2gf = desc->coeff.field_space->d_l2g;
        fvals = desc->coeff.field_values;
    }
    int64_t warps_needed = m->nElt;
    int grid = (int)std::min<int64_t>((warps_needed * 32 + threads - 1) / threads, (int64_t)ctx->sm_count * 64);
    if (grid < 1) grid = 1;
#define LAUNCH(K, PRE)                                                                                           \
    do {                                                                                                         \
        if (sh > 48 * 1024)                                                                                      \
            cudaFuncSetAttribute(k_form2_scatter<K, PRE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh); \
        k_form2_scatter<K, PRE><<<grid, threads, sh, s>>>(dP, T, m->nElt, m->d_xyz, m->d_conn, sp->d_l2g, sp->nd, \
                                                          pat->d_ia, pat->d_pos, l2gf, fvals, a);                \
    } while (0)
    switch (hp.kind) {
    case FEELGPU_FORM2_MASS:
        if (pre) LAUNCH(FEELGPU_FORM2_MASS, true); else LAUNCH(FEELGPU_FORM2_MASS, false);
        break;
    case FEELGPU_FORM2_STIFFNESS:
        if (pre) LAUNCH(FEELGPU_FORM2_STIFFNESS, true); else LAUNCH(FEELGPU_FORM2_STIFFNESS, false);
        break;
    case FEELGPU_FORM2_ELASTICITY:
        LAUNCH(FEELGPU_FORM2_ELASTICITY, true);
        break;
    default:
        ctx->last_error = "form2: unknown kind";
        return FEELGPU_ERR_UNSUPPORTED;
    }
#undef LAUNCH
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    return FEELGPU_OK;
}

int feelgpu_launch_form1_faces(feelgpu_space_t sp, feelgpu_faces_t fc, const FormParams& hp, const feelgpu_form1_desc* desc,
                               double* b, cudaStream_t s)
{
    feelgpu_context_t ctx = sp->ctx;
    const FormParams* dP;
    DevTables T;
    int rc = upload_tables(ctx, hp, desc->tables, s, &dP, &T);
    if (rc) return rc;
    feelgpu_mesh_t m = sp->mesh;
    const size_t sh = sizeof(double) * (hp.n * hp.nq);
    const int64_t total = fc->n * hp.n;
    int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)ctx->sm_count * 16);
    k_form1_faces<<<std::max(grid, 1), 256, sh, s>>>(dP, T, fc->n, fc->d_elt, fc->d_lface, m->d_xyz, m->d_conn, sp->d_l2g,
                                                     sp->nd, b);
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    return FEELGPU_OK;
}

int feelgpu_launch_form2_faces(feelgpu_pattern_t pat, feelgpu_faces_t fc, const FormParams& hp,
                               const feelgpu_form2_desc* desc, double* a, cudaStream_t s)
{
    feelgpu_context_t ctx = pat->ctx;
    const FormParams* dP;
    DevTables T;
    int rc = upload_tables(ctx, hp, desc->tables, s, &dP, &T);
    if (rc) return rc;
    feelgpu_space_t sp = pat->test;
    feelgpu_mesh_t m = sp->mesh;
    const size_t sh = sizeof(double) * (hp.n * hp.nq);
    const int64_t total = fc->n * hp.n * hp.n;
    int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)ctx->sm_count * 16);
    if (sh > 48 * 1024)
        FEELGPU_CHECK_CUDA(ctx, cudaFuncSetAttribute(k_form2_faces_mass, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
    k_form2_faces_mass<<<std::max(grid, 1), 256, sh, s>>>(dP, T, fc->n, fc->d_elt, fc->d_lface, m->d_xyz, m->d_conn,
                                                          sp->d_l2g, sp->nd, pat->d_ia, pat->d_pos, a);
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    return FEELGPU_OK;
}

int feelgpu_launch_form1(feelgpu_space_t sp, const FormParams& hp, const feelgpu_form1_desc* desc, double* b,
                         cudaStream_t s)
{
    feelgpu_context_t ctx = sp->ctx;
    const FormParams* dP;
    DevTables T;
    int rc = upload_tables(ctx, hp, desc->tables, s, &dP, &T);
    if (rc) return rc;
    feelgpu_mesh_t m = sp->mesh;
    const int32_t* l2gf = nullptr;
    const double* fvals = nullptr;
    if (hp.coeff_kind == FEELGPU_COEFF_FIELD) {
        l2gf = desc->coeff.field_space->d_l2g;
        fvals = desc->coeff.field_values;
    }
    size_t sh = sizeof(double) * (hp.n * hp.nq + hp.nf * hp.nq);
    if (hp.coeff_kind != FEELGPU_COEFF_CONST && hp.ncomp == 1) {
        // tiles sized to the default 48 KB of dynamic shared memory (tables <= 20 KB at nq = 64)
        const int tile = (int)std::min<size_t>(64, (48 * 1024 - sh) / (sizeof(double) * 2 * hp.nq));
        if (tile >= 1) {
            const size_t sht = sh + sizeof(double) * 2 * tile * hp.nq;
            const int g2 = (int)std::max<int64_t>(1, std::min<int64_t>((m->nElt + tile - 1) / tile, (int64_t)ctx->sm_count * 16));
            k_form1_coeff<<<g2, 256, sht, s>>>(dP, T, m->nElt, tile, m->d_xyz, m->d_conn, sp->d_l2g, sp->nd, l2gf, fvals, b);
            ctx->launches++;
            FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
            return FEELGPU_OK;
        }
    }
    int64_t total = m->nElt * hp.n;
    int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)ctx->sm_count * 64);
    if (grid < 1) grid = 1;
    k_form1<<<grid, 256, sh, s>>>(dP, T, m->nElt, m->d_xyz, m->d_conn, sp->d_l2g, sp->nd, l2gf, fvals, b);
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    return FEELGPU_OK;
}
