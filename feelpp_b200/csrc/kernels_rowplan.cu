// kernels_rowplan.cu -- owner-computes assembly kernel (see rowplan.h for the design) and the plan's device mirror.
//
// Replaces, for constant-coefficient mass and stiffness forms, the reference's per-element update/integrate/assemble
// + MatSetValues(ADD_VALUES) (feelvf/integrator.hpp:1093-1207, feelalg/matrixpetsc.cpp:917-936) by a gather: every
// CSR value is produced by one thread and stored once, in a fixed summation order.
#include <string>
#include "feelgpu_internal.h"
#include "rowplan.h"

#include <algorithm>
#include <cstring>
#include <thread>

namespace {

struct RowTables {
    double T[3][20][6];   // per kind, per rotated-local column j': stiffness coefficients of {S11,S22,S33,S12,S13,S23},
                          // or T[kind][j'][0] = mass row entry
    int canon[3];
    int pad;
};

__device__ __forceinline__ int pair_idx(int x, int y) { return x + y - 1 + (min(x, y) >= 1 ? 1 : 0); }

template <int ORDER, int FORM>   // FORM 0: mass, 1: stiffness
__global__ void __launch_bounds__(ROWPLAN_WARPS * 32)
k_rowplan(const RowPlanCta* __restrict__ ctas, const RowPlanBundle* __restrict__ bundles,
          const int32_t* __restrict__ visits, const uint32_t* __restrict__ units, const uint16_t* __restrict__ srcs,
          const uint16_t* __restrict__ omap, const int32_t* __restrict__ rows, const double* __restrict__ xyz,
          const int4* __restrict__ conn, const int64_t* __restrict__ ia, double* __restrict__ a, int accumulate,
          int ncomp, int maxchunks, int maxrounds, double cval, const __grid_constant__ RowTables tab)
{
    constexpr int N = (ORDER + 1) * (ORDER + 2) * (ORDER + 3) / 6;
    constexpr int NS = FORM == 1 ? 6 : 1;
    extern __shared__ double sm[];
    double* sS = sm;                                    // [NS][ROWPLAN_VMAX]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int stageSize = maxchunks * N * 32 + 2;       // + zero cell (kept even)
    const int outSize = maxrounds * 32 + 32;            // task results + 32 diagonal slots
    double* stage = sm + NS * ROWPLAN_VMAX + warp * (stageSize + outSize);
    double* out = stage + stageSize;
    const int ZERO = maxchunks * N * 32;
    const int DIAG = maxrounds * 32;
    const RowPlanCta cta = ctas[blockIdx.x];

    // ---- phase A: geometry of every element touched by this CTA's rows, once ----
    for (int v = threadIdx.x; v < (int)cta.nvisit; v += blockDim.x) {
        const int4 c = __ldg(&conn[visits[cta.visit_off + v]]);
        double x[4][3];
        {
            const int id[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const double2* q = reinterpret_cast<const double2*>(xyz + 4 * (int64_t)id[k]);
                const double2 p0 = __ldg(q), p1 = __ldg(q + 1);
                x[k][0] = p0.x;
                x[k][1] = p0.y;
                x[k][2] = p1.x;
            }
        }
        // K = G grad(N) (edge vectors / 2), det, J = |det| (feelpoly/geomap.hpp:2320-2355, 1971-1976)
        double K[3][3];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            K[r][0] = 0.5 * (x[1][r] - x[0][r]);
            K[r][1] = 0.5 * (x[2][r] - x[0][r]);
            K[r][2] = 0.5 * (x[3][r] - x[0][r]);
        }
        const double c00 = K[1][1] * K[2][2] - K[1][2] * K[2][1];
        const double c01 = K[1][2] * K[2][0] - K[1][0] * K[2][2];
        const double c02 = K[1][0] * K[2][1] - K[1][1] * K[2][0];
        const double det = K[0][0] * c00 + K[0][1] * c01 + K[0][2] * c02;
        const double J = fabs(det);
        if (FORM == 0) {
            sS[v] = cval * J;
        } else {
            const double id = 1.0 / det;
            // B = K^{-T} (geomap.hpp:1985-1991): B[r][b] = d xi_b / d x_r ; grad lambda_{b+1} = B[:,b]/2
            double h[4][3];
            h[1][0] = 0.5 * c00 * id;
            h[2][0] = 0.5 * c01 * id;
            h[3][0] = 0.5 * c02 * id;
            h[1][1] = 0.5 * (K[0][2] * K[2][1] - K[0][1] * K[2][2]) * id;
            h[2][1] = 0.5 * (K[0][0] * K[2][2] - K[0][2] * K[2][0]) * id;
            h[3][1] = 0.5 * (K[0][1] * K[2][0] - K[0][0] * K[2][1]) * id;
            h[1][2] = 0.5 * (K[0][1] * K[1][2] - K[0][2] * K[1][1]) * id;
            h[2][2] = 0.5 * (K[0][2] * K[1][0] - K[0][0] * K[1][2]) * id;
            h[3][2] = 0.5 * (K[0][0] * K[1][1] - K[0][1] * K[1][0]) * id;
#pragma unroll
            for (int r = 0; r < 3; ++r) h[0][r] = -(h[1][r] + h[2][r] + h[3][r]);
            const double cj = cval * J;
            // S_kl = c J grad(l_k).grad(l_l), k<l, pair order (0,1),(0,2),(0,3),(1,2),(1,3),(2,3)
            sS[0 * ROWPLAN_VMAX + v] = cj * (h[0][0] * h[1][0] + h[0][1] * h[1][1] + h[0][2] * h[1][2]);
            sS[1 * ROWPLAN_VMAX + v] = cj * (h[0][0] * h[2][0] + h[0][1] * h[2][1] + h[0][2] * h[2][2]);
            sS[2 * ROWPLAN_VMAX + v] = cj * (h[0][0] * h[3][0] + h[0][1] * h[3][1] + h[0][2] * h[3][2]);
            sS[3 * ROWPLAN_VMAX + v] = cj * (h[1][0] * h[2][0] + h[1][1] * h[2][1] + h[1][2] * h[2][2]);
            sS[4 * ROWPLAN_VMAX + v] = cj * (h[1][0] * h[3][0] + h[1][1] * h[3][1] + h[1][2] * h[3][2]);
            sS[5 * ROWPLAN_VMAX + v] = cj * (h[2][0] * h[3][0] + h[2][1] * h[3][1] + h[2][2] * h[3][2]);
        }
    }
    if (lane == 0) stage[ZERO] = 0.0;
    __syncthreads();

    for (uint32_t b = warp; b < cta.nbundle; b += ROWPLAN_WARPS) {
        const RowPlanBundle bd = bundles[cta.bundle_off + b];
        const int kind = bd.kind;
        const int canon = tab.canon[kind];
        out[DIAG + lane] = 0.0;
        __syncwarp();
        // ---- phase 1: lane = unit; its row of the (re-labelled) local matrix goes to stage[entry][lane] ----
        for (int ch = 0; ch < bd.nchunks; ++ch) {
            const uint32_t u = __ldg(&units[(size_t)(bd.unit_off + ch) * 32 + lane]);
            const bool active = u != ROWPLAN_UNIT_NONE;
            const int slot = (int)(u >> 24);
            double dval = 0.0;
            if (active) {
                const int vis = u & 0xFFFF;
                double* st = stage + (size_t)ch * N * 32 + lane;
                if (FORM == 0) {
                    const double cj = sS[vis];
#pragma unroll
                    for (int j = 0; j < N; ++j) {
                        const double v = cj * tab.T[kind][j][0];
                        st[j * 32] = v;
                        if (j == canon) dval = v;
                    }
                } else {
                    const int p0 = (u >> 16) & 3, p1 = (u >> 18) & 3, p2 = (u >> 20) & 3, p3 = (u >> 22) & 3;
                    const double s01 = sS[pair_idx(p0, p1) * ROWPLAN_VMAX + vis];
                    const double s02 = sS[pair_idx(p0, p2) * ROWPLAN_VMAX + vis];
                    const double s03 = sS[pair_idx(p0, p3) * ROWPLAN_VMAX + vis];
                    const double s12 = sS[pair_idx(p1, p2) * ROWPLAN_VMAX + vis];
                    const double s13 = sS[pair_idx(p1, p3) * ROWPLAN_VMAX + vis];
                    const double s23 = sS[pair_idx(p2, p3) * ROWPLAN_VMAX + vis];
                    // rows of the Gram matrix sum to zero: S_kk = -sum_{l != k} S_kl
                    const double s11 = -(s01 + s12 + s13), s22 = -(s02 + s12 + s23), s33 = -(s03 + s13 + s23);
#pragma unroll
                    for (int j = 0; j < N; ++j) {
                        const double* t = tab.T[kind][j];
                        double v = s11 * t[0];
                        v = fma(s22, t[1], v);
                        v = fma(s33, t[2], v);
                        v = fma(s12, t[3], v);
                        v = fma(s13, t[4], v);
                        v = fma(s23, t[5], v);
                        st[j * 32] = v;
                        if (j == canon) dval = v;
                    }
                }
            }
            // diagonal: segmented sum over the lanes of each row (contiguous lanes share a slot)
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const double v = __shfl_down_sync(0xffffffffu, dval, off);
                const int s2 = __shfl_down_sync(0xffffffffu, slot, off);
                if (lane + off < 32 && s2 == slot) dval += v;
            }
            const int sprev = __shfl_up_sync(0xffffffffu, slot, 1);
            if (active && (lane == 0 || sprev != slot)) out[DIAG + slot] += dval;
            __syncwarp();
        }
        // ---- phase 2: lane = a balanced list of output nonzeros, summed from the staging tile in plan order ----
        {
            const uint16_t* sp = srcs + (size_t)bd.src_off * 32 + lane;
            double acc = 0.0;
            int t = 0;
            uint32_t e = bd.niter ? __ldg(sp) : 0;
            for (int it = 0; it < bd.niter; ++it) {
                const uint32_t en = (it + 1 < bd.niter) ? __ldg(sp + (size_t)(it + 1) * 32) : 0;
                int sl = e & ROWPLAN_SRC_MASK;
                sl = sl == ROWPLAN_SRC_MASK ? ZERO : sl;
                acc += stage[sl];
                if (e & ROWPLAN_SRC_END) {
                    out[t * 32 + lane] = acc;
                    acc = 0.0;
                    ++t;
                }
                e = en;
            }
        }
        __syncwarp();
        // ---- phase 3: coalesced store of the rows' value segments ----
        {
            int64_t mybase = 0;
            int mylen = 0;
            int32_t myrow = 0;
            if (lane < bd.nrows) {
                myrow = __ldg(&rows[bd.row_off + lane]);
                mybase = __ldg(&ia[(int64_t)myrow * ncomp]);
                mylen = (int)((__ldg(&ia[(int64_t)myrow * ncomp + 1]) - mybase) / ncomp);
            }
            int off = 0;
            for (int k = 0; k < bd.nrows; ++k) {
                const int64_t base = __shfl_sync(0xffffffffu, mybase, k);
                const int len = __shfl_sync(0xffffffffu, mylen, k);
                const int32_t row = __shfl_sync(0xffffffffu, myrow, k);
                const uint16_t* om = omap + (size_t)bd.omap_off + off;
                if (ncomp == 1) {
                    for (int p = lane; p < len; p += 32) {
                        const uint32_t m = __ldg(om + p);
                        const double v = m == 0xFFFFu ? 0.0 : out[(m & 0x8000u) ? DIAG + (m & 0x7FFFu) : m];
                        if (accumulate) a[base + p] += v;
                        else a[base + p] = v;
                    }
                } else {
                    // product space, block = v * Id: row c1 holds v at column 3p + c1 (zeros elsewhere)
                    for (int c1 = 0; c1 < ncomp; ++c1) {
                        const int64_t b1 = c1 == 0 ? base : __ldg(&ia[(int64_t)row * ncomp + c1]);
                        for (int q = lane; q < len * ncomp; q += 32) {
                            const int p = q / ncomp, c2 = q - p * ncomp;
                            double v = 0.0;
                            if (c2 == c1) {
                                const uint32_t m = __ldg(om + p);
                                v = m == 0xFFFFu ? 0.0 : out[(m & 0x8000u) ? DIAG + (m & 0x7FFFu) : m];
                            }
                            if (!accumulate) a[b1 + q] = v;
                            else if (c2 == c1) a[b1 + q] += v;
                        }
                    }
                }
                off += len;
            }
        }
        __syncwarp();
    }
}

size_t rowplan_smem_bytes(const RowPlan* p, int form)
{
    const int NS = form == 1 ? 6 : 1;
    const size_t stageSize = (size_t)p->maxchunks * p->n * 32 + 2, outSize = (size_t)p->maxrounds * 32 + 32;
    return sizeof(double) * (NS * ROWPLAN_VMAX + ROWPLAN_WARPS * (stageSize + outSize));
}

template <typename T>
cudaError_t upload(T** d, const std::vector<T>& h, int64_t& bytes)
{
    cudaError_t err = cudaMalloc(d, std::max<size_t>(h.size(), 1) * sizeof(T));
    if (err != cudaSuccess) return err;
    bytes += (int64_t)(h.size() * sizeof(T));
    return cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
}

} // namespace

int feelgpu_rowplan_build(feelgpu_pattern_t pat, const int64_t* h_ia, const int64_t* h_ja)
{
    (void)h_ja;
    feelgpu_context_t ctx = pat->ctx;
    feelgpu_space_t sp = pat->test;
    feelgpu_mesh_t m = sp->mesh;
    if (pat->test != pat->trial) return FEELGPU_OK;   // the plan needs test == trial
    const int n = sp->n;
    std::vector<uint16_t> pos((size_t)m->nElt * n * n);
    FEELGPU_CHECK_CUDA(ctx, cudaMemcpy(pos.data(), pat->d_pos, pos.size() * sizeof(uint16_t), cudaMemcpyDeviceToHost));
    RowPlanHost hp;
    if (!rowplan_build_host(sp->order, sp->ncomp, m->nElt, sp->h_l2g.data(), sp->nd, m->nPts, m->h_xyz.data(),
                            m->h_conn.data(), pat->nRows, h_ia, pos.data(), 0, hp)) {
        pat->plan_note = hp.error;   // not an error: the scatter path stays available
        return FEELGPU_OK;
    }
    RowPlan* p = new RowPlan;
    p->order = hp.order;
    p->n = hp.n;
    p->ncomp = hp.ncomp;
    p->maxchunks = hp.maxchunks;
    p->maxrounds = hp.maxrounds;
    p->ncta = (int64_t)hp.ctas.size();
    p->nunits = hp.nunits;
    p->ncontrib = hp.ncontrib;
    p->nsrc_padded = hp.nsrc_padded;
    p->nvisits = hp.nvisits;
    cudaError_t err = upload(&p->d_ctas, hp.ctas, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_bundles, hp.bundles, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_visits, hp.visits, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_units, hp.units, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_srcs, hp.srcs, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_omap, hp.omap, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_rows, hp.rows, p->bytes);
    pat->plan = p;
    if (err != cudaSuccess) {
        ctx->last_error = std::string("rowplan upload: ") + cudaGetErrorString(err);
        feelgpu_rowplan_destroy(pat);
        return FEELGPU_ERR_CUDA;
    }
    if (rowplan_smem_bytes(p, 1) > 220 * 1024) {
        pat->plan_note = "row plan: shared-memory tile too large for this mesh";
        feelgpu_rowplan_destroy(pat);
    }
    return FEELGPU_OK;
}

void feelgpu_rowplan_destroy(feelgpu_pattern_t pat)
{
    RowPlan* p = pat->plan;
    if (!p) return;
    cudaFree(p->d_ctas);
    cudaFree(p->d_bundles);
    cudaFree(p->d_visits);
    cudaFree(p->d_units);
    cudaFree(p->d_srcs);
    cudaFree(p->d_omap);
    cudaFree(p->d_rows);
    delete p;
    pat->plan = nullptr;
}

bool feelgpu_rowplan_supports(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc)
{
    if (!pat->plan) return false;
    if (desc->kind != FEELGPU_FORM2_MASS && desc->kind != FEELGPU_FORM2_STIFFNESS) return false;
    return desc->coeff.kind == FEELGPU_COEFF_CONST;
}

int feelgpu_launch_form2_rowplan(feelgpu_pattern_t pat, const FormParams& hp, const feelgpu_form2_desc* desc,
                                 int accumulate, double* a, cudaStream_t s)
{
    feelgpu_context_t ctx = pat->ctx;
    RowPlan* p = pat->plan;
    feelgpu_space_t sp = pat->test;
    feelgpu_mesh_t m = sp->mesh;
    const feelgpu_tables& t = desc->tables;
    const int n = p->n, nq = t.nq;
    // canonical rows of the reference matrices, from the caller's PreCompute tables
    RowTables tab;
    std::memset(&tab, 0, sizeof(tab));
    RowPlanRot rot[20];
    rowplan_rotation_tables(p->order, rot, tab.canon);
    const int form = desc->kind == FEELGPU_FORM2_STIFFNESS ? 1 : 0;
    for (int kd = 0; kd < 3; ++kd) {
        const int i0 = tab.canon[kd];
        if (i0 < 0) { tab.canon[kd] = 0; continue; }
        for (int j = 0; j < n; ++j) {
            if (form == 0) {
                double mm = 0;
                for (int q = 0; q < nq; ++q) mm += t.w[q] * (t.phi[i0 * nq + q] * t.phi[j * nq + q]);
                tab.T[kd][j][0] = mm;
            } else {
                double M[3][3];
                for (int b = 0; b < 3; ++b)
                    for (int d = 0; d < 3; ++d) {
                        double v = 0;
                        for (int q = 0; q < nq; ++q) v += t.w[q] * (t.dphi[(i0 * nq + q) * 3 + b] * t.dphi[(j * nq + q) * 3 + d]);
                        M[b][d] = v;
                    }
                // A = sum_bd G'_bd M_bd with G'_bd = 4 S'_{b+1,d+1}
                tab.T[kd][j][0] = 4 * M[0][0];
                tab.T[kd][j][1] = 4 * M[1][1];
                tab.T[kd][j][2] = 4 * M[2][2];
                tab.T[kd][j][3] = 4 * (M[0][1] + M[1][0]);
                tab.T[kd][j][4] = 4 * (M[0][2] + M[2][0]);
                tab.T[kd][j][5] = 4 * (M[1][2] + M[2][1]);
            }
        }
    }
    const size_t sh = rowplan_smem_bytes(p, form);
    const int grid = (int)p->ncta, threads = ROWPLAN_WARPS * 32;
    if (grid == 0) return FEELGPU_OK;
#define RLAUNCH(O, F)                                                                                               \
    do {                                                                                                            \
        FEELGPU_CHECK_CUDA(ctx, cudaFuncSetAttribute(k_rowplan<O, F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh)); \
        k_rowplan<O, F><<<grid, threads, sh, s>>>(p->d_ctas, p->d_bundles, p->d_visits, p->d_units, p->d_srcs,       \
                                                  p->d_omap, p->d_rows, m->d_xyz, m->d_conn, pat->d_ia, a,          \
                                                  accumulate, p->ncomp, p->maxchunks, p->maxrounds, hp.cval[0], tab); \
    } while (0)
    switch (p->order * 2 + form) {
    case 2: RLAUNCH(1, 0); break;
    case 3: RLAUNCH(1, 1); break;
    case 4: RLAUNCH(2, 0); break;
    case 5: RLAUNCH(2, 1); break;
    case 6: RLAUNCH(3, 0); break;
    case 7: RLAUNCH(3, 1); break;
    default: ctx->last_error = "rowplan: unsupported order"; return FEELGPU_ERR_UNSUPPORTED;
    }
#undef RLAUNCH
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    return FEELGPU_OK;
}
