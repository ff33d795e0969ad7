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

// Canonical rows of the reference matrices.  They are read with a RUNTIME kind index (LDC through the constant cache,
// all lanes the same address): with a compile-time index ptxas hoists the loads out of the bundle loop, 120 doubles do
// not fit the uniform register file and spill; reading them from shared memory costs 30 LSU wavefronts per bundle on a
// kernel that is LSU-bound.
__constant__ RowTables c_rowtab;

// one unit's row of the re-labelled local matrix -> stage[entry][lane]; returns the canonical (diagonal) entry.
template <int ORDER, int FORM, int KIND>
__device__ __forceinline__ void unit_row(int kind, const double* __restrict__ sS, uint32_t u, double* st)
{
    constexpr int N = (ORDER + 1) * (ORDER + 2) * (ORDER + 3) / 6;
    const int vis = u & 0x3FF;
    if (FORM == 0) {
        const double cj = sS[vis];
#pragma unroll
        for (int j = 0; j < N; ++j) {
            const double v = cj * c_rowtab.T[kind][j][0];
            st[j * 32] = v;
        }
    } else {
        const double s01 = sS[((u >> 10) & 7) * ROWPLAN_VMAX + vis];
        const double s02 = sS[((u >> 13) & 7) * ROWPLAN_VMAX + vis];
        const double s03 = sS[((u >> 16) & 7) * ROWPLAN_VMAX + vis];
        const double s12 = sS[((u >> 19) & 7) * ROWPLAN_VMAX + vis];
        const double s13 = sS[((u >> 22) & 7) * ROWPLAN_VMAX + vis];
        const double s23 = sS[((u >> 25) & 7) * ROWPLAN_VMAX + vis];
        // rows of the barycentric Gram matrix sum to zero: S_kk = -sum_{l != k} S_kl
        const double s11 = -(s01 + s12 + s13), s22 = -(s02 + s12 + s23), s33 = -(s03 + s13 + s23);
        // partially unrolled on purpose: a rolled loop keeps the table loads next to their uses (no hoisting, no spills)
#pragma unroll 2
        for (int j = 0; j < N; ++j) {
            double v = s11 * c_rowtab.T[kind][j][0];
            v = fma(s22, c_rowtab.T[kind][j][1], v);
            v = fma(s33, c_rowtab.T[kind][j][2], v);
            v = fma(s12, c_rowtab.T[kind][j][3], v);
            v = fma(s13, c_rowtab.T[kind][j][4], v);
            v = fma(s23, c_rowtab.T[kind][j][5], v);
            st[j * 32] = v;
        }
    }
}

template <int ORDER, int FORM, int NCOMP>   // FORM 0: mass, 1: stiffness; NCOMP components (block = value * Id)
__global__ void __launch_bounds__(ROWPLAN_WARPS * 32, 3)
k_rowplan(const RowPlanCta* __restrict__ ctas, const uint4* __restrict__ hdrs, const int32_t* __restrict__ visits,
          const uint4* __restrict__ blob, const double* __restrict__ xyz, const int4* __restrict__ conn,
          double* __restrict__ a, int accumulate, int maxchunks, int maxrounds, double cval)
{
    constexpr int N = (ORDER + 1) * (ORDER + 2) * (ORDER + 3) / 6;
    constexpr int NS = FORM == 1 ? 6 : 1;
    extern __shared__ double sm[];
    double* sS = sm;                                    // [NS][ROWPLAN_VMAX]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int stageSize = ROWPLAN_ZOFF + maxchunks * N * 32;   // zero cell (+pad), then the tile
    const int outSize = 32 + maxrounds * 32 + 64;              // {zero, diagonal}, task results, row table
    double* stage = sS + NS * ROWPLAN_VMAX + warp * (stageSize + outSize);
    double* out = stage + stageSize;
    const RowPlanCta cta = ctas[blockIdx.x];
    uint32_t b = warp;
    uint4 hq = __ldg(&hdrs[cta.bundle_off + b]);        // header array is padded: always in bounds
    uint4 hnext = __ldg(&hdrs[cta.bundle_off + b + ROWPLAN_WARPS]);

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
    if (lane == 0) stage[0] = 0.0;
    __syncthreads();

    uint4* rowsm = reinterpret_cast<uint4*>(out + 32 + maxrounds * 32);   // the bundle's row table {base, len, offset}
    const uint32_t stg = (uint32_t)__cvta_generic_to_shared(stage);
    const uint32_t outs = (uint32_t)__cvta_generic_to_shared(out);
    if (lane == 0) out[0] = 0.0;

    while (b < cta.nbundle) {
        // software pipeline: the header two bundles ahead is requested now
        const uint4 hnext2 = __ldg(&hdrs[min(cta.bundle_off + b + 2 * ROWPLAN_WARPS, cta.bundle_off + cta.nbundle)]);
        const int niter4 = hq.y & 0xFFFF, nout = hq.y >> 16;
        const int nrows = hq.z & 0xFF, nchunks = (hq.z >> 8) & 0xFF, kind = (hq.z >> 16) & 0xFF, flags = hq.z >> 24;
        const uint4* B = blob + hq.x;
        // ---- all loads of the bundle are issued up front ----
        const uint32_t u0 = __ldg(reinterpret_cast<const uint32_t*>(B) + lane);
        const uint2* sp = reinterpret_cast<const uint2*>(B + nchunks * 8 + nrows) + (size_t)lane * niter4;
        const uint16_t* om = reinterpret_cast<const uint16_t*>(B + nchunks * 8 + nrows + niter4 * 16);
        uint2 cur = niter4 ? __ldg(sp) : make_uint2(0u, 0u);
        uint32_t omr[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) omr[r] = (lane + 32 * r < nout) ? (uint32_t)__ldg(om + lane + 32 * r) : 0u;
        if (lane < nrows) rowsm[lane] = __ldg(B + nchunks * 8 + lane);

        // ---- phase 1: lane = unit; its row of the (re-labelled) local matrix goes to stage[entry][lane] ----
        // every lane computes (inactive lanes read visit 0 and write a tile column nobody references)
        for (int ch = 0; ch < nchunks; ++ch) {
            const uint32_t u = ch == 0 ? u0 : __ldg(reinterpret_cast<const uint32_t*>(B) + ch * 32 + lane);
            double* st = stage + ROWPLAN_ZOFF + ch * N * 32 + lane;
            if (ORDER == 1 || kind == 0) unit_row<ORDER, FORM, 0>(kind, sS, u, st);
            else if (ORDER == 2 || kind == 1) unit_row<ORDER, FORM, (ORDER >= 2 ? 1 : 0)>(kind, sS, u, st);
            else unit_row<ORDER, FORM, (ORDER >= 3 ? 2 : 0)>(kind, sS, u, st);
            if (flags & ROWPLAN_FLAG_WARPDIAG) {
                // single row: its diagonal is the warp sum of the canonical entries of all units
                const int canon = kind == 0 ? 0 : (kind == 1 ? 4 + 2 * (ORDER - 1) : 4 + 6 * (ORDER - 1) + 3);
                double dv = (u & ROWPLAN_UNIT_ACTIVE) ? st[canon * 32] : 0.0;
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) dv += __shfl_xor_sync(0xffffffffu, dv, off);
                if (lane == 0) out[1] = ch == 0 ? dv : out[1] + dv;
            }
        }
        __syncwarp();
        // ---- phase 2: lane = a balanced list of output nonzeros, summed from the staging tile in plan order ----
        {
            double acc = 0.0;
            uint32_t op = outs + (32 + lane) * 8;
            for (int it = 0; it < niter4; ++it) {
                const uint2 nxt = (it + 1 < niter4) ? __ldg(sp + it + 1) : make_uint2(0u, 0u);
                const uint32_t e[4] = {cur.x & 0xFFFFu, cur.x >> 16, cur.y & 0xFFFFu, cur.y >> 16};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    double v;
                    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(stg + ((e[k] & 0x7FFFu) << 3)));
                    acc += v;
                    // predicated (branch-free) end-of-output: store the sum, advance the result slot, restart
                    const uint32_t end = e[k] >> 15;
                    asm volatile("{ .reg .pred p; setp.ne.u32 p, %0, 0; @p st.shared.f64 [%1], %2; }" ::"r"(end), "r"(op), "d"(acc)
                                 : "memory");
                    op += end << 8;
                    acc = end ? 0.0 : acc;
                }
                cur = nxt;
            }
        }
        __syncwarp();
        // ---- phase 3: coalesced store of the rows' value segments ----
        auto emit = [&](int o, uint32_t m) {
            const uint4 r = rowsm[m >> 11];
            const double v = out[m & 0x3FFu];
            const int64_t base = (int64_t)(((uint64_t)r.y << 32) | r.x);
            const int p = o - (int)r.w;
            if (NCOMP == 1) {
                if (accumulate) a[base + p] += v;
                else a[base + p] = v;
            } else {
                // product space, block = v * Id: component row c1 starts at base + c1*len*NCOMP
                const int len = (int)r.z;
#pragma unroll
                for (int c1 = 0; c1 < NCOMP; ++c1)
#pragma unroll
                    for (int c2 = 0; c2 < NCOMP; ++c2) {
                        double* dst = a + base + (int64_t)c1 * len * NCOMP + (int64_t)p * NCOMP + c2;
                        if (c1 == c2) { if (accumulate) *dst += v; else *dst = v; }
                        else if (!accumulate) *dst = 0.0;
                    }
            }
        };
#pragma unroll
        for (int r = 0; r < 4; ++r)
            if (lane + 32 * r < nout) emit(lane + 32 * r, omr[r]);
        for (int o = lane + 128; o < nout; o += 32) emit(o, (uint32_t)__ldg(om + o));
        __syncwarp();
        hq = hnext;
        hnext = hnext2;
        b += ROWPLAN_WARPS;
    }
}

size_t rowplan_smem_bytes(const RowPlan* p, int form)
{
    const int NS = form == 1 ? 6 : 1;
    const size_t stageSize = ROWPLAN_ZOFF + (size_t)p->maxchunks * p->n * 32, outSize = 32 + (size_t)p->maxrounds * 32 + 64;
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
    p->nbundle = (int64_t)hp.hdrs.size() - ROWPLAN_WARPS;
    cudaError_t err = upload(&p->d_ctas, hp.ctas, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_hdrs, hp.hdrs, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_visits, hp.visits, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_blob, hp.blob, p->bytes);
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
    cudaFree(p->d_hdrs);
    cudaFree(p->d_visits);
    cudaFree(p->d_blob);
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
    // stream-ordered upload of the tables (one context drives one stream at a time, see feelgpu.h)
    FEELGPU_CHECK_CUDA(ctx, cudaMemcpyToSymbolAsync(c_rowtab, &tab, sizeof(RowTables), 0, cudaMemcpyHostToDevice, s));
    const size_t sh = rowplan_smem_bytes(p, form);
    const int grid = (int)p->ncta, threads = ROWPLAN_WARPS * 32;
    if (grid == 0) return FEELGPU_OK;
#define RLAUNCH1(O, F, C)                                                                                          \
    do {                                                                                                            \
        FEELGPU_CHECK_CUDA(ctx, cudaFuncSetAttribute(k_rowplan<O, F, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh)); \
        k_rowplan<O, F, C><<<grid, threads, sh, s>>>(p->d_ctas, reinterpret_cast<const uint4*>(p->d_hdrs), p->d_visits, \
                                                     reinterpret_cast<const uint4*>(p->d_blob), m->d_xyz, m->d_conn, a, \
                                                     accumulate, p->maxchunks, p->maxrounds, hp.cval[0]);          \
    } while (0)
#define RLAUNCH(O, F)                         \
    do {                                      \
        if (p->ncomp == 1) RLAUNCH1(O, F, 1); \
        else RLAUNCH1(O, F, 3);               \
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
#undef RLAUNCH1
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    return FEELGPU_OK;
}
