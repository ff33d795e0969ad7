// kernels_rplan.cu -- row-thread assembly kernel (see rplan.h for the design) and the plan's device mirror.
//
// Replaces, for coefficient-weighted mass / stiffness forms (cst, idv(k), expr("...")) of every order and for the
// vector-elasticity form, the reference's per-element update/integrate/assemble + MatSetValues(ADD_VALUES)
// (feelvf/integrator.hpp:1093-1207, feelvf/bilinearformcontext.hpp:585-589, feelalg/matrixpetsc.cpp:917-936): the CTA
// that owns a patch of CSR rows forms, one thread per (row, element) unit, the rows of the local matrices of the elements
// around the patch's nodes by the quadrature sum (feelpoly/pointsetquadrature.hpp:403-416) and adds them into the row's
// segment of its shared-memory row block; the
// block leaves the SM once, with coalesced stores.  No L2 atomics and no zero fill of the CSR array; the adds into the row
// block are shared-memory atomics (rplan.h), so the summation order inside a value is not fixed (values within 1e-12).
#include <string>
#include "feelgpu_internal.h"
#include "device_common.cuh"
#include "rplan.h"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>

int feelgpu_upload_tables_public(feelgpu_context_t ctx, const FormParams& hp, const feelgpu_tables& t, cudaStream_t s,
                                 const FormParams** dP, const double** phi, const double** dphi, const double** phif,
                                 const double** M9, const double** Mm);

namespace {

struct RArgs {
    const RPlanPatch* patches;
    const int32_t* visits;
    const uint4* rows;
    const uint32_t* units;
    const double* xyz;
    const int4* conn;
    const uint16_t* pos;
    const FormParams* P;
    DevTables T;
    const int32_t* l2gf;
    const double* fvals;
    const double* kq;            // [nElt][nq] coefficient at the quadrature points (variable coefficients)
    double* a;
    int accumulate, maxslots, maxvis, maxrows;
};

// Reference tables in constant memory: every lane of a warp reads the same entry, so they ride as constant-bank operands
// of the DFMAs instead of occupying the shared-memory pipe (one shared-memory double per DFMA caps the kernel at a quarter
// of the FP64 rate).  Layout: stiffness / elasticity [q][c][j] (c = reference direction), mass [q][j]; then the weights.
#define RTAB_MAX (3 * FEELGPU_MAX_NQ * FEELGPU_MAX_NLOC)
__constant__ double c_rtab[RTAB_MAX];
__constant__ double c_rw[FEELGPU_MAX_NQ];
__constant__ double c_rxi[3 * FEELGPU_MAX_NQ];

// shared-memory add of one double (a compare-and-swap loop in SASS, ATOMS.CAST.SPIN: no native fp64 shared atomic)
__device__ __forceinline__ void sadd(double* p, double v) { atomicAdd(p, v); }

// (issuing the compare-and-swap chains of several values together -- loads, sums, first attempts back to back -- was measured:
// slower, 7.6 -> 12.6 ms on the elasticity form, 22.1 -> 22.9 ms on P3: more registers, no gain in the pipe)

// the closed-form coefficient with its program in shared memory (the interpreter of device_common.cuh reads the program
// from global memory: one dependent load per operation)
__device__ double eval_expr_sm(int len, const int* ops, const double* consts, double x, double y, double z)
{
    double st[16];
    int sp = 0;
    for (int k = 0; k < len; ++k) {
        switch (ops[k]) {
        case FEELGPU_OP_X: st[sp++] = x; break;
        case FEELGPU_OP_Y: st[sp++] = y; break;
        case FEELGPU_OP_Z: st[sp++] = z; break;
        case FEELGPU_OP_CONST: st[sp++] = consts[k]; break;
        case FEELGPU_OP_ADD: --sp; st[sp - 1] += st[sp]; break;
        case FEELGPU_OP_SUB: --sp; st[sp - 1] -= st[sp]; break;
        case FEELGPU_OP_MUL: --sp; st[sp - 1] *= st[sp]; break;
        case FEELGPU_OP_DIV: --sp; st[sp - 1] /= st[sp]; break;
        case FEELGPU_OP_NEG: st[sp - 1] = -st[sp - 1]; break;
        case FEELGPU_OP_SIN: st[sp - 1] = sin(st[sp - 1]); break;
        case FEELGPU_OP_COS: st[sp - 1] = cos(st[sp - 1]); break;
        case FEELGPU_OP_EXP: st[sp - 1] = exp(st[sp - 1]); break;
        case FEELGPU_OP_LOG: st[sp - 1] = log(st[sp - 1]); break;
        case FEELGPU_OP_SQRT: st[sp - 1] = sqrt(st[sp - 1]); break;
        case FEELGPU_OP_POW: {
            --sp;
            const double ex = st[sp], b = st[sp - 1];
            // small integer exponents by multiplication (x^2 is what pow returns, correctly rounded; 3 and 4 differ by an ulp)
            if (ex == 2.0) st[sp - 1] = b * b;
            else if (ex == 3.0) st[sp - 1] = b * b * b;
            else if (ex == 4.0) { const double b2 = b * b; st[sp - 1] = b2 * b2; }
            else st[sp - 1] = pow(b, ex);
            break;
        }
        default: break;
        }
    }
    return st[0];
}

// Coefficient pass: k(x_q) at every quadrature point of every element, once (thread = (element, q), full occupancy: the
// interpreter is a chain of dependent loads that the 16 warps of the assembly kernel cannot hide, and a visit-based
// evaluation would repeat it for every patch an element touches).  kq[e][q], read back by phase A of the assembly kernel.
__global__ void __launch_bounds__(256)
k_rplan_coeff(const FormParams* __restrict__ P, int64_t nElt, int nq, const double* __restrict__ xyz, const int4* __restrict__ conn,
              const double* __restrict__ phif, const int32_t* __restrict__ l2gf, const double* __restrict__ fvals,
              double* __restrict__ kq)
{
    __shared__ int sOps[FEELGPU_EXPR_MAX_LEN];
    __shared__ double sCst[FEELGPU_EXPR_MAX_LEN];
    const int len = P->expr_len, kind = P->coeff_kind, nf = P->nf;
    for (int k = threadIdx.x; k < len; k += blockDim.x) { sOps[k] = P->expr_ops[k]; sCst[k] = P->expr_consts[k]; }
    __syncthreads();
    const int64_t total = nElt * nq;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t e = t / nq;
        const int q = (int)(t - e * nq);
        double v;
        if (kind == FEELGPU_COEFF_EXPR) {
            const int4 c = __ldg(&conn[e]);
            double x0[3], x1[3], x2[3], x3[3];
            load_vertex(xyz, c.x, x0);
            load_vertex(xyz, c.y, x1);
            load_vertex(xyz, c.z, x2);
            load_vertex(xyz, c.w, x3);
            const double xi0 = c_rxi[3 * q], xi1 = c_rxi[3 * q + 1], xi2 = c_rxi[3 * q + 2];
            const double N0 = -0.5 * (1.0 + xi0 + xi1 + xi2), N1 = 0.5 * (1.0 + xi0), N2 = 0.5 * (1.0 + xi1), N3 = 0.5 * (1.0 + xi2);
            double x[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) x[a] = x0[a] * N0 + x1[a] * N1 + x2[a] * N2 + x3[a] * N3;
            v = eval_expr_sm(len, sOps, sCst, x[0], x[1], x[2]);
        } else {
            const int32_t* lf = l2gf + e * nf;
            v = 0.0;
            for (int i = 0; i < nf; ++i) v += __ldg(&fvals[__ldg(lf + i)]) * __ldg(&phif[i * nq + q]);
        }
        kq[t] = v;
    }
}

// FORM 0: k*idt(u)*id(v), 1: k*gradt(u)*trans(grad(v)), 2: elasticity.  Thread = unit.  NQ > 0: quadrature points known at
// compile time, tables in constant memory, loops unrolled; NQ = 0: any nq, tables in shared memory (mass / stiffness only).
constexpr int rplan_threads(int form) { return form == 2 ? 512 : 640; }   // elasticity holds 3x3 blocks in registers

template <int ORDER, int FORM, int NQ>
__global__ void __launch_bounds__(rplan_threads(FORM), 1)
k_rplan(const RArgs A)
{
    constexpr int N = (ORDER + 1) * (ORDER + 2) * (ORDER + 3) / 6;
    extern __shared__ double sm[];
    const FormParams* __restrict__ P = A.P;
    const int nq = NQ > 0 ? NQ : P->nq, nf = P->nf, maxvis = A.maxvis;
    const int nrec = FORM == 0 ? nq : FORM == 1 ? 6 + nq : 10;
    double* blk = sm;                                        // [maxslots] the patch's row block
    double* rec = blk + A.maxslots;                          // [nrec][maxvis] per-visit records
    double* tab = rec + (size_t)nrec * maxvis;               // reference tables, [q][c][j] / [q][j] (read with a per-lane row dof)
    const int ntab = FORM == 0 ? N * nq : N * nq * 3;
    double* tphif = tab + ntab;                              // [nf][nq] basis of the coefficient field
    int32_t* sE = reinterpret_cast<int32_t*>(tphif + nf * nq);   // [maxvis] element of every visit
    uint32_t* sRoff = reinterpret_cast<uint32_t*>(sE + maxvis);  // [maxrows][2] row-block offset and length of every row
    const int tid = threadIdx.x;

    const uint4* pq = reinterpret_cast<const uint4*>(A.patches + blockIdx.x);
    const uint4 p0 = __ldg(pq), p1 = __ldg(pq + 1);
    const uint32_t visit_off = p0.x, nvisit = p0.y, row_off = p0.z, nrows = p0.w, unit_off = p1.x, nunits = p1.y, nslots = p1.z;
    const uint4* rows = A.rows + 2 * (size_t)row_off;

    if (FORM == 0) {
        for (int k = tid; k < N * nq; k += blockDim.x) { const int j = k / nq, q = k - j * nq; tab[q * N + j] = A.T.phi[k]; }
    } else {
        for (int k = tid; k < N * nq; k += blockDim.x) {
            const int j = k / nq, q = k - j * nq;
#pragma unroll
            for (int c = 0; c < 3; ++c) tab[(q * 3 + c) * N + j] = A.T.dphi[3 * k + c];
        }
    }
    for (int k = tid; k < nf * nq; k += blockDim.x) tphif[k] = A.T.phif[k];
    const bool varcoef = FORM != 2 && P->coeff_kind != FEELGPU_COEFF_CONST;
    for (uint32_t k = tid; k < nslots; k += blockDim.x) blk[k] = 0.0;
    for (uint32_t r = tid; r < nrows; r += blockDim.x) {
        const uint4 q0 = __ldg(rows + 2 * r);
        sRoff[2 * r] = q0.z;
        sRoff[2 * r + 1] = q0.w;
    }

    // ---- phase A: geometry of every element touching the patch's rows, once (thread = visit) ----
    for (uint32_t v = tid; v < nvisit; v += blockDim.x) {
        const int32_t e = __ldg(&A.visits[visit_off + v]);
        sE[v] = e;
        Geo g;
        geo_update(A.xyz, __ldg(&A.conn[e]), g);
        if (FORM == 2) {
#pragma unroll
            for (int a1 = 0; a1 < 3; ++a1)
#pragma unroll
                for (int b = 0; b < 3; ++b) rec[(a1 * 3 + b) * maxvis + v] = g.B[a1][b];
            rec[9 * maxvis + v] = g.J;
        } else {
            if (FORM == 1) {
                // G = B^T B (the factor J goes with the weights): G_bd = sum_a B_ab B_ad, order 00 01 02 11 12 22
                int k = 0;
#pragma unroll
                for (int b = 0; b < 3; ++b)
#pragma unroll
                    for (int d = b; d < 3; ++d)
                        rec[(k++) * maxvis + v] = g.B[0][b] * g.B[0][d] + g.B[1][b] * g.B[1][d] + g.B[2][b] * g.B[2][d];
            }
            // prod[q] * k(x_q) = w_q J k(x_q)   (feelpoly/pointsetquadrature.hpp:481-490)
            if (!varcoef) {
                const double cj = P->cval[0] * g.J;
                for (int q = 0; q < nq; ++q) rec[((FORM == 1 ? 6 : 0) + q) * maxvis + v] = c_rw[q] * cj;
            } else {
                const double* kp = A.kq + (int64_t)e * nq;
                for (int q = 0; q < nq; ++q) rec[((FORM == 1 ? 6 : 0) + q) * maxvis + v] = (c_rw[q] * g.J) * __ldg(kp + q);
            }
        }
    }
    __syncthreads();

    // ---- phase B: one thread per unit, batches of blockDim.x units in round order ----
    const uint32_t* units = A.units + unit_off;
    for (uint32_t b0 = 0; b0 < nunits; b0 += blockDim.x) {
        const uint32_t b1 = min(b0 + blockDim.x, nunits), u = b0 + tid;
        const bool active = u < b1;
        const uint32_t w = active ? __ldg(&units[u]) : 0u;
        const int v = w & 0xFFF, i = (w >> 12) & 31;
        const uint32_t roff = sRoff[2 * (w >> 17)], rlen = sRoff[2 * (w >> 17) + 1];
        const uint16_t* pp = A.pos + ((int64_t)sE[v] * N + i) * N;
        // the unit's positions in its row (N 16-bit values, 4-byte aligned: N is even) are requested before the quadrature loop
        uint32_t pw[N / 2];
        if (active) {
#pragma unroll
            for (int j = 0; j < N / 2; ++j) pw[j] = __ldg(reinterpret_cast<const uint32_t*>(pp) + j);
        }
        auto posj = [&](int j) { return (pw[j >> 1] >> ((j & 1) * 16)) & 0xFFFFu; };
        if (FORM != 2) {
            double acc[N];
#pragma unroll
            for (int j = 0; j < N; ++j) acc[j] = 0.0;
            if (active) {
                if (FORM == 0) {
                    if (NQ > 0) {
#pragma unroll
                        for (int q = 0; q < NQ; ++q) {
                            const double s = rec[q * maxvis + v] * tab[q * N + i];
#pragma unroll
                            for (int j = 0; j < N; ++j) acc[j] = fma(s, c_rtab[q * N + j], acc[j]);
                        }
                    } else {
                        for (int q = 0; q < nq; ++q) {
                            const double s = rec[q * maxvis + v] * tab[q * N + i];
                            const double2* tq = reinterpret_cast<const double2*>(tab + q * N);
#pragma unroll
                            for (int j = 0; j < N / 2; ++j) {
                                const double2 t = tq[j];
                                acc[2 * j] = fma(s, t.x, acc[2 * j]);
                                acc[2 * j + 1] = fma(s, t.y, acc[2 * j + 1]);
                            }
                        }
                    }
                } else {
                    const double G00 = rec[0 * maxvis + v], G01 = rec[1 * maxvis + v], G02 = rec[2 * maxvis + v],
                                 G11 = rec[3 * maxvis + v], G12 = rec[4 * maxvis + v], G22 = rec[5 * maxvis + v];
                    auto point = [&](int q, double& s0, double& s1, double& s2) {
                        const double wk = rec[(6 + q) * maxvis + v];
                        const double* tq = tab + q * 3 * N;
                        const double a0 = tq[i], a1 = tq[N + i], a2 = tq[2 * N + i];
                        // s = w_q J k_q G grad^(phi_i)(q)
                        s0 = wk * (G00 * a0 + G01 * a1 + G02 * a2);
                        s1 = wk * (G01 * a0 + G11 * a1 + G12 * a2);
                        s2 = wk * (G02 * a0 + G12 * a1 + G22 * a2);
                    };
                    if (NQ > 0) {
#pragma unroll
                        for (int q = 0; q < NQ; ++q) {
                            double s0, s1, s2;
                            point(q, s0, s1, s2);
#pragma unroll
                            for (int j = 0; j < N; ++j)
                                acc[j] = fma(s2, c_rtab[(q * 3 + 2) * N + j], fma(s1, c_rtab[(q * 3 + 1) * N + j], fma(s0, c_rtab[q * 3 * N + j], acc[j])));
                        }
                    } else {
                        for (int q = 0; q < nq; ++q) {
                            double s0, s1, s2;
                            point(q, s0, s1, s2);
                            const double* tq = tab + q * 3 * N;
                            const double2* t0 = reinterpret_cast<const double2*>(tq);
                            const double2* t1 = reinterpret_cast<const double2*>(tq + N);
                            const double2* t2 = reinterpret_cast<const double2*>(tq + 2 * N);
#pragma unroll
                            for (int j = 0; j < N / 2; ++j) {
                                const double2 x = t0[j], y = t1[j], z = t2[j];
                                acc[2 * j] = fma(s2, z.x, fma(s1, y.x, fma(s0, x.x, acc[2 * j])));
                                acc[2 * j + 1] = fma(s2, z.y, fma(s1, y.y, fma(s0, x.y, acc[2 * j + 1])));
                            }
                        }
                    }
                }
            }
            if (active) {
                double* rb = blk + roff;
#pragma unroll
                for (int j = 0; j < N; ++j) sadd(rb + posj(j), acc[j]);
            }
        } else if (NQ > 0) {
            // elasticity: physical gradients of the row function at the quadrature points, then column by column the 3x3 block
            // a_cd = sum_q w_q J (grad phi_i)_c (grad phi_j)_d ; block(c1,c2) = lambda a_c1c2 + mu a_c2c1 + mu delta tr a
            // (feelvf/inner.hpp:374-420).  The blocks of a unit (N x 9 values) are added round by round as they are formed:
            // one round loop per column.
            double B[3][3], pi[NQ > 0 ? NQ : 1][3];
            double J = 0.0;
            if (active) {
#pragma unroll
                for (int k = 0; k < 9; ++k) B[k / 3][k % 3] = rec[k * maxvis + v];
                J = rec[9 * maxvis + v];
#pragma unroll
                for (int q = 0; q < NQ; ++q) {
                    const double* tq = tab + q * 3 * N;
                    const double a0 = tq[i], a1 = tq[N + i], a2 = tq[2 * N + i];
                    const double wj = c_rw[q] * J;
#pragma unroll
                    for (int c = 0; c < 3; ++c) pi[q][c] = wj * (B[c][0] * a0 + B[c][1] * a1 + B[c][2] * a2);
                }
            }
            const double lambda = P->lambda, mu = P->mu;
            // columns in chunks of JC (register budget)
            constexpr int JC = 3;
#pragma unroll
            for (int j0 = 0; j0 < N; j0 += JC) {
                double blkv[JC][9];
                if (active) {
#pragma unroll
                    for (int jj = 0; jj < JC; ++jj) {
                        const int j = j0 + jj;
                        if (j >= N) break;
                        double a[3][3];
#pragma unroll
                        for (int c = 0; c < 9; ++c) a[c / 3][c % 3] = 0.0;
#pragma unroll
                        for (int q = 0; q < NQ; ++q) {
                            double pj[3];
#pragma unroll
                            for (int d = 0; d < 3; ++d)
                                pj[d] = B[d][0] * c_rtab[q * 3 * N + j] + B[d][1] * c_rtab[(q * 3 + 1) * N + j] + B[d][2] * c_rtab[(q * 3 + 2) * N + j];
#pragma unroll
                            for (int c = 0; c < 3; ++c)
#pragma unroll
                                for (int d = 0; d < 3; ++d) a[c][d] = fma(pi[q][c], pj[d], a[c][d]);
                        }
                        const double tr = a[0][0] + a[1][1] + a[2][2];
#pragma unroll
                        for (int c1 = 0; c1 < 3; ++c1)
#pragma unroll
                            for (int c2 = 0; c2 < 3; ++c2)
                                blkv[jj][c1 * 3 + c2] = lambda * a[c1][c2] + mu * a[c2][c1] + (c1 == c2 ? mu * tr : 0.0);
                    }
                }
                if (active) {
#pragma unroll
                    for (int jj = 0; jj < JC; ++jj) {
                        const int j = j0 + jj;
                        if (j >= N) break;
                        const uint32_t p0c = posj(j);
#pragma unroll
                        for (int c1 = 0; c1 < 3; ++c1) {
                            double* rb = blk + roff + c1 * rlen + p0c;
#pragma unroll
                            for (int c2 = 0; c2 < 3; ++c2) sadd(rb + c2, blkv[jj][c1 * 3 + c2]);
                        }
                    }
                }
            }
        }
    }
    __syncthreads();

    // ---- phase C: the row block goes to the CSR value array: one contiguous piece per node row, a warp per row ----
    {
        const int lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
        constexpr int NC = FORM == 2 ? 3 : 1;
        for (uint32_t r = warp; r < nrows; r += nwarp) {
            const uint4 q0 = __ldg(rows + 2 * r);
            double* dst = A.a + (int64_t)(((uint64_t)q0.y << 32) | q0.x);
            const double* src = blk + q0.z;
            const uint32_t cnt = q0.w * NC;
            if (A.accumulate) {
                for (uint32_t k = lane; k < cnt; k += 32) dst[k] += src[k];
            } else {
                for (uint32_t k = lane; k < cnt; k += 32) dst[k] = src[k];
            }
        }
    }
}

template <typename T>
cudaError_t upload(T** d, const std::vector<T>& h, int64_t& bytes)
{
    cudaError_t err = cudaMalloc(d, std::max<size_t>(h.size(), 1) * sizeof(T));
    if (err != cudaSuccess) return err;
    bytes += (int64_t)(h.size() * sizeof(T));
    return cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
}

size_t rplan_smem_bytes(const RPlan* p, int form, int nq, int nf)
{
    const int n = p->n;
    const size_t nrec = form == 0 ? nq : form == 1 ? 6 + nq : 10;
    const size_t ntab = form == 0 ? (size_t)n * nq : (size_t)n * nq * 3;
    return sizeof(double) * ((size_t)p->maxslots + nrec * p->maxvisits + ntab + (size_t)nf * nq) +
           sizeof(int32_t) * ((size_t)p->maxvisits + 2 * (size_t)p->maxrows) + 16;
}

int env_int(const char* name, int dflt)
{
    const char* e = std::getenv(name);
    return e && *e ? std::atoi(e) : dflt;
}

} // namespace

int feelgpu_rplan_build(feelgpu_pattern_t pat, const int64_t* h_ia)
{
    feelgpu_context_t ctx = pat->ctx;
    feelgpu_space_t sp = pat->test;
    feelgpu_mesh_t m = sp->mesh;
    if (pat->test != pat->trial) return FEELGPU_OK;   // the plan needs test == trial
    // patch limits: the row block and the visit records share the SM's 227 KB with the reference tables
    int maxslots, maxvis;
    if (sp->ncomp == 3) { maxslots = 20480; maxvis = 256; }
    else if (sp->order == 1) { maxslots = 8192; maxvis = 512; }
    else if (sp->order == 2) { maxslots = 12288; maxvis = 384; }
    else { maxslots = 16384; maxvis = 256; }
    maxslots = env_int("FEELGPU_RPLAN_SLOTS", maxslots);
    maxvis = env_int("FEELGPU_RPLAN_VISITS", maxvis);
    const int maxrows = env_int("FEELGPU_RPLAN_ROWS", 1 << 20);
    RPlanHost hp;
    const auto t0 = std::chrono::steady_clock::now();
    if (!rplan_build_host(sp->order, sp->ncomp, m->nElt, sp->h_l2g.data(), sp->nd, m->nPts, m->h_xyz.data(), m->h_conn.data(),
                          pat->nRows, h_ia, maxslots, maxvis, maxrows, feelgpu_builder_threads(ctx), hp)) {
        pat->rplan_note = hp.error;   // not an error: another path stays available
        return FEELGPU_OK;
    }
    RPlan* p = new RPlan;
    p->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    p->order = hp.order;
    p->n = hp.n;
    p->ncomp = hp.ncomp;
    p->maxslots = (hp.maxpatch_slots + 1) & ~1;      // what the kernel has to provide: the largest patch, not the limit;
    p->maxvisits = (std::max(hp.maxpatch_visits, 1) + 1) & ~1;   // even counts keep the shared arrays behind them 16-byte aligned
    p->maxrows = hp.maxpatch_rows;
    p->npatch = (int64_t)hp.patches.size();
    p->nunits = hp.nunits;
    p->nvisits = hp.nvisits;
    p->maxpatch_visits = hp.maxpatch_visits;
    p->maxpatch_rows = hp.maxpatch_rows;
    p->maxpatch_slots = hp.maxpatch_slots;
    cudaError_t err = upload(&p->d_patches, hp.patches, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_visits, hp.visits, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_rows, hp.rows, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_units, hp.units, p->bytes);
    pat->rplan = p;
    if (err != cudaSuccess) {
        ctx->last_error = std::string("row-thread plan upload: ") + cudaGetErrorString(err);
        feelgpu_rplan_destroy(pat);
        return FEELGPU_ERR_CUDA;
    }
    return FEELGPU_OK;
}

void feelgpu_rplan_destroy(feelgpu_pattern_t pat)
{
    RPlan* p = pat->rplan;
    if (!p) return;
    cudaFree(p->d_patches);
    cudaFree(p->d_visits);
    cudaFree(p->d_rows);
    cudaFree(p->d_units);
    cudaFree(p->d_kq);
    delete p;
    pat->rplan = nullptr;
}

int feelgpu_rplan_stats(feelgpu_pattern_t pat, int64_t stats[10])
{
    for (int k = 0; k < 10; ++k) stats[k] = 0;
    const RPlan* p = pat->rplan;
    if (!p) return FEELGPU_OK;
    stats[0] = 1;
    stats[1] = p->npatch;
    stats[2] = p->nunits;
    stats[3] = p->nvisits;
    stats[4] = p->bytes;
    stats[5] = p->maxpatch_slots;
    stats[6] = p->maxpatch_visits;
    stats[7] = p->maxpatch_rows;
    stats[8] = (int64_t)(p->build_seconds * 1e3);
    return FEELGPU_OK;
}

static int rplan_form(const feelgpu_form2_desc* desc)
{
    return desc->kind == FEELGPU_FORM2_MASS ? 0 : desc->kind == FEELGPU_FORM2_STIFFNESS ? 1 : 2;
}

bool feelgpu_rplan_supports(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc)
{
    const RPlan* p = pat->rplan;
    if (!p || !pat->d_pos) return false;
    const int form = rplan_form(desc);
    if (desc->tables.n != p->n || desc->tables.nq < 1 || desc->tables.nq > FEELGPU_MAX_NQ) return false;
    if (form == 2 ? (p->ncomp != 3 || p->order > 2) : p->ncomp != 1) return false;    // product spaces: elasticity; mass/stiffness blocks go through the entity plan
    if (form != 0 && !desc->tables.dphi) return false;
    if (form == 2 && desc->tables.nq != 1 && desc->tables.nq != 4) return false;   // elasticity: compile-time rules only (P1: 1 point, P2: 4)
    const int nf = desc->coeff.kind == FEELGPU_COEFF_FIELD ? desc->tables.nf : 0;
    return rplan_smem_bytes(p, form, desc->tables.nq, nf) <= 227 * 1024;
}

int feelgpu_launch_form2_rplan(feelgpu_pattern_t pat, const FormParams& hp, const feelgpu_form2_desc* desc, int accumulate,
                               double* a, cudaStream_t s)
{
    feelgpu_context_t ctx = pat->ctx;
    RPlan* p = pat->rplan;
    feelgpu_space_t sp = pat->test;
    feelgpu_mesh_t m = sp->mesh;
    const int form = rplan_form(desc);
    RArgs A;
    int rc = feelgpu_upload_tables_public(ctx, hp, desc->tables, s, &A.P, &A.T.phi, &A.T.dphi, &A.T.phif, &A.T.M9, &A.T.Mm);
    if (rc) return rc;
    A.patches = p->d_patches;
    A.visits = p->d_visits;
    A.rows = reinterpret_cast<const uint4*>(p->d_rows);
    A.units = p->d_units;
    A.xyz = m->d_xyz;
    A.conn = m->d_conn;
    A.pos = pat->d_pos;
    A.l2gf = nullptr;
    A.fvals = nullptr;
    if (hp.coeff_kind == FEELGPU_COEFF_FIELD && form != 2) {
        A.l2gf = desc->coeff.field_space->d_l2g;
        A.fvals = desc->coeff.field_values;
    }
    A.a = a;
    A.accumulate = accumulate;
    A.maxslots = p->maxslots;
    A.maxvis = p->maxvisits;
    A.maxrows = p->maxrows;
    const size_t sh = rplan_smem_bytes(p, form, hp.nq, hp.nf);
    const int grid = (int)p->npatch;
    if (grid == 0) return FEELGPU_OK;
    const int threads = std::min(rplan_threads(form), std::max(64, env_int("FEELGPU_RPLAN_THREADS", rplan_threads(form))));
    // constant-memory copy of the tables in the kernel's layout (stream-ordered; one context drives one stream at a time)
    const int n = p->n, nq = hp.nq;
    const bool fixed = !env_int("FEELGPU_RPLAN_NO_CONST", 0) && (nq == 1 || nq == 4 || nq == 5 || nq == 11 || nq == 14 || nq == 24);
    if (fixed || form == 2) {
        std::vector<double> ct((size_t)(form == 0 ? n * nq : 3 * n * nq));
        for (int j = 0; j < n; ++j)
            for (int q = 0; q < nq; ++q) {
                if (form == 0) ct[(size_t)q * n + j] = desc->tables.phi[j * nq + q];
                else
                    for (int c = 0; c < 3; ++c) ct[((size_t)q * 3 + c) * n + j] = desc->tables.dphi[(j * nq + q) * 3 + c];
            }
        FEELGPU_CHECK_CUDA(ctx, cudaMemcpyToSymbolAsync(c_rtab, ct.data(), ct.size() * sizeof(double), 0, cudaMemcpyHostToDevice, s));
    }
    FEELGPU_CHECK_CUDA(ctx, cudaMemcpyToSymbolAsync(c_rw, desc->tables.w, nq * sizeof(double), 0, cudaMemcpyHostToDevice, s));
    FEELGPU_CHECK_CUDA(ctx, cudaMemcpyToSymbolAsync(c_rxi, desc->tables.xi, 3 * nq * sizeof(double), 0, cudaMemcpyHostToDevice, s));
    A.kq = nullptr;
    if (form != 2 && hp.coeff_kind != FEELGPU_COEFF_CONST) {
        const size_t need = sizeof(double) * (size_t)m->nElt * nq;
        if (p->kq_bytes < need) {
            if (p->d_kq) FEELGPU_CHECK_CUDA(ctx, cudaFree(p->d_kq));
            p->d_kq = nullptr;
            p->kq_bytes = 0;
            FEELGPU_CHECK_CUDA(ctx, cudaMalloc(&p->d_kq, std::max<size_t>(need, 8)));
            p->kq_bytes = need;
        }
        const int64_t total = m->nElt * nq;
        const int cgrid = (int)std::max<int64_t>(1, std::min<int64_t>((total + 255) / 256, (int64_t)ctx->sm_count * 32));
        k_rplan_coeff<<<cgrid, 256, 0, s>>>(A.P, m->nElt, nq, m->d_xyz, m->d_conn, A.T.phif, A.l2gf, A.fvals, p->d_kq);
        ctx->launches++;
        FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
        A.kq = p->d_kq;
    }
#define PLAUNCH(O, F, Q)                                                                                                       \
    do {                                                                                                                       \
        FEELGPU_CHECK_CUDA(ctx, cudaFuncSetAttribute(k_rplan<O, F, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));  \
        k_rplan<O, F, Q><<<grid, threads, sh, s>>>(A);                                                                         \
    } while (0)
#define PQ(O, F)                                                  \
    do {                                                          \
        if (!fixed) PLAUNCH(O, F, 0);                             \
        else switch (nq) {                                        \
            case 1: PLAUNCH(O, F, 1); break;                      \
            case 4: PLAUNCH(O, F, 4); break;                      \
            case 5: PLAUNCH(O, F, 5); break;                      \
            case 11: PLAUNCH(O, F, 11); break;                    \
            case 14: PLAUNCH(O, F, 14); break;                    \
            default: PLAUNCH(O, F, 24); break;                    \
        }                                                         \
    } while (0)
    switch (p->order * 3 + form) {
    case 3: PQ(1, 0); break;
    case 4: PQ(1, 1); break;
    case 5: PLAUNCH(1, 2, 1); break;
    case 6: PQ(2, 0); break;
    case 7: PQ(2, 1); break;
    case 8: if (nq == 4) PLAUNCH(2, 2, 4); else PLAUNCH(2, 2, 1); break;
    case 9: PQ(3, 0); break;
    case 10: PQ(3, 1); break;
    default: ctx->last_error = "row-thread plan: unsupported order / form"; return FEELGPU_ERR_UNSUPPORTED;
    }
#undef PQ
#undef PLAUNCH
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    return FEELGPU_OK;
}

// ---- allocation-free variant (feelgpu_form2_assemble_noalloc, include/feelgpu.h) ---------------------------------------
// The coefficient pass writes k(x_q) of every (element, point) into scratch memory that the assembly kernel reads back; the
// regular path keeps that scratch (and the parameter block + reference tables) in cudaMalloc'ed buffers.  This variant keeps
// both in ONE file-scope device array of fixed size, so that the call allocates nothing: [0, NA_PAR_DOUBLES) the FormParams
// block followed by phi | dphi | phif, then FEELGPU_NOALLOC_MAX_KQ doubles of kq.  Every GPU that loads the library holds
// the array (16.05 MiB) until exit.  All calls share it: it keeps the previous call's contents (nothing in it has to start at
// zero -- the copy and the coefficient pass overwrite every entry the assembly kernel reads) and calls that overlap on the
// GPU (different streams, no ordering between them) race on it.
namespace {

constexpr size_t NA_TAB_DOUBLES = (size_t)5 * FEELGPU_MAX_NQ * FEELGPU_MAX_NLOC;           // phi (1) + dphi (3) + phif (1)
constexpr size_t NA_PAR_DOUBLES = ((sizeof(FormParams) / 8 + NA_TAB_DOUBLES + 31) / 32) * 32;   // kq starts 256-byte aligned
static_assert(sizeof(FormParams) % 8 == 0, "the tables follow the parameter block as doubles");

static __device__ __align__(256) double g_noalloc_scratch[NA_PAR_DOUBLES + FEELGPU_NOALLOC_MAX_KQ];

// In host code g_noalloc_scratch is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current.
constexpr int NA_MAX_DEVICES = 64;
double* g_noalloc_ptr[NA_MAX_DEVICES] = {};

int noalloc_scratch(feelgpu_context_t ctx, double** ptr)
{
    int dev = -1;
    FEELGPU_CHECK_CUDA(ctx, cudaGetDevice(&dev));
    if (dev < 0 || dev >= NA_MAX_DEVICES) {
        ctx->last_error = "form2_assemble_noalloc: device ordinal outside [0,64)";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    double* p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        FEELGPU_CHECK_CUDA(ctx, cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) {
            ctx->last_error = "form2_assemble_noalloc: cudaGetSymbolAddress returned a null address";
            return FEELGPU_ERR_CUDA;
        }
        __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);   // the same value whichever thread gets here first
    }
    *ptr = p;
    return FEELGPU_OK;
}

template <int O, int F, int Q>
cudaError_t na_launch(const RArgs& A, int grid, int threads, size_t sh, cudaStream_t s)
{
    cudaError_t err = cudaFuncSetAttribute(k_rplan<O, F, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
    if (err != cudaSuccess) return err;
    k_rplan<O, F, Q><<<grid, threads, sh, s>>>(A);
    return cudaGetLastError();
}

template <int O, int F>
cudaError_t na_launch_q(bool fixed, int nq, const RArgs& A, int grid, int threads, size_t sh, cudaStream_t s)
{
    if (!fixed) return na_launch<O, F, 0>(A, grid, threads, sh, s);
    switch (nq) {
    case 1: return na_launch<O, F, 1>(A, grid, threads, sh, s);
    case 4: return na_launch<O, F, 4>(A, grid, threads, sh, s);
    case 5: return na_launch<O, F, 5>(A, grid, threads, sh, s);
    case 11: return na_launch<O, F, 11>(A, grid, threads, sh, s);
    case 14: return na_launch<O, F, 14>(A, grid, threads, sh, s);
    default: return na_launch<O, F, 24>(A, grid, threads, sh, s);
    }
}

} // namespace

bool feelgpu_rplan_noalloc_supports(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc)
{
    // the forms that use the scratch: mass / stiffness of a scalar space with a point-dependent coefficient
    if (desc->kind != FEELGPU_FORM2_MASS && desc->kind != FEELGPU_FORM2_STIFFNESS) return false;
    if (desc->coeff.kind != FEELGPU_COEFF_FIELD && desc->coeff.kind != FEELGPU_COEFF_EXPR) return false;
    return feelgpu_rplan_supports(pat, desc);
}

int feelgpu_launch_form2_rplan_noalloc(feelgpu_pattern_t pat, const FormParams& hp, const feelgpu_form2_desc* desc,
                                       int accumulate, double* a, cudaStream_t s)
{
    feelgpu_context_t ctx = pat->ctx;
    const RPlan* p = pat->rplan;
    feelgpu_space_t sp = pat->test;
    feelgpu_mesh_t m = sp->mesh;
    const int form = rplan_form(desc), n = p->n, nq = hp.nq, nf = hp.nf;
    const int64_t total = m->nElt * nq;
    if (total > (int64_t)FEELGPU_NOALLOC_MAX_KQ) {
        ctx->last_error = "form2_assemble_noalloc: elements x quadrature points = " + std::to_string(total) +
                          " exceeds the static scratch (FEELGPU_NOALLOC_MAX_KQ = " + std::to_string((int64_t)FEELGPU_NOALLOC_MAX_KQ) +
                          " doubles); use feelgpu_form2_assemble";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    if (n > FEELGPU_MAX_NLOC || nq > FEELGPU_MAX_NQ || nf > FEELGPU_MAX_NLOC) {
        ctx->last_error = "form2_assemble_noalloc: tables larger than the static parameter block";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    const int grid = (int)p->npatch;
    if (grid == 0) return FEELGPU_OK;
    double* scratch = nullptr;
    int rc = noalloc_scratch(ctx, &scratch);
    if (rc) return rc;
    double* d_kq = scratch + NA_PAR_DOUBLES;

    // parameter block + tables, staged on the stack (the copy from pageable memory has left the buffer when the call returns)
    alignas(16) double hb[sizeof(FormParams) / 8 + NA_TAB_DOUBLES];
    std::memcpy(hb, &hp, sizeof(FormParams));
    double* hphi = hb + sizeof(FormParams) / 8;
    double* hdphi = hphi + (size_t)n * nq;
    double* hphif = hdphi + (size_t)3 * n * nq;
    std::memcpy(hphi, desc->tables.phi, sizeof(double) * n * nq);
    if (desc->tables.dphi) std::memcpy(hdphi, desc->tables.dphi, sizeof(double) * 3 * n * nq);
    else std::memset(hdphi, 0, sizeof(double) * 3 * n * nq);
    if (nf) std::memcpy(hphif, desc->tables.phif, sizeof(double) * nf * nq);
    const size_t used = sizeof(FormParams) / 8 + (size_t)4 * n * nq + (size_t)nf * nq;
    FEELGPU_CHECK_CUDA(ctx, cudaMemcpyAsync(scratch, hb, used * sizeof(double), cudaMemcpyHostToDevice, s));

    RArgs A;
    A.P = reinterpret_cast<const FormParams*>(scratch);
    A.T.phi = scratch + sizeof(FormParams) / 8;
    A.T.dphi = A.T.phi + (size_t)n * nq;
    A.T.phif = A.T.dphi + (size_t)3 * n * nq;
    A.T.M9 = nullptr;                 // not read by the row-thread kernels
    A.T.Mm = nullptr;
    A.patches = p->d_patches;
    A.visits = p->d_visits;
    A.rows = reinterpret_cast<const uint4*>(p->d_rows);
    A.units = p->d_units;
    A.xyz = m->d_xyz;
    A.conn = m->d_conn;
    A.pos = pat->d_pos;
    A.l2gf = nullptr;
    A.fvals = nullptr;
    if (hp.coeff_kind == FEELGPU_COEFF_FIELD) {
        A.l2gf = desc->coeff.field_space->d_l2g;
        A.fvals = desc->coeff.field_values;
    }
    A.kq = d_kq;
    A.a = a;
    A.accumulate = accumulate;
    A.maxslots = p->maxslots;
    A.maxvis = p->maxvisits;
    A.maxrows = p->maxrows;
    const size_t sh = rplan_smem_bytes(p, form, nq, nf);
    const int threads = rplan_threads(form);
    const bool fixed = nq == 1 || nq == 4 || nq == 5 || nq == 11 || nq == 14 || nq == 24;
    if (fixed) {
        double ct[RTAB_MAX];
        for (int j = 0; j < n; ++j)
            for (int q = 0; q < nq; ++q) {
                if (form == 0) ct[(size_t)q * n + j] = desc->tables.phi[j * nq + q];
                else
                    for (int c = 0; c < 3; ++c) ct[((size_t)q * 3 + c) * n + j] = desc->tables.dphi[(j * nq + q) * 3 + c];
            }
        const size_t cnt = (size_t)(form == 0 ? n * nq : 3 * n * nq);
        FEELGPU_CHECK_CUDA(ctx, cudaMemcpyToSymbolAsync(c_rtab, ct, cnt * sizeof(double), 0, cudaMemcpyHostToDevice, s));
    }
    FEELGPU_CHECK_CUDA(ctx, cudaMemcpyToSymbolAsync(c_rw, desc->tables.w, nq * sizeof(double), 0, cudaMemcpyHostToDevice, s));
    FEELGPU_CHECK_CUDA(ctx, cudaMemcpyToSymbolAsync(c_rxi, desc->tables.xi, 3 * nq * sizeof(double), 0, cudaMemcpyHostToDevice, s));

    const int cgrid = (int)std::max<int64_t>(1, std::min<int64_t>((total + 255) / 256, (int64_t)ctx->sm_count * 32));
    k_rplan_coeff<<<cgrid, 256, 0, s>>>(A.P, m->nElt, nq, m->d_xyz, m->d_conn, A.T.phif, A.l2gf, A.fvals, d_kq);
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    cudaError_t err;
    switch (p->order * 2 + form) {
    case 2: err = na_launch_q<1, 0>(fixed, nq, A, grid, threads, sh, s); break;
    case 3: err = na_launch_q<1, 1>(fixed, nq, A, grid, threads, sh, s); break;
    case 4: err = na_launch_q<2, 0>(fixed, nq, A, grid, threads, sh, s); break;
    case 5: err = na_launch_q<2, 1>(fixed, nq, A, grid, threads, sh, s); break;
    case 6: err = na_launch_q<3, 0>(fixed, nq, A, grid, threads, sh, s); break;
    case 7: err = na_launch_q<3, 1>(fixed, nq, A, grid, threads, sh, s); break;
    default: ctx->last_error = "form2_assemble_noalloc: unsupported order / form"; return FEELGPU_ERR_UNSUPPORTED;
    }
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, err);
    return FEELGPU_OK;
}

// ---- host-side view of the plan (include/feelgpu.h): no device needed, used by the CPU tests ----------------------------
extern "C" {

int feelgpu_rplan_host_build(int order, int ncomp, int64_t nElt, const int32_t* l2g, int64_t nPts, const double* xyz,
                             const int32_t* conn, int64_t nRows, const int64_t* ia, int maxslots, int maxvisits, int maxrows,
                             void** plan, int64_t sizes[8])
{
    if (!l2g || !xyz || !conn || !ia || !plan || !sizes || order < 1 || order > 3) return FEELGPU_ERR_ARG;
    RPlanHost* hp = new RPlanHost;
    *plan = hp;
    const int n = (order + 1) * (order + 2) * (order + 3) / 6;
    const bool ok = rplan_build_host(order, ncomp, nElt, l2g, n * ncomp, nPts, xyz, conn, nRows, ia, maxslots, maxvisits, maxrows, 0, *hp);
    sizes[0] = (int64_t)hp->patches.size();
    sizes[1] = (int64_t)hp->visits.size();
    sizes[2] = (int64_t)hp->rows.size();
    sizes[3] = (int64_t)hp->units.size();
    sizes[4] = hp->maxpatch_slots;
    sizes[5] = hp->maxpatch_visits;
    sizes[6] = hp->maxpatch_rows;
    sizes[7] = 0;
    return ok ? FEELGPU_OK : FEELGPU_ERR_UNSUPPORTED;
}

int feelgpu_rplan_host_copy(void* plan, void* patches /*32 B each*/, int32_t* visits, void* rows /*32 B each*/, uint32_t* units)
{
    if (!plan) return FEELGPU_ERR_ARG;
    const RPlanHost* hp = static_cast<const RPlanHost*>(plan);
    if (patches) std::memcpy(patches, hp->patches.data(), hp->patches.size() * sizeof(RPlanPatch));
    if (visits) std::memcpy(visits, hp->visits.data(), hp->visits.size() * sizeof(int32_t));
    if (rows) std::memcpy(rows, hp->rows.data(), hp->rows.size() * sizeof(RPlanRow));
    if (units) std::memcpy(units, hp->units.data(), hp->units.size() * sizeof(uint32_t));
    return FEELGPU_OK;
}

const char* feelgpu_rplan_host_error(void* plan) { return plan ? static_cast<const RPlanHost*>(plan)->error.c_str() : ""; }

int feelgpu_rplan_host_free(void* plan)
{
    delete static_cast<RPlanHost*>(plan);
    return FEELGPU_OK;
}

} // extern "C"
