// device_common.cuh -- device-side pieces shared by the assembly kernels: the affine geomap of one tetrahedron
// (K = G*grad(N), J = |det K|, B = K^{-T}: feelpoly/geomap.hpp:2320-2355,1971-1991), xReal (geomap.hpp:1964-1969), the
// closed-form coefficient interpreter (feelvf/detail/ginacexvf.hpp:750-848) and the coefficient at a quadrature point.
#pragma once
#include "feelgpu_internal.h"

namespace {

struct DevTables {
    const double* phi;   // [n][nq]
    const double* dphi;  // [n][nq][3]
    const double* phif;  // [nf][nq]
    const double* M9;    // [n][n][9]  sum_q w_q dphi_i^b dphi_j^d
    const double* Mm;    // [n][n]     sum_q w_q phi_i phi_j
};

__device__ __forceinline__ void load_vertex(const double* __restrict__ xyz, int v, double* p)
{
    const double2* q = reinterpret_cast<const double2*>(xyz + 4 * (int64_t)v);
    double2 a = __ldg(q), b = __ldg(q + 1);
    p[0] = a.x;
    p[1] = a.y;
    p[2] = b.x;
}

// affine geomap of one element: K = G*grad(N), J = |det K|, B = K^{-T}
struct Geo {
    double B[3][3];
    double J;
    double x0[3], x1[3], x2[3], x3[3];
};

__device__ __forceinline__ void geo_update(const double* __restrict__ xyz, int4 c, Geo& g)
{
    load_vertex(xyz, c.x, g.x0);
    load_vertex(xyz, c.y, g.x1);
    load_vertex(xyz, c.z, g.x2);
    load_vertex(xyz, c.w, g.x3);
    double K[3][3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        K[a][0] = 0.5 * (g.x1[a] - g.x0[a]);
        K[a][1] = 0.5 * (g.x2[a] - g.x0[a]);
        K[a][2] = 0.5 * (g.x3[a] - g.x0[a]);
    }
    double c00 = K[1][1] * K[2][2] - K[1][2] * K[2][1];
    double c01 = K[1][2] * K[2][0] - K[1][0] * K[2][2];
    double c02 = K[1][0] * K[2][1] - K[1][1] * K[2][0];
    double det = K[0][0] * c00 + K[0][1] * c01 + K[0][2] * c02;
    double id = 1.0 / det;
    g.J = fabs(det);
    // B = K^{-T}: B[a][b] = inv[b][a] = cofactor(a,b)/det
    g.B[0][0] = c00 * id;
    g.B[0][1] = c01 * id;
    g.B[0][2] = c02 * id;
    g.B[1][0] = (K[0][2] * K[2][1] - K[0][1] * K[2][2]) * id;
    g.B[1][1] = (K[0][0] * K[2][2] - K[0][2] * K[2][0]) * id;
    g.B[1][2] = (K[0][1] * K[2][0] - K[0][0] * K[2][1]) * id;
    g.B[2][0] = (K[0][1] * K[1][2] - K[0][2] * K[1][1]) * id;
    g.B[2][1] = (K[0][2] * K[1][0] - K[0][0] * K[1][2]) * id;
    g.B[2][2] = (K[0][0] * K[1][1] - K[0][1] * K[1][0]) * id;
}

__device__ __forceinline__ void xreal(const Geo& g, const double* xi, double* x)
{
    double N0 = -0.5 * (1.0 + xi[0] + xi[1] + xi[2]), N1 = 0.5 * (1.0 + xi[0]), N2 = 0.5 * (1.0 + xi[1]),
           N3 = 0.5 * (1.0 + xi[2]);
#pragma unroll
    for (int a = 0; a < 3; ++a) x[a] = g.x0[a] * N0 + g.x1[a] * N1 + g.x2[a] * N2 + g.x3[a] * N3;
}

__device__ double eval_expr(const FormParams* P, double x, double y, double z)
{
    double st[16];
    int sp = 0;
    for (int k = 0; k < P->expr_len; ++k) {
        switch (P->expr_ops[k]) {
        case FEELGPU_OP_X: st[sp++] = x; break;
        case FEELGPU_OP_Y: st[sp++] = y; break;
        case FEELGPU_OP_Z: st[sp++] = z; break;
        case FEELGPU_OP_CONST: st[sp++] = P->expr_consts[k]; break;
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
        case FEELGPU_OP_POW: --sp; st[sp - 1] = pow(st[sp - 1], st[sp]); break;
        default: break;
        }
    }
    return st[0];
}

// coefficient at quadrature point q of element e (cst / idv(field) / expr at xReal)
__device__ __forceinline__ double coeff_at(const FormParams* P, const double* s_phif, const Geo& g, int64_t e, int q,
                                           const int32_t* __restrict__ l2gf, const double* __restrict__ fvals)
{
    if (P->coeff_kind == FEELGPU_COEFF_CONST) return P->cval[0];
    if (P->coeff_kind == FEELGPU_COEFF_FIELD) {
        double v = 0;
        for (int i = 0; i < P->nf; ++i) v += fvals[l2gf[e * P->nf + i]] * s_phif[i * P->nq + q];
        return v;
    }
    double x[3];
    xreal(g, &P->xi[3 * q], x);
    return eval_expr(P, x[0], x[1], x[2]);
}


} // namespace
