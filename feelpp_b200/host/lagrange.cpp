#include "lagrange.hpp"

#include <array>
#include <cmath>
#include <stdexcept>

#include "../../include/feelgpu_quadrature_tables.h"
#include "doftable.hpp"

namespace feelpp_b200 {

Quadrature im(int order)
{
    if (order < 0 || order > FEELGPU_TET_QUAD_MAX_ORDER) throw std::out_of_range("im: quadrature order not tabulated");
    Quadrature q;
    q.order = order;
    q.nq = feelgpu_tet_quad_npts[order];
    const double* t = feelgpu_tet_quad_rule[order];
    q.xi.resize(3 * q.nq);
    q.w.resize(q.nq);
    for (int k = 0; k < q.nq; ++k) {
        q.xi[3 * k + 0] = t[4 * k + 0];
        q.xi[3 * k + 1] = t[4 * k + 1];
        q.xi[3 * k + 2] = t[4 * k + 2];
        q.w[k] = t[4 * k + 3];
    }
    return q;
}

static const double kRefVertex[4][3] = {{-1, -1, -1}, {1, -1, -1}, {-1, 1, -1}, {-1, -1, 1}};

// Face rules: IMTriangle<order> carried onto face f of the reference tet by the face's boundary map (vertices in f2p
// order, feelmesh/simplexordering.hpp:341-347), weights times the map's area factor (feelpoly/pointsetquadrature.hpp:
// 696-760, identity permutation).  Result: 4 blocks of nq points, block f = local face f.
Quadrature imFaces(int order)
{
    if (order < 0 || order > FEELGPU_TET_QUAD_MAX_ORDER) throw std::out_of_range("imFaces: quadrature order not tabulated");
    static const int f2p[4][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}};
    Quadrature q;
    q.order = order;
    const int nqf = feelgpu_tri_quad_npts[order];
    const double* t = feelgpu_tri_quad_rule[order];
    q.nq = 4 * nqf;
    q.xi.resize(3 * q.nq);
    q.w.resize(q.nq);
    for (int f = 0; f < 4; ++f) {
        const double *a = kRefVertex[f2p[f][0]], *b = kRefVertex[f2p[f][1]], *c = kRefVertex[f2p[f][2]];
        double u[3], v[3];
        for (int d = 0; d < 3; ++d) { u[d] = 0.5 * (b[d] - a[d]); v[d] = 0.5 * (c[d] - a[d]); }
        const double uu = u[0] * u[0] + u[1] * u[1] + u[2] * u[2], uv = u[0] * v[0] + u[1] * v[1] + u[2] * v[2],
                     vv = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
        const double area = std::sqrt(uu * vv - uv * uv);
        for (int k = 0; k < nqf; ++k) {
            const double s = t[3 * k], r = t[3 * k + 1];
            const double l0 = -0.5 * (s + r), l1 = 0.5 * (1.0 + s), l2 = 0.5 * (1.0 + r);
            const int p = f * nqf + k;
            for (int d = 0; d < 3; ++d) q.xi[3 * p + d] = a[d] * l0 + b[d] * l1 + c[d] * l2;
            q.w[p] = t[3 * k + 2] * area;
        }
    }
    return q;
}

std::vector<double> lagrangePoints(int order)
{
    if (order < 1 || order > 3) throw std::invalid_argument("lagrangePoints: order must be 1, 2 or 3");
    std::vector<double> p;
    auto push = [&](double x, double y, double z) { p.push_back(x); p.push_back(y); p.push_back(z); };
    for (int v = 0; v < 4; ++v) push(kRefVertex[v][0], kRefVertex[v][1], kRefVertex[v][2]);
    std::vector<double> s;
    if (order == 2) s = {0.5};
    if (order == 3) {
        const double g = std::sqrt(0.2); // Gauss-Lobatto interior nodes of the 4-point rule: +-sqrt(1/5)
        s = {0.5 * (1.0 - g), 0.5 * (1.0 + g)};
    }
    for (int e = 0; e < 6; ++e)
        for (double t : s) {
            const double* a = kRefVertex[kEdgeToPoint[e][0]];
            const double* b = kRefVertex[kEdgeToPoint[e][1]];
            push(a[0] + t * (b[0] - a[0]), a[1] + t * (b[1] - a[1]), a[2] + t * (b[2] - a[2]));
        }
    if (order == 3)
        for (int f = 0; f < 4; ++f) {
            const double* a = kRefVertex[kFaceToPoint[f][0]];
            const double* b = kRefVertex[kFaceToPoint[f][1]];
            const double* c = kRefVertex[kFaceToPoint[f][2]];
            push((a[0] + b[0] + c[0]) / 3.0, (a[1] + b[1] + c[1]) / 3.0, (a[2] + b[2] + c[2]) / 3.0);
        }
    return p;
}

namespace {

// P1 / P2: closed forms in barycentric coordinates
//   lambda_0 = -(1+x+y+z)/2, lambda_a = (1+xi_{a-1})/2 ; d lambda_0/d xi = -1/2, d lambda_a / d xi_{a-1} = 1/2
void tabulateLow(int order, const double* xi, double* phi, double* dphi /*[n][3]*/)
{
    double l[4] = {-0.5 * (1.0 + xi[0] + xi[1] + xi[2]), 0.5 * (1.0 + xi[0]), 0.5 * (1.0 + xi[1]), 0.5 * (1.0 + xi[2])};
    double dl[4][3] = {{-0.5, -0.5, -0.5}, {0.5, 0, 0}, {0, 0.5, 0}, {0, 0, 0.5}};
    if (order == 1) {
        for (int i = 0; i < 4; ++i) {
            phi[i] = l[i];
            for (int c = 0; c < 3; ++c) dphi[3 * i + c] = dl[i][c];
        }
        return;
    }
    for (int i = 0; i < 4; ++i) {
        phi[i] = l[i] * (2.0 * l[i] - 1.0);
        for (int c = 0; c < 3; ++c) dphi[3 * i + c] = (4.0 * l[i] - 1.0) * dl[i][c];
    }
    for (int e = 0; e < 6; ++e) {
        const int a = kEdgeToPoint[e][0], b = kEdgeToPoint[e][1];
        phi[4 + e] = 4.0 * l[a] * l[b];
        for (int c = 0; c < 3; ++c) dphi[3 * (4 + e) + c] = 4.0 * (l[a] * dl[b][c] + l[b] * dl[a][c]);
    }
}

// P3: generalised Vandermonde in the Cartesian monomials x^a y^b z^c (a+b+c<=3), solved by LU in long double.
struct P3Basis {
    static constexpr int N = 20;
    int ex[N][3];
    long double coef[N][N]; // coef[i][m]: coefficient of monomial m in basis function i
    P3Basis()
    {
        int m = 0;
        for (int d = 0; d <= 3; ++d)
            for (int a = d; a >= 0; --a)
                for (int b = d - a; b >= 0; --b) {
                    ex[m][0] = a;
                    ex[m][1] = b;
                    ex[m][2] = d - a - b;
                    ++m;
                }
        std::vector<double> pts = lagrangePoints(3);
        long double V[N][N];
        for (int j = 0; j < N; ++j)
            for (int k = 0; k < N; ++k)
                V[j][k] = powl(pts[3 * j], ex[k][0]) * powl(pts[3 * j + 1], ex[k][1]) * powl(pts[3 * j + 2], ex[k][2]);
        // LU with partial pivoting, then solve V^T-systems: basis i satisfies sum_m coef[i][m] mono_m(p_j) = delta_ij
        int piv[N];
        for (int i = 0; i < N; ++i) piv[i] = i;
        for (int c = 0; c < N; ++c) {
            int p = c;
            for (int r = c + 1; r < N; ++r)
                if (fabsl(V[r][c]) > fabsl(V[p][c])) p = r;
            if (p != c) {
                for (int k = 0; k < N; ++k) std::swap(V[p][k], V[c][k]);
                std::swap(piv[p], piv[c]);
            }
            for (int r = c + 1; r < N; ++r) {
                V[r][c] /= V[c][c];
                for (int k = c + 1; k < N; ++k) V[r][k] -= V[r][c] * V[c][k];
            }
        }
        for (int i = 0; i < N; ++i) {
            long double y[N];
            for (int r = 0; r < N; ++r) y[r] = (piv[r] == i) ? 1.0L : 0.0L;
            for (int r = 0; r < N; ++r)
                for (int k = 0; k < r; ++k) y[r] -= V[r][k] * y[k];
            for (int r = N - 1; r >= 0; --r) {
                for (int k = r + 1; k < N; ++k) y[r] -= V[r][k] * y[k];
                y[r] /= V[r][r];
            }
            for (int m2 = 0; m2 < N; ++m2) coef[i][m2] = y[m2];
        }
    }
    void eval(const double* xi, double* phi, double* dphi) const
    {
        long double pw[3][4];
        for (int c = 0; c < 3; ++c) {
            pw[c][0] = 1;
            for (int k = 1; k < 4; ++k) pw[c][k] = pw[c][k - 1] * xi[c];
        }
        for (int i = 0; i < N; ++i) {
            long double v = 0, g[3] = {0, 0, 0};
            for (int m = 0; m < N; ++m) {
                const int a = ex[m][0], b = ex[m][1], c = ex[m][2];
                const long double cf = coef[i][m];
                v += cf * pw[0][a] * pw[1][b] * pw[2][c];
                if (a) g[0] += cf * a * pw[0][a - 1] * pw[1][b] * pw[2][c];
                if (b) g[1] += cf * b * pw[0][a] * pw[1][b - 1] * pw[2][c];
                if (c) g[2] += cf * c * pw[0][a] * pw[1][b] * pw[2][c - 1];
            }
            phi[i] = (double)v;
            for (int c = 0; c < 3; ++c) dphi[3 * i + c] = (double)g[c];
        }
    }
};

} // namespace

BasisTables tabulate(int order, int nq, const double* xi)
{
    if (order < 1 || order > 3) throw std::invalid_argument("tabulate: order must be 1, 2 or 3");
    BasisTables T;
    T.n = nLocalDofTet(order);
    T.nq = nq;
    T.phi.resize((size_t)T.n * nq);
    T.dphi.resize((size_t)T.n * nq * 3);
    std::vector<double> p(T.n), dp(3 * T.n);
    for (int q = 0; q < nq; ++q) {
        if (order <= 2) tabulateLow(order, xi + 3 * q, p.data(), dp.data());
        else {
            static const P3Basis B;
            B.eval(xi + 3 * q, p.data(), dp.data());
        }
        for (int i = 0; i < T.n; ++i) {
            T.phi[(size_t)i * nq + q] = p[i];
            for (int c = 0; c < 3; ++c) T.dphi[((size_t)i * nq + q) * 3 + c] = dp[3 * i + c];
        }
    }
    return T;
}

} // namespace feelpp_b200
