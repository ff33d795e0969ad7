// feel_oracle.cpp -- CPU restatement of the reference's assembly path.  TEST INFRASTRUCTURE ONLY (see feel_oracle.h).
//
// Every function cites the reference file:line (relative to /root/reference/feelpp/feel/) whose algorithm it
// restates.  Nothing here is copied from the reference; the quadrature literals come from the generated data
// header include/feelgpu_quadrature_tables.h (tools/gen_quadrature_tables.py).
//
// Parity status: "parity unpinned" (no golden matrices exist in the reference test-suite and the reference cannot
// be built here).  Pinned by invariants and by exactly integrated known answers (barycentric monomials) in
// tests/test_oracle_invariants.py.
#include "feel_oracle.h"
#include "../include/feelgpu_quadrature_tables.h"

#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <thread>
#include <tuple>
#include <vector>

namespace {

// ---------------------------------------------------------------------------------------------------------------
// Reference element: tet (-1,-1,-1),(1,-1,-1),(-1,1,-1),(-1,-1,1)  (feelmesh/refsimplex.hpp:151-162)
// local entity numbering (feelmesh/simplexordering.hpp:283-291, 341-347)
// ---------------------------------------------------------------------------------------------------------------
const int E2P[6][2] = {{1, 2}, {2, 0}, {0, 1}, {0, 3}, {1, 3}, {2, 3}};
const int F2P[4][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}};
const double REFV[4][3] = {{-1, -1, -1}, {1, -1, -1}, {-1, 1, -1}, {-1, -1, 1}};

int nloc(int order) { return (order + 1) * (order + 2) * (order + 3) / 6; }

// barycentric coordinates of a reference point: lambda_0 = -(1+x+y+z)/2, lambda_{a+1} = (1+xi_a)/2
void bary(const double* xi, long double* l)
{
    l[1] = (1.0L + xi[0]) / 2;
    l[2] = (1.0L + xi[1]) / 2;
    l[3] = (1.0L + xi[2]) / 2;
    l[0] = 1.0L - l[1] - l[2] - l[3];
}

// Nodal points of fem::Lagrange<3,k,Scalar,Continuous,double,Simplex,PointSetFekete>:
//   k<=2 : equispaced (feelpoly/equispaced.hpp:99-140); k==3 : warp&blend, which moves only the edge-interior points
//   (feelpoly/warpblend.hpp:96-121, max_entity_dim=1 for Order==3) onto the Gauss-Lobatto points of the edge.
// Order of points: vertices, then edge-interior points edge by edge (directed from the first to the second vertex
// of simplexordering.hpp:283-291), then face-interior points face by face, then volume points.
void lagrange_points(int order, std::vector<std::array<double, 3>>& pts)
{
    pts.clear();
    for (int v = 0; v < 4; ++v) pts.push_back({REFV[v][0], REFV[v][1], REFV[v][2]});
    if (order == 1) return;
    std::vector<double> s; // parametric positions in (0,1) of the edge-interior points
    if (order == 2) s = {0.5};
    else if (order == 3) {
        const double g = 1.0 / std::sqrt(5.0); // interior Gauss-Lobatto-Legendre points of 4-point rule: +-1/sqrt(5)
        s = {(1.0 - g) / 2, (1.0 + g) / 2};
    }
    for (int e = 0; e < 6; ++e)
        for (double t : s) {
            std::array<double, 3> p;
            for (int c = 0; c < 3; ++c) p[c] = (1.0 - t) * REFV[E2P[e][0]][c] + t * REFV[E2P[e][1]][c];
            pts.push_back(p);
        }
    if (order == 3)
        for (int f = 0; f < 4; ++f) {
            std::array<double, 3> p;
            for (int c = 0; c < 3; ++c) p[c] = (REFV[F2P[f][0]][c] + REFV[F2P[f][1]][c] + REFV[F2P[f][2]][c]) / 3.0;
            pts.push_back(p);
        }
}

// exponents of the barycentric monomial basis lambda^alpha, |alpha| = k
void exponents(int k, std::vector<std::array<int, 4>>& al)
{
    al.clear();
    for (int a = 0; a <= k; ++a)
        for (int b = 0; a + b <= k; ++b)
            for (int c = 0; a + b + c <= k; ++c) al.push_back({a, b, c, k - a - b - c});
}

long double ipow(long double x, int p)
{
    long double r = 1;
    for (int i = 0; i < p; ++i) r *= x;
    return r;
}

// The nodal (Lagrange) basis is the dual of the point evaluations on P_k (feelpoly/lagrange.hpp:65-300 builds it as
// the dual of the orthonormal Dubiner set; it is the unique nodal basis, so any well-conditioned construction
// agrees to rounding).  Here: generalised Vandermonde in the barycentric monomials, inverted in long double.
struct Basis {
    int order = 0, n = 0;
    std::vector<std::array<int, 4>> al;
    std::vector<long double> C; // C[m*n+i]: coefficient of monomial m in basis function i
    void build(int k)
    {
        order = k;
        n = nloc(k);
        exponents(k, al);
        std::vector<std::array<double, 3>> pts;
        lagrange_points(k, pts);
        // V[j][m] = monomial m at point j ; solve V C = I
        std::vector<long double> V(n * n), I(n * n, 0);
        for (int j = 0; j < n; ++j) {
            long double l[4];
            bary(pts[j].data(), l);
            for (int m = 0; m < n; ++m)
                V[j * n + m] = ipow(l[0], al[m][0]) * ipow(l[1], al[m][1]) * ipow(l[2], al[m][2]) * ipow(l[3], al[m][3]);
            I[j * n + j] = 1;
        }
        // Gauss-Jordan with partial pivoting
        for (int c = 0; c < n; ++c) {
            int p = c;
            for (int r = c + 1; r < n; ++r)
                if (fabsl(V[r * n + c]) > fabsl(V[p * n + c])) p = r;
            if (p != c)
                for (int m = 0; m < n; ++m) {
                    std::swap(V[p * n + m], V[c * n + m]);
                    std::swap(I[p * n + m], I[c * n + m]);
                }
            long double d = V[c * n + c];
            for (int m = 0; m < n; ++m) {
                V[c * n + m] /= d;
                I[c * n + m] /= d;
            }
            for (int r = 0; r < n; ++r)
                if (r != c) {
                    long double f = V[r * n + c];
                    if (f != 0)
                        for (int m = 0; m < n; ++m) {
                            V[r * n + m] -= f * V[c * n + m];
                            I[r * n + m] -= f * I[c * n + m];
                        }
                }
        }
        C = I; // C[m*n + i]
    }
    // value and reference gradient (d/dxi_a) of basis i at xi
    void eval(const double* xi, double* phi, double* dphi /*[n][3]*/) const
    {
        long double l[4];
        bary(xi, l);
        std::vector<long double> mv(n), md(n * 4);
        for (int m = 0; m < n; ++m) {
            long double p[4], dp[4];
            for (int k = 0; k < 4; ++k) {
                p[k] = ipow(l[k], al[m][k]);
                dp[k] = al[m][k] > 0 ? al[m][k] * ipow(l[k], al[m][k] - 1) : 0.0L;
            }
            mv[m] = p[0] * p[1] * p[2] * p[3];
            md[m * 4 + 0] = dp[0] * p[1] * p[2] * p[3];
            md[m * 4 + 1] = p[0] * dp[1] * p[2] * p[3];
            md[m * 4 + 2] = p[0] * p[1] * dp[2] * p[3];
            md[m * 4 + 3] = p[0] * p[1] * p[2] * dp[3];
        }
        for (int i = 0; i < n; ++i) {
            long double v = 0, g[4] = {0, 0, 0, 0};
            for (int m = 0; m < n; ++m) {
                long double c = C[m * n + i];
                v += c * mv[m];
                for (int k = 0; k < 4; ++k) g[k] += c * md[m * 4 + k];
            }
            phi[i] = (double)v;
            // d lambda_0/d xi_a = -1/2, d lambda_{a+1}/d xi_a = 1/2
            for (int a = 0; a < 3; ++a) dphi[i * 3 + a] = (double)((g[a + 1] - g[0]) / 2);
        }
    }
};

const Basis& basis(int order)
{
    static Basis B[4];
    static bool done[4] = {false, false, false, false};
    if (!done[order]) {
        B[order].build(order);
        done[order] = true;
    }
    return B[order];
}

// PreCompute tables at the quadrature points (feelpoly/polynomialset.hpp:1052-1083): phi[n][Q], dphi[n][Q][3]
struct Tables {
    int n, Q;
    std::vector<double> w, xi, phi, dphi;
};

Tables make_tables(int order, int quad_order)
{
    Tables T;
    T.n = nloc(order);
    T.Q = feelgpu_tet_quad_npts[quad_order];
    const double* q = feelgpu_tet_quad_rule[quad_order];
    T.w.resize(T.Q);
    T.xi.resize(3 * T.Q);
    T.phi.resize(T.n * T.Q);
    T.dphi.resize(T.n * T.Q * 3);
    std::vector<double> p(T.n), dp(T.n * 3);
    for (int k = 0; k < T.Q; ++k) {
        T.w[k] = q[4 * k + 3];
        for (int c = 0; c < 3; ++c) T.xi[3 * k + c] = q[4 * k + c];
        basis(order).eval(&q[4 * k], p.data(), dp.data());
        for (int i = 0; i < T.n; ++i) {
            T.phi[i * T.Q + k] = p[i];
            for (int c = 0; c < 3; ++c) T.dphi[(i * T.Q + k) * 3 + c] = dp[i * 3 + c];
        }
    }
    return T;
}

// postfix coefficient expression (stands in for the GiNaC-compiled `expr("...")`, feelvf/detail/ginacexvf.hpp:750-848)
double eval_expr(const oracle_coeff* c, double x, double y, double z)
{
    double st[32];
    int sp = 0;
    for (int k = 0; k < c->expr_len; ++k) {
        switch (c->expr_ops[k]) {
        case 0: st[sp++] = x; break;
        case 1: st[sp++] = y; break;
        case 2: st[sp++] = z; break;
        case 3: st[sp++] = c->expr_consts[k]; break;
        case 4: --sp; st[sp - 1] += st[sp]; break;
        case 5: --sp; st[sp - 1] -= st[sp]; break;
        case 6: --sp; st[sp - 1] *= st[sp]; break;
        case 7: --sp; st[sp - 1] /= st[sp]; break;
        case 8: st[sp - 1] = -st[sp - 1]; break;
        case 9: st[sp - 1] = std::sin(st[sp - 1]); break;
        case 10: st[sp - 1] = std::cos(st[sp - 1]); break;
        case 11: st[sp - 1] = std::exp(st[sp - 1]); break;
        case 12: st[sp - 1] = std::log(st[sp - 1]); break;
        case 13: st[sp - 1] = std::sqrt(st[sp - 1]); break;
        case 14: --sp; st[sp - 1] = std::pow(st[sp - 1], st[sp]); break;
        default: return NAN;
        }
    }
    return st[0];
}

// Affine geomap context of one element (feelpoly/geomap.hpp:2320-2355: K = G * grad(N); :1971-1976 J = |det K|;
// :1985-1991 B = K^{-T}); one point per element for order-1 simplices (geomap.hpp:1142-1143).
struct Geo {
    double K[3][3], B[3][3], J;
    double x0[3]; // image of the reference origin-vertex, for xReal = G*N(xi) (geomap.hpp:1964-1969)
    const double* G[4];
};

inline void geo_update(const double* const G[4], Geo& g)
{
    // geometric P1 basis gradients on the reference tet: grad N_0 = (-1/2,-1/2,-1/2), grad N_{b+1} = e_b/2
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) g.K[a][b] = -0.5 * G[0][a] + 0.5 * G[b + 1][a];
    const double(*K)[3] = g.K;
    double c00 = K[1][1] * K[2][2] - K[1][2] * K[2][1];
    double c01 = K[1][2] * K[2][0] - K[1][0] * K[2][2];
    double c02 = K[1][0] * K[2][1] - K[1][1] * K[2][0];
    double det = K[0][0] * c00 + K[0][1] * c01 + K[0][2] * c02;
    g.J = std::fabs(det);
    double inv[3][3]; // K^{-1}
    inv[0][0] = c00 / det;
    inv[1][0] = c01 / det;
    inv[2][0] = c02 / det;
    inv[0][1] = (K[0][2] * K[2][1] - K[0][1] * K[2][2]) / det;
    inv[1][1] = (K[0][0] * K[2][2] - K[0][2] * K[2][0]) / det;
    inv[2][1] = (K[0][1] * K[2][0] - K[0][0] * K[2][1]) / det;
    inv[0][2] = (K[0][1] * K[1][2] - K[0][2] * K[1][1]) / det;
    inv[1][2] = (K[0][2] * K[1][0] - K[0][0] * K[1][2]) / det;
    inv[2][2] = (K[0][0] * K[1][1] - K[0][1] * K[1][0]) / det;
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) g.B[a][b] = inv[b][a];
    for (int a = 0; a < 4; ++a) g.G[a] = G[a];
}

inline void geo_xreal(const Geo& g, const double* xi, double* x)
{
    double N[4] = {-(1 + xi[0] + xi[1] + xi[2]) / 2, (1 + xi[0]) / 2, (1 + xi[1]) / 2, (1 + xi[2]) / 2};
    for (int a = 0; a < 3; ++a) x[a] = g.G[0][a] * N[0] + g.G[1][a] * N[1] + g.G[2][a] * N[2] + g.G[3][a] * N[3];
}

// coefficient value at quadrature point q of element e
inline double coeff_at(const oracle_coeff* c, const Geo& g, const Tables& T, const Tables* Tf, int64_t e, int q)
{
    if (!c || c->kind == ORACLE_COEFF_CONST) return c ? c->value[0] : 1.0;
    if (c->kind == ORACLE_COEFF_FIELD) {
        // idv(k): sum_i k[l2g(e,i)] * phi_i(xi_q)   (feelvf operators.hpp idv)
        double v = 0;
        for (int i = 0; i < Tf->n; ++i) v += c->field_values[c->field_l2g[e * Tf->n + i]] * Tf->phi[i * Tf->Q + q];
        return v;
    }
    double x[3];
    geo_xreal(g, &T.xi[3 * q], x);
    return eval_expr(c, x[0], x[1], x[2]);
}

// One element: update + integrate (feelvf/bilinearformcontext.hpp:328-374, 455-628).  Ae row-major nd x nd,
// rows = test dofs (bilinearform.hpp:462-464).  Vector spaces: local index I = n*c + i (doftable.hpp:1289-1292).
void local_matrix(int kind, int ncomp, const Tables& T, const Tables* Tf, const oracle_coeff* c, const Geo& g,
                  int64_t e, double* Ae, double* gradbuf /* n*Q*3 */, double* prod /* Q */, double* kq /* Q */)
{
    const int n = T.n, Q = T.Q, nd = n * ncomp;
    // M_integrator.update: prod[q] = w[q] * J   (feelpoly/pointsetquadrature.hpp:481-495)
    for (int q = 0; q < Q; ++q) prod[q] = T.w[q] * g.J;
    const bool need_k = (kind != ORACLE_FORM2_ELASTICITY);
    if (need_k)
        for (int q = 0; q < Q; ++q) kq[q] = coeff_at(c, g, T, Tf, e, q);
    if (kind != ORACLE_FORM2_MASS) {
        // updateGrad: grad phi_i(q) = grad_hat phi_i(q) * B^T   (feelpoly/polynomialset_impl.hpp:274-305)
        for (int i = 0; i < n; ++i)
            for (int q = 0; q < Q; ++q) {
                const double* gh = &T.dphi[(i * Q + q) * 3];
                for (int a = 0; a < 3; ++a)
                    gradbuf[(i * Q + q) * 3 + a] = g.B[a][0] * gh[0] + g.B[a][1] * gh[1] + g.B[a][2] * gh[2];
            }
    }
    // M_rep(i,j) = sum_q prod[q] * expr.evalijq(i,j,q)   (bilinearformcontext.hpp:585-589, pointsetquadrature.hpp:403-416)
    for (int I = 0; I < nd; ++I) {
        const int c1 = I / n, i = I % n;
        for (int Jd = 0; Jd < nd; ++Jd) {
            const int c2 = Jd / n, j = Jd % n;
            double s = 0;
            if (kind == ORACLE_FORM2_MASS) {
                if (c1 == c2)
                    for (int q = 0; q < Q; ++q) s += prod[q] * (kq[q] * (T.phi[j * Q + q] * T.phi[i * Q + q]));
            } else if (kind == ORACLE_FORM2_STIFFNESS) {
                if (c1 == c2)
                    for (int q = 0; q < Q; ++q) {
                        const double* gi = &gradbuf[(i * Q + q) * 3];
                        const double* gj = &gradbuf[(j * Q + q) * 3];
                        s += prod[q] * (kq[q] * (gj[0] * gi[0] + gj[1] * gi[1] + gj[2] * gi[2]));
                    }
            } else { // inner( lambda*trace(sym(gradt(u)))*Id + 2*mu*sym(gradt(u)), grad(v) )
                for (int q = 0; q < Q; ++q) {
                    const double* gi = &gradbuf[(i * Q + q) * 3];
                    const double* gj = &gradbuf[(j * Q + q) * 3];
                    double f = c->lambda * gj[c2] * gi[c1] + c->mu * gj[c1] * gi[c2];
                    if (c1 == c2) f += c->mu * (gj[0] * gi[0] + gj[1] * gi[1] + gj[2] * gi[2]);
                    s += prod[q] * f;
                }
            }
            Ae[I * nd + Jd] = s;
        }
    }
}

// MatSetValues(ADD_VALUES) on a preallocated sorted CSR: per-row binary search (feelalg/matrixpetsc.cpp:917-936)
inline int64_t find_slot(const int64_t* ia, const int64_t* ja, int64_t r, int64_t col)
{
    const int64_t* b = ja + ia[r];
    const int64_t* e = ja + ia[r + 1];
    const int64_t* p = std::lower_bound(b, e, col);
    if (p == e || *p != col) return -1;
    return p - ja;
}

struct Work {
    std::vector<double> Ae, grad, prod, kq;
    Work(int nd, int n, int Q) : Ae((size_t)nd * nd), grad((size_t)n * Q * 3), prod(Q), kq(Q) {}
};

} // namespace

extern "C" {

int oracle_quadrature(int order, double* q)
{
    if (order < 0 || order > FEELGPU_TET_QUAD_MAX_ORDER) return -1;
    int n = feelgpu_tet_quad_npts[order];
    if (q) std::memcpy(q, feelgpu_tet_quad_rule[order], sizeof(double) * 4 * n);
    return n;
}

int oracle_nlocaldof(int order) { return nloc(order); }

void oracle_lagrange_points(int order, double* pts)
{
    std::vector<std::array<double, 3>> p;
    lagrange_points(order, p);
    for (size_t i = 0; i < p.size(); ++i)
        for (int c = 0; c < 3; ++c) pts[3 * i + c] = p[i][c];
}

void oracle_tabulate(int order, int Q, const double* xi, double* phi, double* dphi)
{
    const Basis& B = basis(order);
    std::vector<double> p(B.n), dp(B.n * 3);
    for (int k = 0; k < Q; ++k) {
        B.eval(&xi[3 * k], p.data(), dp.data());
        for (int i = 0; i < B.n; ++i) {
            phi[i * Q + k] = p[i];
            for (int c = 0; c < 3; ++c) dphi[(i * Q + k) * 3 + c] = dp[i * 3 + c];
        }
    }
}

// ExpressionOrder (feelvf/expr.hpp:1046-1079): order = polyOrder * nOrderGeo (nOrderGeo = 1), or nOrderGeo when 0.
// Operator orders (feelvf/operators.hpp:134-153): id -> k, grad -> k-1; product = sum (operations.hpp:437-442);
// inner = sum (inner.hpp:154); GiNaC string: total degree if polynomial, else 2 (detail/ginacexvf.hpp:836-845);
// idv(P_m field) -> m; cst -> 0.  Checked by testsuite/feelintegration/test_auto_order_integrate.cpp:169-211.
int oracle_quad_order_form2(int kind, int order, int coeff_deg)
{
    int cd = coeff_deg < 0 ? 2 : coeff_deg;
    int p = (kind == ORACLE_FORM2_MASS) ? 2 * order : 2 * (order - 1);
    if (kind != ORACLE_FORM2_ELASTICITY) p += cd;
    return p == 0 ? 1 : p;
}
int oracle_quad_order_form1(int order, int coeff_deg)
{
    int cd = coeff_deg < 0 ? 2 : coeff_deg;
    int p = order + cd;
    return p == 0 ? 1 : p;
}

// DofTable::build (feeldiscr/doftable.hpp:1802-2070) for one process: elements in mesh order, per element
// vertices -> edges -> faces -> volume (feeldiscr/doffromelement.hpp:492-530); insertDof gives the next free index to
// a (entity type, entity id * nDofPerEntity + l) key at first touch (doftable.hpp:1233-1241); product spaces:
// local = n*c + l, global = base + c, next_free += ncomp (doftable.hpp:1289-1299, 1238-1241).
// Edge dofs are reversed when pointId(eToP(e,0)) > pointId(eToP(e,1)) (feelmesh/mesh3d.hpp:755-782,
// doffromelement.hpp:300-318).  Entity ids themselves do not influence the numbering (only identity does), so edges
// and faces are keyed by their sorted vertex ids.
int64_t oracle_doftable(int64_t nPts, int64_t nElt, const int32_t* conn, int order, int ncomp, int32_t* l2g)
{
    (void)nPts;
    const int n = nloc(order), nd = n * ncomp;
    const int nPerEdge = order - 1;
    const int nPerFace = (order - 1) * (order - 2) / 2;
    const int nPerVol = (order - 1) * (order - 2) * (order - 3) / 6;
    typedef std::tuple<int, int64_t, int64_t, int64_t, int> key_t; // (entity dim, sorted vertex ids..., l)
    std::map<key_t, int64_t> map_gdof;
    int64_t next_free = 0;
    auto insert = [&](int64_t e, int ldof, const key_t& key) {
        auto it = map_gdof.find(key);
        if (it == map_gdof.end()) {
            it = map_gdof.emplace(key, next_free).first;
            next_free += ncomp;
        }
        for (int c = 0; c < ncomp; ++c) l2g[e * nd + n * c + ldof] = (int32_t)(it->second + c);
    };
    for (int64_t e = 0; e < nElt; ++e) {
        const int32_t* v = conn + 4 * e;
        int lc = 0;
        for (int i = 0; i < 4; ++i, ++lc) insert(e, lc, key_t(0, v[i], -1, -1, 0));
        if (nPerEdge > 0)
            for (int ed = 0; ed < 6; ++ed) {
                int64_t i1 = v[E2P[ed][0]], i2 = v[E2P[ed][1]];
                bool reverse = i1 > i2;
                for (int l = 0; l < nPerEdge; ++l, ++lc) {
                    int gl = reverse ? nPerEdge - 1 - l : l;
                    insert(e, lc, key_t(1, std::min(i1, i2), std::max(i1, i2), -1, gl));
                }
            }
        if (nPerFace > 0)
            for (int f = 0; f < 4; ++f) {
                std::array<int64_t, 3> fv = {v[F2P[f][0]], v[F2P[f][1]], v[F2P[f][2]]};
                std::sort(fv.begin(), fv.end());
                // one dof per face at order 3: no permutation needed (doffromelement.hpp:430-434)
                for (int l = 0; l < nPerFace; ++l, ++lc) insert(e, lc, key_t(2, fv[0], fv[1], fv[2], l));
            }
        for (int l = 0; l < nPerVol; ++l, ++lc) insert(e, lc, key_t(3, e, -1, -1, l));
    }
    return next_free;
}

// ---------------------------------------------------------------------------------------------------------------
// DofTable / DataMap / GraphCSR in MPI mode, restated for one rank of a partitioned mesh (test infrastructure: pins the
// partitioned numbering and pattern of feelpp_b200/parallel.py).
//   * every rank builds its table by first touch over its ACTIVE elements in local element order (same loop as above,
//     feeldiscr/doftable.hpp:1802-1940);
//   * a dof is a ghost on rank p when a lower rank has an element containing its entity: procMin over the ranks sharing
//     the point / edge / face (feeldiscr/doftablempi.hpp:138-200 updateDofOnVertices/Edges, :548 "IdProcessOfGhost<myRank");
//   * M_first_df_globalcluster = prefix sums of the active counts (doftablempi.hpp:593-621); active dofs get
//     firstGlobIndex + running index in local order (:623-635); ghosts receive their owner's cluster id;
//   * the local ids are then reordered: active dofs before, ghost dofs after, each group keeping its order
//     (feeldiscr/doftable.hpp:1975-1992).
// ---------------------------------------------------------------------------------------------------------------
namespace {
typedef std::tuple<int, int64_t, int64_t, int64_t, int> ekey_t;   // (entity dim, sorted vertex ids..., l)

void element_keys(const int32_t* v, int64_t e, int order, std::vector<ekey_t>& out)
{
    const int nPerEdge = order - 1, nPerFace = (order - 1) * (order - 2) / 2, nPerVol = (order - 1) * (order - 2) * (order - 3) / 6;
    out.clear();
    for (int i = 0; i < 4; ++i) out.emplace_back(0, v[i], -1, -1, 0);
    for (int ed = 0; ed < 6 && nPerEdge > 0; ++ed) {
        const int64_t i1 = v[E2P[ed][0]], i2 = v[E2P[ed][1]];
        for (int l = 0; l < nPerEdge; ++l) out.emplace_back(1, std::min(i1, i2), std::max(i1, i2), -1, i1 > i2 ? nPerEdge - 1 - l : l);
    }
    for (int f = 0; f < 4 && nPerFace > 0; ++f) {
        std::array<int64_t, 3> fv = {v[F2P[f][0]], v[F2P[f][1]], v[F2P[f][2]]};
        std::sort(fv.begin(), fv.end());
        for (int l = 0; l < nPerFace; ++l) out.emplace_back(2, fv[0], fv[1], fv[2], l);
    }
    for (int l = 0; l < nPerVol; ++l) out.emplace_back(3, e, -1, -1, l);
}

struct MpiTables {
    std::map<ekey_t, int> owner;                         // procMin
    std::vector<std::vector<ekey_t>> touch;              // per rank: entities in first-touch order
    std::vector<std::map<ekey_t, int64_t>> active_idx;   // per rank: running index among its active entities
    std::vector<int64_t> first;                          // per rank: first cluster NODE id
};

void mpi_tables(int64_t nElt, const int32_t* conn, const int32_t* part, int nparts, int order, MpiTables& M)
{
    std::vector<ekey_t> keys;
    M.touch.assign(nparts, {});
    M.active_idx.assign(nparts, {});
    std::vector<std::set<ekey_t>> seen(nparts);
    for (int64_t e = 0; e < nElt; ++e) {
        const int p = part[e];
        element_keys(conn + 4 * e, e, order, keys);
        for (const ekey_t& k : keys) {
            auto it = M.owner.find(k);
            if (it == M.owner.end()) M.owner.emplace(k, p);
            else if (p < it->second) it->second = p;
            if (seen[p].insert(k).second) M.touch[p].push_back(k);
        }
    }
    M.first.assign(nparts + 1, 0);
    for (int p = 0; p < nparts; ++p) {
        int64_t run = 0;
        for (const ekey_t& k : M.touch[p])
            if (M.owner[k] == p) M.active_idx[p].emplace(k, run++);
        M.first[p + 1] = M.first[p] + run;
    }
}
} // namespace

extern "C" {

// local table of `rank`: l2g[nMine][nd] (elements of the rank in global element order; local ids: active first, ghosts
// after), p2c[nWithGhost] global-cluster ids, ghost_owner[nWithGhost - nOwned] (dof level), first_dof[nparts+1].
// Arrays sized by the caller (nMine*nd bounds the dofs).  Returns nWithGhost.
int64_t oracle_doftable_mpi(int64_t nElt, const int32_t* conn, const int32_t* part, int nparts, int rank, int order, int ncomp,
                            int32_t* l2g, int64_t* nOwned, int64_t* p2c, int32_t* ghost_owner, int64_t* first_dof)
{
    MpiTables M;
    mpi_tables(nElt, conn, part, nparts, order, M);
    const int n = nloc(order), nd = n * ncomp;
    for (int p = 0; p <= nparts; ++p) first_dof[p] = M.first[p] * ncomp;
    std::map<ekey_t, int64_t> local;                     // entity -> local node id
    int64_t nact = 0, ngh = 0;
    for (const ekey_t& k : M.touch[rank])
        if (M.owner[k] == rank) local[k] = nact++;
    for (const ekey_t& k : M.touch[rank])
        if (M.owner[k] != rank) {
            const int64_t id = nact + ngh++;
            local[k] = id;
            const int q = M.owner[k];
            for (int c = 0; c < ncomp; ++c) {
                p2c[id * ncomp + c] = (M.first[q] + M.active_idx[q][k]) * ncomp + c;
                ghost_owner[(id - nact) * ncomp + c] = q;
            }
        }
    for (const auto& kv : M.active_idx[rank])
        for (int c = 0; c < ncomp; ++c) p2c[kv.second * ncomp + c] = (M.first[rank] + kv.second) * ncomp + c;
    std::vector<ekey_t> keys;
    int64_t m = 0;
    for (int64_t e = 0; e < nElt; ++e) {
        if (part[e] != rank) continue;
        element_keys(conn + 4 * e, e, order, keys);
        for (int i = 0; i < n; ++i)
            for (int c = 0; c < ncomp; ++c) l2g[m * nd + n * c + i] = (int32_t)(local[keys[i]] * ncomp + c);
        ++m;
    }
    *nOwned = nact * ncomp;
    return (nact + ngh) * ncomp;
}

// local graph of `rank` (Stencil::computeGraph + GraphCSR::close in MPI mode, feelalg/graphcsr.cpp:800-951): rows = local
// dofs; an OWNED row holds the columns of every element of the mesh containing its entity (the neighbours' ghost rows were
// shipped to it), a GHOST row those of the rank's own elements; columns = global-cluster ids, ascending.
int64_t oracle_graph_mpi(int64_t nElt, const int32_t* conn, const int32_t* part, int nparts, int rank, int order, int ncomp,
                         int64_t nWithGhost, const int32_t* l2g, int64_t* ia, int64_t** ja_out)
{
    MpiTables M;
    mpi_tables(nElt, conn, part, nparts, order, M);
    const int n = nloc(order), nd = n * ncomp;
    std::map<ekey_t, int64_t> local;                     // as above, rebuilt from the rank's l2g
    std::vector<ekey_t> keys;
    {
        int64_t m = 0;
        for (int64_t e = 0; e < nElt; ++e) {
            if (part[e] != rank) continue;
            element_keys(conn + 4 * e, e, order, keys);
            for (int i = 0; i < n; ++i) local[keys[i]] = l2g[m * nd + i] / ncomp;
            ++m;
        }
    }
    std::vector<std::set<int64_t>> rows((size_t)nWithGhost);
    std::vector<int64_t> cols(nd);
    for (int64_t e = 0; e < nElt; ++e) {
        element_keys(conn + 4 * e, e, order, keys);
        for (int j = 0; j < n; ++j) {
            const int q = M.owner[keys[j]];
            for (int c = 0; c < ncomp; ++c) cols[n * c + j] = (M.first[q] + M.active_idx[q][keys[j]]) * ncomp + c;
        }
        for (int i = 0; i < n; ++i) {
            auto it = local.find(keys[i]);
            if (it == local.end()) continue;
            const bool owned = M.owner[keys[i]] == rank;
            if (!owned && part[e] != rank) continue;     // ghost rows only carry the rank's own contributions
            for (int c = 0; c < ncomp; ++c) rows[(size_t)(it->second * ncomp + c)].insert(cols.begin(), cols.end());
        }
    }
    ia[0] = 0;
    for (int64_t r = 0; r < nWithGhost; ++r) ia[r + 1] = ia[r] + (int64_t)rows[(size_t)r].size();
    int64_t* ja = (int64_t*)std::malloc(sizeof(int64_t) * (size_t)std::max<int64_t>(ia[nWithGhost], 1));
    int64_t k = 0;
    for (int64_t r = 0; r < nWithGhost; ++r)
        for (int64_t c : rows[(size_t)r]) ja[k++] = c;
    *ja_out = ja;
    return ia[nWithGhost];
}

} // extern "C"

// Stencil::computeGraph, Pattern::COUPLED (feeldiscr/stencil.hpp:1560-1634): for every element, every test dof row
// receives the sorted set of all trial dofs of the element (std::set per row, feelalg/graphcsr.hpp:77);
// addMissingZeroEntriesDiagonal (feelalg/graphcsr.cpp:633-659); close() lays rows out with ascending columns
// (graphcsr.cpp:994-1019).
int64_t oracle_graph(int64_t nElt, int nd, const int32_t* l2g, int64_t ndof, int64_t* ia, int64_t** ja_out)
{
    std::vector<std::set<int64_t>> rows(ndof);
    std::vector<int64_t> cols(nd);
    for (int64_t e = 0; e < nElt; ++e) {
        for (int j = 0; j < nd; ++j) cols[j] = l2g[e * nd + j];
        std::sort(cols.begin(), cols.end());
        for (int i = 0; i < nd; ++i) rows[l2g[e * nd + i]].insert(cols.begin(), cols.end());
    }
    for (int64_t r = 0; r < ndof; ++r) rows[r].insert(r);
    ia[0] = 0;
    for (int64_t r = 0; r < ndof; ++r) ia[r + 1] = ia[r] + (int64_t)rows[r].size();
    int64_t* ja = (int64_t*)std::malloc(sizeof(int64_t) * (size_t)std::max<int64_t>(ia[ndof], 1));
    int64_t k = 0;
    for (int64_t r = 0; r < ndof; ++r)
        for (int64_t c : rows[r]) ja[k++] = c;
    *ja_out = ja;
    return ia[ndof];
}

void oracle_free(void* p) { std::free(p); }

int oracle_local_matrix(int kind, int order, int ncomp, int quad_order, const double* G, const oracle_coeff* coeff,
                        double* Ae)
{
    Tables T = make_tables(order, quad_order);
    Tables Tf;
    if (coeff && coeff->kind == ORACLE_COEFF_FIELD) {
        Tf = make_tables(coeff->field_order, quad_order);
    }
    Work W(T.n * ncomp, T.n, T.Q);
    const double* Gp[4] = {G, G + 3, G + 6, G + 9};
    Geo g;
    geo_update(Gp, g);
    local_matrix(kind, ncomp, T, &Tf, coeff, g, 0, Ae, W.grad.data(), W.prod.data(), W.kq.data());
    return 0;
}

// Integrator::assemble element loop (feelvf/integrator.hpp:1093-1207): per element geomap update, form context
// update, integrate, assemble (localToGlobalIndices + addMatrix -> MatSetValues ADD_VALUES).
int oracle_form2(int kind, int order, int ncomp, int quad_order, int64_t nPts, const double* xyz, int64_t nElt,
                 const int32_t* conn, const int32_t* l2g, const int64_t* ia, const int64_t* ja,
                 const oracle_coeff* coeff, int64_t eb, int64_t ee, double* a)
{
    (void)nPts;
    (void)nElt;
    // ORACLE_ABS_FLAG: accumulate |a^e_ij| instead (the cancellation scale used by the parity tolerance)
    const bool absmode = (kind & ORACLE_ABS_FLAG) != 0;
    kind &= ~ORACLE_ABS_FLAG;
    Tables T = make_tables(order, quad_order);
    Tables Tf;
    if (coeff && coeff->kind == ORACLE_COEFF_FIELD) Tf = make_tables(coeff->field_order, quad_order);
    const int n = T.n, nd = n * ncomp;
    Work W(nd, n, T.Q);
    for (int64_t e = eb; e < ee; ++e) {
        const int32_t* v = conn + 4 * e;
        const double* Gp[4] = {xyz + 3 * (int64_t)v[0], xyz + 3 * (int64_t)v[1], xyz + 3 * (int64_t)v[2],
                               xyz + 3 * (int64_t)v[3]};
        Geo g;
        geo_update(Gp, g);
        local_matrix(kind, ncomp, T, &Tf, coeff, g, e, W.Ae.data(), W.grad.data(), W.prod.data(), W.kq.data());
        const int32_t* idx = l2g + e * nd;
        for (int i = 0; i < nd; ++i)
            for (int j = 0; j < nd; ++j) {
                int64_t s = find_slot(ia, ja, idx[i], idx[j]);
                if (s < 0) return 1; // new nonzero outside the preallocated pattern (matrixpetsc.cpp:2688)
                a[s] += absmode ? std::fabs(W.Ae[i * nd + j]) : W.Ae[i * nd + j];
            }
    }
    return 0;
}

// LinearForm::Context::integrate/assemble (feelvf/linearformcontext.hpp:294-309, 391-408): b_i = sum_q prod[q] f(q) phi_i(q)
// scalar: f*id(v); vector: inner(f_vec, id(v)) with a constant vector.  VecSetValues ADD (vectorpetsc.cpp:177-190).
int oracle_form1(int kind, int order, int ncomp, int quad_order, int64_t nPts, const double* xyz, int64_t nElt,
                 const int32_t* conn, const int32_t* l2g, const oracle_coeff* coeff, int64_t eb, int64_t ee, double* b)
{
    (void)nPts;
    (void)nElt;
    (void)kind;
    Tables T = make_tables(order, quad_order);
    Tables Tf;
    if (coeff && coeff->kind == ORACLE_COEFF_FIELD) Tf = make_tables(coeff->field_order, quad_order);
    const int n = T.n, nd = n * ncomp, Q = T.Q;
    std::vector<double> prod(Q), fq(Q), be(nd);
    for (int64_t e = eb; e < ee; ++e) {
        const int32_t* v = conn + 4 * e;
        const double* Gp[4] = {xyz + 3 * (int64_t)v[0], xyz + 3 * (int64_t)v[1], xyz + 3 * (int64_t)v[2],
                               xyz + 3 * (int64_t)v[3]};
        Geo g;
        geo_update(Gp, g);
        for (int q = 0; q < Q; ++q) prod[q] = T.w[q] * g.J;
        for (int c = 0; c < ncomp; ++c) {
            for (int q = 0; q < Q; ++q) {
                if (ncomp == 1) fq[q] = coeff_at(coeff, g, T, &Tf, e, q);
                else fq[q] = coeff ? coeff->value[c] : 1.0;
            }
            for (int i = 0; i < n; ++i) {
                double s = 0;
                for (int q = 0; q < Q; ++q) s += prod[q] * (fq[q] * T.phi[i * Q + q]);
                be[n * c + i] = s;
            }
        }
        for (int i = 0; i < nd; ++i) b[l2g[e * nd + i]] += be[i];
    }
    return 0;
}

} // extern "C"

namespace {

// ---- boundary-face integrals (feelvf/integrator.hpp:3242-3704) -----------------------------------------------------
// Face quadrature: the triangle rule IMTriangle<k> on the reference triangle (-1,-1),(1,-1),(-1,1) is mapped onto face f of
// the reference tet through its boundary map, vertices in f2p order (feelmesh/simplexordering.hpp:341-347), and the weights
// are multiplied by that map's Jacobian sqrt(det(K^T K)) (feelpoly/pointsetquadrature.hpp:696-760).  Identity permutation:
// for a boundary face the permutation only reorders the points of these symmetric rules.  In an element the weights become
//   prod[q] = w_f[q] * J * |B n_f|    (pointsetquadrature.hpp:481-490; n_f the reference normal, feelmesh/refsimplex.hpp:1045-1083,
//   |B n_f| from geomap.hpp:2021-2034)
struct FaceTables {
    int n, Qf;
    std::vector<double> w, xi, phi;   // [4][Qf], [4][Qf][3], [n][4*Qf]
};

void ref_normal(int f, double* nrm)
{
    if (f == 0) nrm[0] = nrm[1] = nrm[2] = std::sqrt(1.0 / 3.0);
    else {
        nrm[0] = nrm[1] = nrm[2] = 0.0;
        nrm[f - 1] = -1.0;
    }
}

FaceTables make_face_tables(int order, int quad_order)
{
    FaceTables T;
    T.n = nloc(order);
    T.Qf = feelgpu_tri_quad_npts[quad_order];
    const double* q = feelgpu_tri_quad_rule[quad_order];
    const int Q = 4 * T.Qf;
    T.w.resize(Q);
    T.xi.resize(3 * Q);
    T.phi.resize(T.n * Q);
    std::vector<double> p(T.n), dp(T.n * 3);
    for (int f = 0; f < 4; ++f) {
        const double* a = REFV[F2P[f][0]];
        const double* b = REFV[F2P[f][1]];
        const double* c = REFV[F2P[f][2]];
        // boundary map of the face: x = a N0 + b N1 + c N2, K = [ (b-a)/2 | (c-a)/2 ], J = sqrt(det(K^T K))
        double k0[3], k1[3];
        for (int d = 0; d < 3; ++d) { k0[d] = 0.5 * (b[d] - a[d]); k1[d] = 0.5 * (c[d] - a[d]); }
        const double g00 = k0[0] * k0[0] + k0[1] * k0[1] + k0[2] * k0[2];
        const double g01 = k0[0] * k1[0] + k0[1] * k1[1] + k0[2] * k1[2];
        const double g11 = k1[0] * k1[0] + k1[1] * k1[1] + k1[2] * k1[2];
        const double Jf = std::sqrt(g00 * g11 - g01 * g01);
        for (int k = 0; k < T.Qf; ++k) {
            const double s = q[3 * k], t = q[3 * k + 1];
            const double N0 = -0.5 * (s + t), N1 = 0.5 * (1.0 + s), N2 = 0.5 * (1.0 + t);
            const int qq = f * T.Qf + k;
            for (int d = 0; d < 3; ++d) T.xi[3 * qq + d] = a[d] * N0 + b[d] * N1 + c[d] * N2;
            T.w[qq] = q[3 * k + 2] * Jf;
            basis(order).eval(&T.xi[3 * qq], p.data(), dp.data());
            for (int i = 0; i < T.n; ++i) T.phi[i * Q + qq] = p[i];
        }
    }
    return T;
}

double face_measure_factor(const Geo& g, int f)
{
    double nrm[3], N[3];
    ref_normal(f, nrm);
    for (int a = 0; a < 3; ++a) N[a] = g.B[a][0] * nrm[0] + g.B[a][1] * nrm[1] + g.B[a][2] * nrm[2];
    return g.J * std::sqrt(N[0] * N[0] + N[1] * N[1] + N[2] * N[2]);
}

inline double face_coeff_at(const oracle_coeff* c, const Geo& g, const FaceTables& T, int q)
{
    if (!c || c->kind == ORACLE_COEFF_CONST) return c ? c->value[0] : 1.0;
    double x[3];
    geo_xreal(g, &T.xi[3 * q], x);
    return eval_expr(c, x[0], x[1], x[2]);
}

} // namespace

extern "C" {

int oracle_face_quadrature(int quad_order, double* xi, double* w)
{
    if (quad_order < 0 || quad_order > FEELGPU_TET_QUAD_MAX_ORDER) return -1;
    FaceTables T = make_face_tables(1, quad_order);
    if (xi) std::memcpy(xi, T.xi.data(), sizeof(double) * T.xi.size());
    if (w) std::memcpy(w, T.w.data(), sizeof(double) * T.w.size());
    return T.Qf;
}

// form1 on boundary faces: b_i += sum_q prod[q] g(x_q) phi_i(x_q); faces = (element, local face) pairs
int oracle_form1_faces(int order, int ncomp, int quad_order, const double* xyz, const int32_t* conn, const int32_t* l2g,
                       int64_t nFaces, const int32_t* felt, const int32_t* flocal, const oracle_coeff* coeff, double* b)
{
    if (coeff && coeff->kind == ORACLE_COEFF_FIELD) return -1;
    FaceTables T = make_face_tables(order, quad_order);
    const int n = T.n, nd = n * ncomp, Qf = T.Qf, Q = 4 * Qf;
    for (int64_t k = 0; k < nFaces; ++k) {
        const int64_t e = felt[k];
        const int f = flocal[k];
        const int32_t* v = conn + 4 * e;
        const double* Gp[4] = {xyz + 3 * (int64_t)v[0], xyz + 3 * (int64_t)v[1], xyz + 3 * (int64_t)v[2],
                               xyz + 3 * (int64_t)v[3]};
        Geo g;
        geo_update(Gp, g);
        const double mf = face_measure_factor(g, f);
        for (int c = 0; c < ncomp; ++c)
            for (int i = 0; i < n; ++i) {
                double s = 0;
                for (int q = f * Qf; q < (f + 1) * Qf; ++q) {
                    const double fq = ncomp == 1 ? face_coeff_at(coeff, g, T, q) : (coeff ? coeff->value[c] : 1.0);
                    s += (T.w[q] * mf) * (fq * T.phi[i * Q + q]);
                }
                b[l2g[e * nd + n * c + i]] += s;
            }
    }
    return 0;
}

// form2 on boundary faces, k idt(u) id(v) (Robin term): a(row_i, col_j) += sum_q prod[q] k(x_q) phi_j phi_i, all i, j
int oracle_form2_faces(int order, int ncomp, int quad_order, const double* xyz, const int32_t* conn, const int32_t* l2g,
                       int64_t nFaces, const int32_t* felt, const int32_t* flocal, const oracle_coeff* coeff,
                       const int64_t* ia, const int64_t* ja, double* a)
{
    if (coeff && coeff->kind == ORACLE_COEFF_FIELD) return -1;
    FaceTables T = make_face_tables(order, quad_order);
    const int n = T.n, nd = n * ncomp, Qf = T.Qf, Q = 4 * Qf;
    for (int64_t k = 0; k < nFaces; ++k) {
        const int64_t e = felt[k];
        const int f = flocal[k];
        const int32_t* v = conn + 4 * e;
        const double* Gp[4] = {xyz + 3 * (int64_t)v[0], xyz + 3 * (int64_t)v[1], xyz + 3 * (int64_t)v[2],
                               xyz + 3 * (int64_t)v[3]};
        Geo g;
        geo_update(Gp, g);
        const double mf = face_measure_factor(g, f);
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) {
                double s = 0;
                for (int q = f * Qf; q < (f + 1) * Qf; ++q)
                    s += (T.w[q] * mf) * (face_coeff_at(coeff, g, T, q) * (T.phi[j * Q + q] * T.phi[i * Q + q]));
                for (int c = 0; c < ncomp; ++c) {
                    const int64_t row = l2g[e * nd + n * c + i], col = l2g[e * nd + n * c + j];
                    const int64_t* lo = std::lower_bound(ja + ia[row], ja + ia[row + 1], col);
                    if (lo == ja + ia[row + 1] || *lo != col) return -2;
                    a[lo - ja] += s;
                }
            }
    }
    return 0;
}

// CPU baseline with P worker threads: contiguous element blocks ("partitions"); a row is owned by the lowest block
// touching it (feeldiscr/doftablempi.hpp:138-200); contributions to rows owned by another block go to a private
// stash and are added by the owner after the loop -- the shape of MatSetValuesLocal + MatAssemblyBegin/End
// (feelalg/matrixpetsc.cpp:2875-2880, 750-752).  Returns seconds spent in element loop + stash merge.
double oracle_form2_threads(int kind, int order, int ncomp, int quad_order, int64_t nPts, const double* xyz,
                            int64_t nElt, const int32_t* conn, const int32_t* l2g, const int64_t* ia,
                            const int64_t* ja, const oracle_coeff* coeff, int P, int64_t eb, int64_t ee, double* a)
{
    (void)nPts;
    (void)nElt;
    const bool absmode = (kind & ORACLE_ABS_FLAG) != 0;   // as in oracle_form2: the cancellation scale sum_e |a^e_ij|
    kind &= ~ORACLE_ABS_FLAG;
    Tables T = make_tables(order, quad_order);
    Tables Tf;
    if (coeff && coeff->kind == ORACLE_COEFF_FIELD) Tf = make_tables(coeff->field_order, quad_order);
    const int n = T.n, nd = n * ncomp;
    if (P < 1) P = 1;
    const int64_t N = ee - eb;
    std::vector<int64_t> cut(P + 1);
    for (int p = 0; p <= P; ++p) cut[p] = eb + N * p / P;
    // setup (untimed, as the DofTable/ghost setup in the reference): owner block of every touched row
    int64_t maxrow = 0;
    for (int64_t e = eb; e < ee; ++e)
        for (int i = 0; i < nd; ++i) maxrow = std::max<int64_t>(maxrow, l2g[e * nd + i]);
    std::vector<int16_t> owner(maxrow + 1, -1);
    for (int p = 0; p < P; ++p)
        for (int64_t e = cut[p]; e < cut[p + 1]; ++e)
            for (int i = 0; i < nd; ++i) {
                int32_t r = l2g[e * nd + i];
                if (owner[r] < 0) owner[r] = (int16_t)p;
            }
    std::vector<std::vector<std::pair<int64_t, double>>> stash(P);
    std::vector<int> status(P, 0);
    auto t0 = std::chrono::steady_clock::now();
    auto worker = [&](int p) {
        Work W(nd, n, T.Q);
        auto& st = stash[p];
        for (int64_t e = cut[p]; e < cut[p + 1]; ++e) {
            const int32_t* v = conn + 4 * e;
            const double* Gp[4] = {xyz + 3 * (int64_t)v[0], xyz + 3 * (int64_t)v[1], xyz + 3 * (int64_t)v[2],
                                   xyz + 3 * (int64_t)v[3]};
            Geo g;
            geo_update(Gp, g);
            local_matrix(kind, ncomp, T, &Tf, coeff, g, e, W.Ae.data(), W.grad.data(), W.prod.data(), W.kq.data());
            if (absmode)
                for (int k = 0; k < nd * nd; ++k) W.Ae[k] = std::fabs(W.Ae[k]);
            const int32_t* idx = l2g + e * nd;
            for (int i = 0; i < nd; ++i) {
                const bool mine = owner[idx[i]] == p;
                for (int j = 0; j < nd; ++j) {
                    int64_t s = find_slot(ia, ja, idx[i], idx[j]);
                    if (s < 0) { status[p] = 1; return; }
                    if (mine) a[s] += W.Ae[i * nd + j];
                    else st.emplace_back(s, W.Ae[i * nd + j]);
                }
            }
        }
    };
    {
        std::vector<std::thread> th;
        for (int p = 1; p < P; ++p) th.emplace_back(worker, p);
        worker(0);
        for (auto& t : th) t.join();
    }
    // MatAssemblyEnd: owners add the stashed triplets (serial per stash here; the stash is O(interface))
    for (int p = 0; p < P; ++p)
        for (auto& sv : stash[p]) a[sv.first] += sv.second;
    auto t1 = std::chrono::steady_clock::now();
    for (int p = 0; p < P; ++p)
        if (status[p]) return -1.0;
    return std::chrono::duration<double>(t1 - t0).count();
}

} // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// on(): Dirichlet elimination.  feelvf/on.hpp:395-449 collects (dof, value) pairs and calls MatrixPetsc::zeroRows
// (feelalg/matrixpetsc.cpp:1762-1860), which is PETSc's MatZeroRows(A, rows, diag, x, b) -- rows zeroed, diag on the
// diagonal, b_r = diag*x_r -- or, with ContextOn::SYMMETRIC, MatZeroRowsColumns -- additionally b_i -= a_ir*x_r and
// a_ir = 0 for the other rows.  ContextOn::KEEP_DIAGONAL restores the assembled diagonal afterwards
// (MatDiagonalSet(diag, INSERT_VALUES)) and sets b_r = x_r * a_rr (matrixpetsc.cpp:1817-1829).
// strategy = ContextOn bits (feelalg/enums.hpp:56-66): 0x1 ELIMINATION (nothing happens without it, matrixpetsc.cpp:1784),
// 0x2 KEEP_DIAGONAL, 0x4 SYMMETRIC.  Sequential restatement, test infrastructure only.
// ---------------------------------------------------------------------------------------------------------------
extern "C" int oracle_zero_rows(int64_t nRows, const int64_t* ia, const int64_t* ja, int64_t n, const int32_t* rows,
                                const double* x, int strategy, double value_on_diagonal, double* a, double* b)
{
    if (!(strategy & 1)) return 0;                       // matrixpetsc.cpp:1784: only ContextOn::ELIMINATION acts
    std::vector<int64_t> mark((size_t)nRows, -1);
    for (int64_t k = 0; k < n; ++k) {
        if (rows[k] < 0 || rows[k] >= nRows) return 1;
        mark[rows[k]] = k;
    }
    std::vector<double> olddiag((size_t)n, 0.0);
    for (int64_t k = 0; k < n; ++k)
        for (int64_t p = ia[rows[k]]; p < ia[rows[k] + 1]; ++p)
            if (ja[p] == rows[k]) olddiag[k] = a[p];
    if (strategy & 4)
        for (int64_t i = 0; i < nRows; ++i) {          // columns first: every other row loses its coupling to the rows
            for (int64_t p = ia[i]; p < ia[i + 1]; ++p) {
                const int64_t k = mark[ja[p]];
                if (k < 0 || ja[p] == i) continue;
                if (mark[i] < 0 && b) b[i] -= a[p] * x[k];
                a[p] = 0.0;
            }
        }
    for (int64_t k = 0; k < n; ++k) {
        const int64_t r = rows[k];
        const double d = (strategy & 2) ? olddiag[k] : value_on_diagonal;
        for (int64_t p = ia[r]; p < ia[r + 1]; ++p) a[p] = (ja[p] == r) ? d : 0.0;
        if (b) b[r] = x[k] * d;
    }
    return 0;
}
