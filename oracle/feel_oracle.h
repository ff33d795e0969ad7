/* feel_oracle.h -- C interface of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a CPU restatement of the reference's (feelpp/feelpp) element-loop
 * assembly for `form2/form1 += integrate(elements(mesh), ...)` over Lagrange Pk tetrahedra.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.  Nothing under
 * feelpp_b200/ includes, links or calls it.
 *
 * Parity status: "parity unpinned" -- the reference cannot be built in this container (no Boost/Eigen/PETSc/MPI,
 * SURVEY.md section 8c) and its test-suite holds no golden matrices for this path; the oracle is pinned by the
 * reference's own analytic invariants (testsuite/feelvf/test_forms.cpp:135-139, test_laplaciant.cpp:108-126,
 * testsuite/feeldiscr/test_wire_basket.cpp:141-142, testsuite/feelpoly/testlagrange.hpp:100-137) and by the
 * quadrature literals extracted from feelpoly/imsimplex.cpp (tests/golden/tet_quadrature.json).
 */
#ifndef FEEL_ORACLE_H
#define FEEL_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* form kinds (same numeric values as include/feelgpu.h, restated here so that the oracle has no product include) */
enum { ORACLE_FORM2_MASS = 0, ORACLE_FORM2_STIFFNESS = 1, ORACLE_FORM2_ELASTICITY = 2 };
enum { ORACLE_FORM1_SOURCE = 0 };
#define ORACLE_ABS_FLAG 0x100  /* oracle_form2: accumulate |a^e_ij| (cancellation scale for tolerances) */
enum { ORACLE_COEFF_CONST = 0, ORACLE_COEFF_FIELD = 1, ORACLE_COEFF_EXPR = 2 };

typedef struct {
    int kind;              /* ORACLE_COEFF_*                                                        */
    double value[3];       /* CONST: value[0] (scalar) or value[0..2] (constant vector, form1)      */
    int field_order;       /* FIELD: Lagrange order of the scalar field                              */
    const int32_t* field_l2g;    /* FIELD: [nElt][nloc(field_order)]                                 */
    const double* field_values;  /* FIELD: [nDofField]                                               */
    int expr_len;          /* EXPR: postfix program length                                           */
    const int32_t* expr_ops;     /* EXPR: opcodes (see feelgpu.h FEELGPU_OP_*)                       */
    const double* expr_consts;   /* EXPR: one constant per op slot (used by PUSH_CONST / POWI)       */
    double lambda, mu;     /* ELASTICITY: Lame coefficients                                          */
} oracle_coeff;

/* quadrature rule IMTetrahedra<order> (imsimplex.cpp); returns number of points, fills q[4*npts] if q != NULL */
int oracle_quadrature(int order, double* q);
/* number of scalar local dofs of Lagrange Pk on a tet */
int oracle_nlocaldof(int order);
/* nodal points of fem::Lagrange<3,k> on the reference tet, Feel++ local order: pts[n][3] */
void oracle_lagrange_points(int order, double* pts);
/* PreCompute: phi[n][Q], dphi[n][Q][3] at reference points xi[Q][3] (polynomialset.hpp:1052-1083) */
void oracle_tabulate(int order, int Q, const double* xi, double* phi, double* dphi);
/* quadrature order the reference's integrate() selects (expr.hpp:1075, operators.hpp:134-153, operations.hpp:437-442) */
int oracle_quad_order_form2(int kind, int order, int coeff_poly_degree /* -1: non polynomial -> 2 */);
int oracle_quad_order_form1(int order, int coeff_poly_degree);

/* DofTable::build: first-touch numbering in element order; returns number of dofs; l2g[nElt][ncomp*n] */
int64_t oracle_doftable(int64_t nPts, int64_t nElt, const int32_t* conn, int order, int ncomp, int32_t* l2g);
/* Stencil::computeGraph (COUPLED) + GraphCSR::close, one process: ia[ndof+1], ja[nnz] ascending. ja malloc'd */
int64_t oracle_graph(int64_t nElt, int nd, const int32_t* l2g, int64_t ndof, int64_t* ia, int64_t** ja);
void oracle_free(void* p);
/* the same in MPI mode for one rank of an element partition (feeldiscr/doftablempi.hpp:138-200,548,593-636,
 * feeldiscr/doftable.hpp:1975-1992, feelalg/graphcsr.cpp:800-951): see feel_oracle.cpp */
int64_t oracle_doftable_mpi(int64_t nElt, const int32_t* conn, const int32_t* part, int nparts, int rank, int order, int ncomp,
                            int32_t* l2g, int64_t* nOwned, int64_t* p2c, int32_t* ghost_owner, int64_t* first_dof);
int64_t oracle_graph_mpi(int64_t nElt, const int32_t* conn, const int32_t* part, int nparts, int rank, int order, int ncomp,
                         int64_t nWithGhost, const int32_t* l2g, int64_t* ia, int64_t** ja);

/* element loop: a[nnz] += local matrices (MatSetValues ADD_VALUES semantics; binary search per row)           */
int oracle_form2(int kind, int order, int ncomp, int quad_order,
                 int64_t nPts, const double* xyz, int64_t nElt, const int32_t* conn,
                 const int32_t* l2g, const int64_t* ia, const int64_t* ja,
                 const oracle_coeff* coeff, int64_t elt_begin, int64_t elt_end, double* a);
int oracle_form1(int kind, int order, int ncomp, int quad_order,
                 int64_t nPts, const double* xyz, int64_t nElt, const int32_t* conn,
                 const int32_t* l2g, const oracle_coeff* coeff, int64_t elt_begin, int64_t elt_end, double* b);
/* local matrix of ONE element (row-major nd x nd, bilinearform.hpp:462-464), for unit tests */
/* on(): MatrixPetsc::zeroRows (feelalg/matrixpetsc.cpp:1762-1860); strategy = ContextOn bits (feelalg/enums.hpp:56-66): 1 elimination (required), 2 keep diagonal, 4 symmetric */
int oracle_zero_rows(int64_t nRows, const int64_t* ia, const int64_t* ja, int64_t n, const int32_t* rows, const double* x,
                     int strategy, double value_on_diagonal, double* a, double* b);
int oracle_local_matrix(int kind, int order, int ncomp, int quad_order, const double* G /*[4][3]*/,
                        const oracle_coeff* coeff, double* Ae);
/* boundary-face integrals (feelvf/integrator.hpp:3242-3704): faces = (element, local face 0..3) pairs; face rule =
 * IMTriangle<quad_order> mapped onto the faces of the reference tet (pointsetquadrature.hpp:696-760).
 * oracle_face_quadrature returns the points per face and fills xi[4][Qf][3], w[4][Qf] (weights include the
 * reference-face Jacobian). */
int oracle_face_quadrature(int quad_order, double* xi, double* w);
int oracle_form1_faces(int order, int ncomp, int quad_order, const double* xyz, const int32_t* conn, const int32_t* l2g,
                       int64_t nFaces, const int32_t* felt, const int32_t* flocal, const oracle_coeff* coeff, double* b);
int oracle_form2_faces(int order, int ncomp, int quad_order, const double* xyz, const int32_t* conn, const int32_t* l2g,
                       int64_t nFaces, const int32_t* felt, const int32_t* flocal, const oracle_coeff* coeff,
                       const int64_t* ia, const int64_t* ja, double* a);
/* multi-threaded CPU baseline: P threads each on a contiguous element block with a private value array
 * merged at the end (shape of the reference's `mpirun -np P` run); returns seconds of the element loop */
double oracle_form2_threads(int kind, int order, int ncomp, int quad_order,
                            int64_t nPts, const double* xyz, int64_t nElt, const int32_t* conn,
                            const int32_t* l2g, const int64_t* ia, const int64_t* ja,
                            const oracle_coeff* coeff, int nthreads, int64_t elt_begin, int64_t elt_end, double* a);

#ifdef __cplusplus
}
#endif
#endif
