// lagrange.hpp -- reference-element tables: quadrature rules and Lagrange Pk tabulation on the reference tet.
//
// Stands where IMGeneral<3,double,Simplex>(order) (feelpoly/im.hpp:143-147, tables feelpoly/imsimplex.cpp) and
// PolynomialSet::PreCompute (feelpoly/polynomialset.hpp:1052-1083) stand: it tabulates, once per form on the host,
// phi_i(xi_q) and d phi_i / d xi_c (xi_q) for fem::Lagrange<3,k> (feelpoly/lagrange.hpp:326) in Feel++'s local dof
// order (vertices, edges e0..e5, faces f0..f3: feelmesh/simplexordering.hpp:283-365).  Nodes: equispaced for k<=2
// (feelpoly/equispaced.hpp:99-140); for k=3 the edge nodes sit at the Gauss-Lobatto points (warp&blend,
// feelpoly/warpblend.hpp:96-121).
#pragma once
#include <vector>

namespace feelpp_b200 {

struct Quadrature {
    int order = 0, nq = 0;
    std::vector<double> xi;  // [nq][3]
    std::vector<double> w;   // [nq]
};
// im(3, order, simplex) (feelpoly/imsimplex.cpp:8375-8396); throws std::out_of_range beyond the tabulated orders
Quadrature im(int order);
// the triangle rule of that order on the 4 faces of the reference tet: 4 blocks of nq/4 points (pointsetquadrature.hpp:696-760)
Quadrature imFaces(int order);

struct BasisTables {
    int n = 0, nq = 0;
    std::vector<double> phi;   // [n][nq]
    std::vector<double> dphi;  // [n][nq][3]
};
// nodal points of the Pk Lagrange element, [n][3]
std::vector<double> lagrangePoints(int order);
// tabulate the Pk basis at reference points xi[nq][3]
BasisTables tabulate(int order, int nq, const double* xi);

} // namespace feelpp_b200
