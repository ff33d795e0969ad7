// ginac.hpp -- closed-form coefficient expressions `expr("1+x*y+z^2")`.
//
// Feel++ parses such strings with GiNaC and compiles them to native code (feelvf/detail/ginacexvf.hpp:750-848);
// the polynomial degree of the result drives the quadrature order (total degree if polynomial in x,y,z, else 2:
// ginacexvf.hpp:836-845).  Here the string is parsed by a small recursive-descent parser into a postfix program
// (FEELGPU_OP_*) which the device evaluates at xReal, together with its polynomial degree (-1: not a polynomial).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace feelpp_b200 {

struct GinacProgram {
    std::string source;
    std::vector<int32_t> ops;
    std::vector<double> consts;   // one slot per op (read for FEELGPU_OP_CONST)
    int polyDegree = 0;           // total degree in (x,y,z); -1 when not polynomial
    double eval(double x, double y, double z) const;   // host evaluation (used by tests / interpolation)
};

// grammar: sum := term (('+'|'-') term)* ; term := unary (('*'|'/') unary)* ; unary := '-' unary | power ;
// power := atom ('^' unary)? ; atom := number | x | y | z | pi | f '(' sum [',' sum] ')' | '(' sum ')'
// functions: sin cos exp log sqrt pow.  Throws std::invalid_argument on a syntax error.
GinacProgram parseExpr(std::string const& s);

} // namespace feelpp_b200
