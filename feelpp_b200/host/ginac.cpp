#include "ginac.hpp"

#include <cctype>
#include <cmath>
#include <cstdlib>
#include <stdexcept>

#include "../../include/feelgpu.h"

namespace feelpp_b200 {

namespace {

struct Parser {
    const std::string& s;
    size_t p = 0;
    GinacProgram& out;
    Parser(const std::string& str, GinacProgram& o) : s(str), out(o) {}
    void ws() { while (p < s.size() && std::isspace((unsigned char)s[p])) ++p; }
    bool eat(char c)
    {
        ws();
        if (p < s.size() && s[p] == c) { ++p; return true; }
        return false;
    }
    [[noreturn]] void fail(const std::string& m) { throw std::invalid_argument("expr(\"" + s + "\"): " + m + " at position " + std::to_string(p)); }
    void emit(int op, double c = 0.0) { out.ops.push_back(op); out.consts.push_back(c); }
    // every rule returns the polynomial degree of what it parsed (-1: not polynomial); constVal set when constant
    struct Info { int deg; bool isConst; double val; };
    static int addDeg(int a, int b) { return (a < 0 || b < 0) ? -1 : (a > b ? a : b); }
    Info sum()
    {
        Info a = term();
        for (;;) {
            if (eat('+')) { Info b = term(); emit(FEELGPU_OP_ADD); a = {addDeg(a.deg, b.deg), a.isConst && b.isConst, a.val + b.val}; }
            else if (eat('-')) { Info b = term(); emit(FEELGPU_OP_SUB); a = {addDeg(a.deg, b.deg), a.isConst && b.isConst, a.val - b.val}; }
            else return a;
        }
    }
    Info term()
    {
        Info a = unary();
        for (;;) {
            if (eat('*')) {
                Info b = unary();
                emit(FEELGPU_OP_MUL);
                a = {(a.deg < 0 || b.deg < 0) ? -1 : a.deg + b.deg, a.isConst && b.isConst, a.val * b.val};
            } else if (eat('/')) {
                Info b = unary();
                emit(FEELGPU_OP_DIV);
                a = {(a.deg < 0 || b.deg != 0) ? -1 : a.deg, a.isConst && b.isConst, a.val / b.val};
            } else return a;
        }
    }
    Info unary()
    {
        if (eat('-')) { Info a = unary(); emit(FEELGPU_OP_NEG); a.val = -a.val; return a; }
        if (eat('+')) return unary();
        return power();
    }
    Info power()
    {
        Info a = atom();
        if (eat('^')) {
            Info b = unary();
            emit(FEELGPU_OP_POW);
            int deg = -1;
            if (a.deg == 0 && b.deg == 0) deg = 0;
            else if (a.deg >= 0 && b.isConst && b.val >= 0 && b.val == std::floor(b.val) && b.val <= 64) deg = a.deg * (int)b.val;
            return {deg, a.isConst && b.isConst, std::pow(a.val, b.val)};
        }
        return a;
    }
    Info atom()
    {
        ws();
        if (p >= s.size()) fail("unexpected end");
        if (eat('(')) { Info a = sum(); if (!eat(')')) fail("expected ')'"); return a; }
        if (std::isdigit((unsigned char)s[p]) || s[p] == '.') {
            char* end = nullptr;
            double v = std::strtod(s.c_str() + p, &end);
            if (end == s.c_str() + p) fail("bad number");
            p = end - s.c_str();
            emit(FEELGPU_OP_CONST, v);
            return {0, true, v};
        }
        if (std::isalpha((unsigned char)s[p])) {
            size_t b = p;
            while (p < s.size() && (std::isalnum((unsigned char)s[p]) || s[p] == '_')) ++p;
            std::string id = s.substr(b, p - b);
            if (id == "x") { emit(FEELGPU_OP_X); return {1, false, 0}; }
            if (id == "y") { emit(FEELGPU_OP_Y); return {1, false, 0}; }
            if (id == "z") { emit(FEELGPU_OP_Z); return {1, false, 0}; }
            if (id == "pi" || id == "Pi") { emit(FEELGPU_OP_CONST, M_PI); return {0, true, M_PI}; }
            int op = -1;
            if (id == "sin") op = FEELGPU_OP_SIN;
            else if (id == "cos") op = FEELGPU_OP_COS;
            else if (id == "exp") op = FEELGPU_OP_EXP;
            else if (id == "log") op = FEELGPU_OP_LOG;
            else if (id == "sqrt") op = FEELGPU_OP_SQRT;
            else if (id == "pow") op = FEELGPU_OP_POW;
            else fail("unknown symbol '" + id + "'");
            if (!eat('(')) fail("expected '(' after " + id);
            Info a = sum();
            if (op == FEELGPU_OP_POW) {
                if (!eat(',')) fail("pow needs two arguments");
                Info e = sum();
                if (!eat(')')) fail("expected ')'");
                emit(op);
                int deg = -1;
                if (a.deg == 0 && e.deg == 0) deg = 0;
                else if (a.deg >= 0 && e.isConst && e.val >= 0 && e.val == std::floor(e.val) && e.val <= 64) deg = a.deg * (int)e.val;
                return {deg, a.isConst && e.isConst, std::pow(a.val, e.val)};
            }
            if (!eat(')')) fail("expected ')'");
            emit(op);
            double v = a.val;
            switch (op) {
            case FEELGPU_OP_SIN: v = std::sin(v); break;
            case FEELGPU_OP_COS: v = std::cos(v); break;
            case FEELGPU_OP_EXP: v = std::exp(v); break;
            case FEELGPU_OP_LOG: v = std::log(v); break;
            default: v = std::sqrt(v); break;
            }
            return {a.deg == 0 ? 0 : -1, a.isConst, v};
        }
        fail(std::string("unexpected character '") + s[p] + "'");
    }
};

} // namespace

GinacProgram parseExpr(std::string const& s)
{
    GinacProgram g;
    g.source = s;
    Parser P(s, g);
    auto info = P.sum();
    P.ws();
    if (P.p != s.size()) P.fail("trailing input");
    if ((int)g.ops.size() > FEELGPU_EXPR_MAX_LEN) throw std::invalid_argument("expr: program longer than FEELGPU_EXPR_MAX_LEN");
    g.polyDegree = info.deg;
    return g;
}

double GinacProgram::eval(double x, double y, double z) const
{
    double st[64];
    int sp = 0;
    for (size_t k = 0; k < ops.size(); ++k) {
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
        case FEELGPU_OP_SIN: st[sp - 1] = std::sin(st[sp - 1]); break;
        case FEELGPU_OP_COS: st[sp - 1] = std::cos(st[sp - 1]); break;
        case FEELGPU_OP_EXP: st[sp - 1] = std::exp(st[sp - 1]); break;
        case FEELGPU_OP_LOG: st[sp - 1] = std::log(st[sp - 1]); break;
        case FEELGPU_OP_SQRT: st[sp - 1] = std::sqrt(st[sp - 1]); break;
        case FEELGPU_OP_POW: --sp; st[sp - 1] = std::pow(st[sp - 1], st[sp]); break;
        }
    }
    return st[0];
}

} // namespace feelpp_b200
