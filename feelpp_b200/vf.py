"""Variational-form language of the harness: the subset of Feel++'s feelvf DSEL that the engine accelerates.

    a = form2(_test=Xh, _trial=Xh, engine=eng)
    a.assign(integrate(_range=elements(mesh), _expr=gradt(u) * trans(grad(v))))      # a  = integrate(...)
    a += integrate(_range=elements(mesh), _expr=idt(u) * id(v))                       # a += integrate(...)

Mirrors feelvf/form.hpp:92-147 (form1/form2), integrate.hpp:64-98 (integrate, quadrature orders from the
expression), operators.hpp:134-153 (id/idt/grad/gradt and their polynomial orders), operations.hpp:437-442
(product order = sum), inner.hpp:154, expr.hpp:1046-1079 (ExpressionOrder).  An expression outside the supported set
raises NotImplementedError -- on the Feel++ side the adapter would fall through to the reference loop; the engine
itself has no CPU path.
"""
import numpy as np

from . import hostmirror as hm

FORM2_MASS, FORM2_STIFFNESS, FORM2_ELASTICITY = 0, 1, 2
FORM1_SOURCE = 0
COEFF_CONST, COEFF_FIELD, COEFF_EXPR = 0, 1, 2


class Element:
    """an element u of a function space (only its space and, for idv, its values matter here)."""

    def __init__(self, space, values=None, name="u"):
        self.space, self.values, self.name = space, values, name


class Expr:
    def __mul__(self, o):
        return Mul(self, _lift(o))

    def __rmul__(self, o):
        return Mul(_lift(o), self)

    def __add__(self, o):
        return Add(self, _lift(o))


def _lift(x):
    return x if isinstance(x, Expr) else Cst(float(x))


class Cst(Expr):
    def __init__(self, v):
        self.v = v


class CstVec(Expr):
    def __init__(self, v):
        self.v = [float(t) for t in v]


class Idv(Expr):
    def __init__(self, elt):
        self.elt = elt


class Ginac(Expr):
    def __init__(self, src):
        self.prog = hm.GinacProgram(src)


class Op(Expr):
    def __init__(self, kind, elt):
        self.kind, self.elt = kind, elt   # kind in id, idt, grad, gradt


class Trans(Expr):
    def __init__(self, e):
        self.e = e


class Sym(Expr):
    def __init__(self, e):
        self.e = e


class Trace(Expr):
    def __init__(self, e):
        self.e = e


class IdMat(Expr):
    pass


class Mul(Expr):
    def __init__(self, a, b):
        self.a, self.b = a, b


class Add(Expr):
    def __init__(self, a, b):
        self.a, self.b = a, b


class Inner(Expr):
    def __init__(self, a, b):
        self.a, self.b = a, b


def cst(v): return Cst(float(v))
def vec(*v): return CstVec(v)
def idv(k): return Idv(k)
def expr(s): return Ginac(s)
def id(v): return Op("id", v)          # noqa: A001  (Feel++ name)
def idt(u): return Op("idt", u)
def grad(v): return Op("grad", v)
def gradt(u): return Op("gradt", u)
def trans(e): return Trans(e)
def sym(e): return Sym(e)
def trace(e): return Trace(e)
def Id(): return IdMat()
def inner(a, b): return Inner(a, b)


class Coefficient:
    """product of scalar factors cst * idv(k) * expr(...); at most one non-constant factor is supported."""

    def __init__(self):
        self.scale, self.field, self.prog, self.vector = 1.0, None, None, None

    def mul(self, e):
        if isinstance(e, Cst):
            self.scale *= e.v
        elif isinstance(e, CstVec):
            if self.vector is not None:
                raise NotImplementedError("product of two vector constants")
            self.vector = list(e.v)
        elif isinstance(e, Idv):
            if self.field is not None or self.prog is not None:
                raise NotImplementedError("more than one non-constant coefficient factor")
            self.field = e.elt
        elif isinstance(e, Ginac):
            if self.field is not None or self.prog is not None:
                raise NotImplementedError("more than one non-constant coefficient factor")
            self.prog = e.prog
        else:
            raise NotImplementedError("unsupported coefficient factor %r" % type(e).__name__)

    def kind(self):
        return COEFF_FIELD if self.field is not None else COEFF_EXPR if self.prog is not None else COEFF_CONST

    def polyOrder(self):
        # cst: 0; idv(P_m): m; expr: total degree if polynomial else 2 (feelvf/detail/ginacexvf.hpp:836-845)
        if self.field is not None:
            return self.field.space.order
        if self.prog is not None:
            return self.prog.polyDegree if self.prog.polyDegree >= 0 else 2
        return 0


def _factors(e, out):
    """flatten a product tree into its factors"""
    if isinstance(e, Mul):
        _factors(e.a, out)
        _factors(e.b, out)
    else:
        out.append(e)
    return out


def _is(e, kind):
    return isinstance(e, Op) and e.kind == kind


class Classified:
    def __init__(self, kind, coeff, order, lam=0.0, mu=0.0):
        self.kind, self.coeff, self.quad_order, self.lam, self.mu = kind, coeff, order, lam, mu


def _quad(poly):
    # ExpressionOrder: order = polyOrder * nOrderGeo, or nOrderGeo when 0 (feelvf/expr.hpp:1075), nOrderGeo = 1
    return poly if poly > 0 else 1


def classify_form2(e, k):
    """map an expression onto (kind, coefficient, quadrature order) -- the adapter's pattern match."""
    fs = _factors(e, [])
    coeff = Coefficient()
    core = []
    for f in fs:
        if isinstance(f, (Cst, Idv, Ginac)):
            coeff.mul(f)
        else:
            core.append(f)
    if len(core) == 2:
        a, b = core
        # idt(u)*id(v) or id(v)*idt(u)
        if (_is(a, "idt") and _is(b, "id")) or (_is(a, "id") and _is(b, "idt")):
            return Classified(FORM2_MASS, coeff, _quad(2 * k + coeff.polyOrder()))
        # gradt(u)*trans(grad(v))
        if _is(a, "gradt") and isinstance(b, Trans) and _is(b.e, "grad"):
            return Classified(FORM2_STIFFNESS, coeff, _quad(2 * (k - 1) + coeff.polyOrder()))
    if len(core) == 1 and isinstance(core[0], Inner):
        a, b = core[0].a, core[0].b
        if (_is(a, "gradt") and _is(b, "grad")) or (_is(a, "grad") and _is(b, "gradt")):
            return Classified(FORM2_STIFFNESS, coeff, _quad(2 * (k - 1) + coeff.polyOrder()))
        if (_is(a, "idt") and _is(b, "id")) or (_is(a, "id") and _is(b, "idt")):
            return Classified(FORM2_MASS, coeff, _quad(2 * k + coeff.polyOrder()))
        # inner( lambda*trace(sym(gradt(u)))*Id() + 2*mu*sym(gradt(u)), grad(v) )
        if isinstance(a, Add) and _is(b, "grad") and coeff.kind() == COEFF_CONST:
            lam = mu = None
            for term in (a.a, a.b):
                tf = _factors(term, [])
                sc = 1.0
                rest = []
                for f in tf:
                    if isinstance(f, Cst):
                        sc *= f.v
                    else:
                        rest.append(f)
                if (len(rest) == 2 and isinstance(rest[0], Trace) and isinstance(rest[0].e, Sym)
                        and _is(rest[0].e.e, "gradt") and isinstance(rest[1], IdMat)):
                    lam = sc
                elif len(rest) == 1 and isinstance(rest[0], Sym) and _is(rest[0].e, "gradt"):
                    mu = sc / 2.0
            if lam is not None and mu is not None:
                return Classified(FORM2_ELASTICITY, coeff, _quad(2 * (k - 1)), lam * coeff.scale, mu * coeff.scale)
    raise NotImplementedError("bilinear expression outside the accelerated set")


def classify_form1(e, k):
    fs = _factors(e, [])
    coeff = Coefficient()
    core = []
    for f in fs:
        if isinstance(f, (Cst, Idv, Ginac)):
            coeff.mul(f)
        else:
            core.append(f)
    if len(core) == 1 and _is(core[0], "id"):
        return Classified(FORM1_SOURCE, coeff, _quad(k + coeff.polyOrder()))
    if len(core) == 1 and isinstance(core[0], Inner):
        a, b = core[0].a, core[0].b
        if isinstance(b, CstVec):
            a, b = b, a
        if isinstance(a, CstVec) and _is(b, "id"):
            coeff.mul(a)
            return Classified(FORM1_SOURCE, coeff, _quad(k))
    raise NotImplementedError("linear expression outside the accelerated set")


class Integrator:
    def __init__(self, rng, e, quad=None):
        self.range, self.expr, self.quad = rng, e, quad


def integrate(_range, _expr, _quad=None):
    """integrate(_range=elements(mesh) | boundaryfaces(mesh) | markedfaces(mesh, m), _expr=..., [_quad=order]) (feelvf/integrate.hpp:64-98)."""
    if not isinstance(_range, (hm.ElementRange, hm.BoundaryFaceRange)):
        raise NotImplementedError("only elements(mesh) and boundaryfaces(mesh) ranges are accelerated")
    return Integrator(_range, _expr, _quad)


# ContextOn (feelalg/enums.hpp:56-66) and ContextOnMap (feelalg/enums.cpp:28-35), numerically as in the reference
ON_PENALISATION, ON_ELIMINATION, ON_KEEP_DIAGONAL, ON_SYMMETRIC = 0, 1, 2, 4
_ON_TYPES = {"elimination": ON_ELIMINATION, "elimination_keep_diagonal": ON_ELIMINATION | ON_KEEP_DIAGONAL,
             "elimination_symmetric": ON_ELIMINATION | ON_SYMMETRIC,
             "elimination_symmetric_keep_diagonal": ON_ELIMINATION | ON_SYMMETRIC | ON_KEEP_DIAGONAL}


class On:
    """on(_range=boundaryfaces(mesh) | markedfaces(mesh, m), _rhs=l, _element=u, _expr=g, _type="elimination", _value_on_diagonal=1.0)
    (feelvf/on.hpp:1126-1190): the Dirichlet dofs and their values are collected on the host by nodal interpolation of
    g (on.hpp:395-449: fe->interpolate(expr)); the elimination itself runs on the device CSR (feelgpu_form2_zero_rows)."""

    def __init__(self, _range, _rhs, _element, _expr, _type="elimination", _value_on_diagonal=1.0):
        if not isinstance(_range, hm.BoundaryFaceRange):
            raise NotImplementedError("on(): only boundaryfaces(mesh) ranges")
        if _type not in _ON_TYPES:
            raise NotImplementedError("on(): type %r stays on the host path (penalisation)" % _type)
        self.range, self.rhs, self.element, self.expr = _range, _rhs, _element, _expr
        self.strategy, self.value_on_diagonal = _ON_TYPES[_type], float(_value_on_diagonal)

    def dofs_and_values(self):
        Xh = self.element.space
        nodes = Xh.boundaryDofs(self.range.faces() if isinstance(self.range, hm.MarkedFaceRange) else None)
        p = Xh.dofPoints()[nodes]
        g = self.expr
        if callable(g):
            val = np.asarray(g(p[:, 0], p[:, 1], p[:, 2]), dtype=np.float64)
        elif isinstance(g, Cst):
            val = np.full(len(nodes), float(g.v))
        elif isinstance(g, Ginac):
            val = g.prog.evaluate(p)
        else:
            val = np.full(len(nodes), float(g))
        if Xh.ncomp == 1:
            return nodes.astype(np.int32), np.broadcast_to(val, (len(nodes),)).astype(np.float64).copy()
        val = np.asarray(val, dtype=np.float64)
        if val.ndim == 1:
            val = np.repeat(val[:, None], Xh.ncomp, axis=1)
        dofs = (nodes[:, None] * Xh.ncomp + np.arange(Xh.ncomp)[None, :]).ravel()
        return dofs.astype(np.int32), val.ravel().copy()


def on(_range, _rhs, _element, _expr, _type="elimination", _value_on_diagonal=1.0):
    return On(_range, _rhs, _element, _expr, _type, _value_on_diagonal)
