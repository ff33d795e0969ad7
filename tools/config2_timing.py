#!/usr/bin/env python3
"""kernel times (CUDA events recorded by the library) of config 2's whole sequence on the bench mesh:
a = mass ; a += stiffness ; l = 1*id(v) ; l = f*id(v)"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import feelpp_b200 as fpp
from feelpp_b200 import cst, elements, expr, grad, gradt, id, idt, integrate, trans
from feelpp_b200._capi import check
from feelpp_b200 import engine as E

m = int(sys.argv[1]) if len(sys.argv) > 1 else 110
eng = fpp.Engine(0)
mesh = fpp.unitCube(m, 0.15, False)
Xh = fpp.Pch(mesh, 2)
a = fpp.form2(_test=Xh, _trial=Xh, engine=eng)
l = fpp.form1(_test=Xh, engine=eng)
u = v = fpp.Element(Xh)
lib = eng.lib
keep = E._Keep()
dm = a._desc(integrate(_range=elements(mesh), _expr=idt(u) * id(v)), keep)
dk = a._desc(integrate(_range=elements(mesh), _expr=gradt(u) * trans(grad(v))), keep)
check(lib.feelgpu_set_timing(eng.ctx, 1), eng.ctx)
tm = (C.c_double * 3)()


def form2_ms(desc, acc, n=10):
    out = []
    for _ in range(n):
        check(lib.feelgpu_form2_assemble(a.pat, C.byref(desc), acc, a.d_values, None), eng.ctx)
        check(lib.feelgpu_last_timing(eng.ctx, tm), eng.ctx)
        out.append(tm[1])
    return sorted(out)[len(out) // 2]


print("a  = mass (Q=11)       kernel ms", round(form2_ms(dm, 0), 4))
print("a += stiffness (Q=4)   kernel ms", round(form2_ms(dk, 1), 4))
print("a  = stiffness         kernel ms", round(form2_ms(dk, 0), 4))
print("a += mass              kernel ms", round(form2_ms(dm, 1), 4))
check(lib.feelgpu_set_timing(eng.ctx, 0), eng.ctx)
for name, ex in (("l = 1*id(v) (Q=4)", cst(1.0) * id(v)), ("l = sin(pi x)cos(pi y)z*id(v) (Q=11)", expr("sin(pi*x)*cos(pi*y)*z") * id(v))):
    integ = integrate(_range=elements(mesh), _expr=ex)
    l.assign(integ)
    eng.sync()
    t0 = time.perf_counter()
    for _ in range(10):
        l.assign(integ)
    eng.sync()
    print(name, "ms per assembly (host clock, 10 back to back)", round(1e2 * (time.perf_counter() - t0), 4))

# y = A x on the assembled matrix (device hand-off)
import numpy as np
n = Xh.nDof()
dx = eng.to_device(np.ones(n))
dy = eng.malloc(8 * n)
check(lib.feelgpu_matvec(a.pat, a.d_values, dx, dy, None), eng.ctx)
eng.sync()
t0 = time.perf_counter()
for _ in range(20):
    check(lib.feelgpu_matvec(a.pat, a.d_values, dx, dy, None), eng.ctx)
eng.sync()
ms = 1e3 * (time.perf_counter() - t0) / 20
nnz = int(a.nnz)
print("y = A x  ms", round(ms, 4), " GB/s (12 B per nonzero + 24 B per row)", round((12.0 * nnz + 24.0 * n) / ms / 1e6, 1))
