#!/usr/bin/env python3
"""where the end-to-end (host buffers) step of bench.py spends its time: coordinate upload, assembly + D2H, bare D2H"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import feelpp_b200 as fpp
from feelpp_b200 import elements, grad, gradt, integrate, trans
from feelpp_b200._capi import check, c_dblp, ptr
from feelpp_b200 import engine as E

m = int(sys.argv[1]) if len(sys.argv) > 1 else 110
eng = fpp.Engine(0)
mesh = fpp.unitCube(m, 0.15, False)
Xh = fpp.Pch(mesh, 2)
a = fpp.form2(_test=Xh, _trial=Xh, engine=eng)
u = v = fpp.Element(Xh)
keep = E._Keep()
desc = a._desc(integrate(_range=elements(mesh), _expr=gradt(u) * trans(grad(v))), keep)
lib = eng.lib
nnz = int(a.nnz) if hasattr(a, "nnz") else int(Xh.graph()[0][-1])
hp = C.c_void_p()
check(lib.feelgpu_host_alloc_pinned(eng.ctx, 8 * nnz, C.byref(hp)), eng.ctx)
hv = C.cast(hp, c_dblp)
mh = eng.mesh(mesh)


def timed(f, n=5):
    f()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        f()
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    return 1e3 * min(ts), 1e3 * sorted(ts)[len(ts) // 2]


print("nnz", nnz, "bytes", 8 * nnz)
hx = C.c_void_p()
check(lib.feelgpu_host_alloc_pinned(eng.ctx, mesh.xyz.nbytes, C.byref(hx)), eng.ctx)
C.memmove(hx, mesh.xyz.ctypes.data, mesh.xyz.nbytes)
px = C.cast(hx, c_dblp)
print("update_points (pageable) min/median ms", timed(lambda: check(lib.feelgpu_mesh_update_points(mh, ptr(mesh.xyz, c_dblp), None), eng.ctx)))
print("update_points (pinned)   min/median ms", timed(lambda: check(lib.feelgpu_mesh_update_points(mh, px, None), eng.ctx)))


def step():
    check(lib.feelgpu_mesh_update_points(mh, px, None), eng.ctx)
    check(lib.feelgpu_form2_assemble_host(a.pat, C.byref(desc), 0, hv), eng.ctx)


print("whole step     min/median ms", timed(step))
print("assemble_host  min/median ms", timed(lambda: check(lib.feelgpu_form2_assemble_host(a.pat, C.byref(desc), 0, hv), eng.ctx)))
print("bare d2h       min/median ms", timed(lambda: check(lib.feelgpu_memcpy_d2h(eng.ctx, hp, a.d_values, 8 * nnz), eng.ctx)))
print("assemble (dev) min/median ms", timed(lambda: check(lib.feelgpu_form2_assemble(a.pat, C.byref(desc), 0, a.d_values, None), eng.ctx)))
