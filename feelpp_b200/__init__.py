"""feelpp_b200 -- B200-native assembly engine for Feel++'s form2/form1 += integrate(elements(mesh), ...) path.

The product is the CUDA library behind include/feelgpu.h (feelpp_b200/csrc -> feelpp_b200/lib/libfeelgpu.so) and the
C++ host mirror (feelpp_b200/host).  This Python package is the test/bench harness around them (ctypes only).
"""
from .hostmirror import (Mesh, unitCube, loadMesh, elements, boundaryfaces, markedfaces, FunctionSpace, Pch, Pchv, im, tabulate, lagrangePoints,
                         nLocalDof, GinacProgram)
from .vf import (Element, cst, vec, idv, expr, id, idt, grad, gradt, trans, sym, trace, Id, inner, integrate, on,
                 classify_form1, classify_form2)
from .engine import (Engine, form1, form2, ALGO_AUTO, ALGO_SCATTER, ALGO_ROWPLAN, ALGO_EPLAN, ALGO_RPLAN,
                     PATTERN_SLOTMAP, PATTERN_ROWPLAN, PATTERN_EPLAN, PATTERN_RPLAN, PATTERN_ASYNC)
from ._capi import FeelGpuError
