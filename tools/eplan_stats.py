#!/usr/bin/env python3
"""Statistics of the entity plan built on the host (no GPU): tasks, iterations, padding, visits, modelled bank conflicts."""
import sys, time
import numpy as np
sys.path.insert(0, __file__.rsplit("/tools", 1)[0])
sys.path.insert(0, __file__.rsplit("/tools", 1)[0] + "/tests")
import feelpp_b200 as fpp
from test_eplan_host import build_plan

m = int(sys.argv[1]) if len(sys.argv) > 1 else 30
order = int(sys.argv[2]) if len(sys.argv) > 2 else 2
shuffle = len(sys.argv) > 3 and sys.argv[3] == "shuffle"
mesh = fpp.unitCube(m, 0.15, shuffle)
Xh = fpp.Pch(mesh, order)
ia, ja = Xh.graph()
t0 = time.time()
P = build_plan(mesh, Xh, ia, ja)
t1 = time.time()
ne = mesh.numElements()
st = P["stats"]
nbytes = sum(P[k].nbytes for k in ("ctas", "verts", "visits", "chunks", "blob", "rows"))
print("elements %d  nnz %d  CTAs %d  build %.1fs" % (ne, ia[-1], len(P["ctas"]), t1 - t0))
print("per element: tasks %.2f  iters %.2f  padded %.2f  visits %.2f  plan bytes %.1f (blob %.1f)" % (
    st["tasks"] / ne, st["iters"] / ne, st["iters_padded"] / ne, st["visits"] / ne, nbytes / ne, P["blob"].nbytes / ne))
print("S-load wavefronts per element: %.2f  ideal %.2f  conflict factor %.2f" % (
    st["wavefronts"] / ne, st["wavefronts_ideal"] / ne, st["wavefronts"] / st["wavefronts_ideal"]))
c = P["ctas"]
print("per CTA: rows %.0f  visits %.0f  verts %.0f  out %.0f  chunks %.0f" % (
    c["nrows"].mean(), c["nvisit"].mean(), c["nverts"].mean(), (c["nout"] & 0x7FFF).mean(), c["nchunk"].mean()))
ty = P["chunks"]["type"]
for t in range(6):
    sel = ty == t
    if sel.any():
        print("  type %d: chunks %d  mean niter %.1f" % (t, sel.sum(), P["chunks"]["niter"][sel].mean()))
print("phase A1 coordinate reads: modelled passes after / before levelling the vertex slots: %.3f" % st.get("a1_ratio", 0))
r = P["rows"]
mis = ((r["base"] ^ r["off"].astype(np.int64)) & 1) == 1
print("row segments: %d (%.1f per CTA, mean length %.1f); parity mismatch (copied by warps, not by TMA): %d = %.2f %% of the values" % (
    len(r), len(r) / len(P["ctas"]), r["len"].mean(), mis.sum(), 100.0 * r["len"][mis].sum() / r["len"].sum()))
