#!/usr/bin/env python3
"""host plan builder on the CPU: wall time and a digest of the plan (a builder change that keeps the digest is a pure
speed change: the kernel interprets the same bytes)"""
import hashlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import feelpp_b200 as fpp
import test_eplan_host as T

m = int(sys.argv[1]) if len(sys.argv) > 1 else 32
order = int(sys.argv[2]) if len(sys.argv) > 2 else 2
mesh = fpp.unitCube(m, 0.15, False)
Xh = fpp.Pch(mesh, order)
ia, ja = Xh.graph()
t0 = time.time()
pos = T.position_map(Xh.l2g, ia, ja, 1)
t1 = time.time()
T.position_map = lambda *a: pos
t2 = time.time()
P = T.build_plan(mesh, Xh, ia, ja)
t3 = time.time()
h = hashlib.sha1()
for k in ("ctas", "verts", "visits", "chunks", "blob", "rows"):
    h.update(P[k].tobytes())
print("elements", mesh.numElements(), "build_s", round(t3 - t2, 3), "(position map in numpy %.2f s)" % (t1 - t0),
      "ctas", len(P["ctas"]), "digest", h.hexdigest()[:16])
st = P["stats"]
print("iterations padded/real %.4f  modelled S-load wavefronts / ideal %.4f  (%.0f)" % (st["iters_padded"] / st["iters"], st["wavefronts"] / st["wavefronts_ideal"], st["wavefronts"]))
