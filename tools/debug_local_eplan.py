"""One GPU: the local problem of ONE rank of a partitioned mesh assembled through the entity plan and through the scatter
kernel (ghost rows stay in the local tail, no exchange); rows that differ are listed.  Debug aid for the multi-GPU path."""
import sys, os, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import feelpp_b200 as fpp
from feelpp_b200 import parallel as par, distributed as dd
from feelpp_b200 import elements, grad, gradt, integrate, trans
from test_partitioned_numbering import ThreadComm

m, world, rank = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
part = np.load(sys.argv[4])["part"].astype(np.int32)
mesh = fpp.unitCube(m, 0.15, False)
comm = ThreadComm(world)
out = [None] * world
def run(r):
    out[r] = par.LocalProblem(mesh, part, 2, 1, comm.view(r))
th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
[x.start() for x in th]; [x.join() for x in th]
lp = out[rank]
eng = fpp.Engine(0)
Xh = dd.LocalSpace(lp)
u = v = fpp.Element(Xh)
form = integrate(_range=elements(lp.mesh), _expr=gradt(u) * trans(grad(v)))
res = {}
for name, algo, flags in (("eplan", fpp.ALGO_EPLAN, 5), ("scatter", fpp.ALGO_SCATTER, 1)):
    a = fpp.form2(_test=Xh, _trial=Xh, engine=eng, algorithm=algo, pattern_flags=flags)
    a.assign(form)
    res[name] = a.values()
    a.close()
ia = lp.ia
d = np.abs(res["eplan"] - res["scatter"])
sc = np.abs(res["scatter"]).max()
bad = d > 1e-9 * sc
rows = np.repeat(np.arange(len(ia) - 1), np.diff(ia))
brow = np.unique(rows[bad])
print("rows", len(ia) - 1, "owned", lp.nDofOwned, "values differing", int(bad.sum()), "rows differing", len(brow))
print("first differing rows", brow[:40])
zero = np.add.reduceat(np.abs(res["eplan"]), ia[:-1]) == 0
print("all-zero rows in the entity-plan result:", int(zero.sum()), "in scatter:", int((np.add.reduceat(np.abs(res["scatter"]), ia[:-1]) == 0).sum()))
for r in brow[:6]:
    lo, hi = ia[r], ia[r + 1]
    print("row", r, "owned" if r < lp.nDofOwned else "ghost", "len", hi - lo, "eplan nz", int((res["eplan"][lo:hi] != 0).sum()), "scatter nz", int((res["scatter"][lo:hi] != 0).sum()))
