"""The row-thread plan (feelpp_b200/csrc/rplan.h) interpreted on the CPU: every (row, element) unit of the mesh occurs
exactly once, in the row of its node; patches respect their limits; and walking the plan exactly as the kernel does --
one thread per unit in the plan's round order, each unit's row of the oracle's local matrix added through the element->CSR position
map into the row's segment of the patch's row block, the block copied to the CSR array -- reproduces the oracle matrix.
The GPU parity tests then only have to show that the kernel is a faithful interpreter with its own quadrature rows."""
import ctypes as C

import numpy as np
import pytest

import feelpp_b200 as fpp
from feelpp_b200 import _capi as capi
from feelpp_b200._capi import c_dblp, c_i32p, c_i64p, ptr
from oracle import oracle as orc

PATCH = np.dtype([("visit_off", "u4"), ("nvisit", "u4"), ("row_off", "u4"), ("nrows", "u4"), ("unit_off", "u4"),
                  ("nunits", "u4"), ("nslots", "u4"), ("pad", "u4")])
ROW = np.dtype([("base", "i8"), ("off", "u4"), ("len", "u4"), ("unit0", "u4"), ("nunits", "u4"), ("pad", "u4", 2)])


def build(Xh, ia, maxslots, maxvisits, maxrows=1 << 20):
    lib = capi.gpu()
    mesh = Xh.mesh
    plan, sizes = C.c_void_p(), np.zeros(8, dtype=np.int64)
    rc = lib.feelgpu_rplan_host_build(Xh.order, Xh.ncomp, mesh.numElements(), ptr(Xh.l2g, c_i32p), mesh.numPoints(),
                                      ptr(mesh.xyz, c_dblp), ptr(mesh.conn, c_i32p), Xh.nDof(), ptr(ia, c_i64p), maxslots,
                                      maxvisits, maxrows, C.byref(plan), ptr(sizes, c_i64p))
    if rc:
        msg = lib.feelgpu_rplan_host_error(plan).decode()
        lib.feelgpu_rplan_host_free(plan)
        raise RuntimeError(msg)
    patches, rows = np.zeros(sizes[0], dtype=PATCH), np.zeros(sizes[2], dtype=ROW)
    visits, units = np.zeros(sizes[1], dtype=np.int32), np.zeros(sizes[3], dtype=np.uint32)
    lib.feelgpu_rplan_host_copy(plan, patches.ctypes.data_as(C.c_void_p), ptr(visits, c_i32p), rows.ctypes.data_as(C.c_void_p),
                                units.ctypes.data_as(C.c_void_p))
    lib.feelgpu_rplan_host_free(plan)
    return patches, visits, rows, units, sizes


def position_map(Xh, ia, ja):
    """pos[e][i][j]: position of the (component-0) column of local node j inside the (component-0) row of local node i"""
    n, nc = Xh.n, Xh.ncomp
    pos = np.zeros((Xh.mesh.numElements(), n, n), dtype=np.int64)
    for e in range(Xh.mesh.numElements()):
        for i in range(n):
            r = Xh.l2g[e, i]
            seg = ja[ia[r]:ia[r + 1]]
            pos[e, i] = np.searchsorted(seg, Xh.l2g[e, :n])
    return pos


@pytest.mark.parametrize("order,ncomp,kind,m,maxslots,maxvisits", [(1, 1, 1, 4, 200, 60), (2, 1, 1, 3, 700, 50), (2, 1, 0, 3, 4000, 400),
                                                                   (3, 1, 1, 3, 1500, 40), (2, 3, 2, 3, 3000, 64)])
def test_plan_interpreted_on_the_host_gives_the_oracle_matrix(order, ncomp, kind, m, maxslots, maxvisits):
    mesh = fpp.unitCube(m, 0.15, True)
    Xh = fpp.FunctionSpace(mesh, order, ncomp)
    ia, ja = Xh.graph()
    n, nd = Xh.n, Xh.nd
    patches, visits, rows, units, sizes = build(Xh, ia, maxslots, maxvisits)
    assert len(patches) > 1, "the limits should cut the mesh into several patches"
    assert sizes[4] <= maxslots and sizes[5] <= maxvisits
    pos = position_map(Xh, ia, ja)
    q = orc.quad_order_form2(kind, order)
    co = orc.Coeff(lam=2.0, mu=1.5)
    a = np.full(len(ja), np.nan)
    seen_units = set()
    seen_nodes = set()
    for P in patches:
        vis = visits[P["visit_off"]:P["visit_off"] + P["nvisit"]]
        assert len(np.unique(vis)) == len(vis)
        blk = np.zeros(P["nslots"])
        used = np.zeros(P["nslots"], dtype=bool)
        prow = rows[P["row_off"]:P["row_off"] + P["nrows"]]
        assert (np.diff(prow["nunits"].astype(np.int64)) <= 0).all(), "rows are sorted by descending unit count"
        local = {}
        rownode = []
        for R in prow:
            node = int(np.searchsorted(ia, R["base"])) // ncomp
            assert ia[node * ncomp] == R["base"] and node not in seen_nodes
            seen_nodes.add(node)
            rownode.append(node)
            assert R["len"] == ia[node * ncomp + 1] - ia[node * ncomp]
            seg = slice(int(R["off"]), int(R["off"]) + int(R["len"]) * ncomp)
            assert not used[seg].any(), "row segments of a patch overlap"
            used[seg] = True
        assert int(prow["nunits"].sum()) == P["nunits"]
        per_row = np.zeros(len(prow), dtype=np.int64)
        pu = units[P["unit_off"]:P["unit_off"] + P["nunits"]]
        # round order: the lanes of a warp (32 consecutive units) work on different rows while enough rows have a k-th unit
        for k0 in range(0, len(pu), 32):
            rr = (pu[k0:k0 + 32] >> 17).astype(np.int64)
            if int((prow["nunits"] > per_row.max()).sum()) >= 64:
                assert len(np.unique(rr)) == len(rr)
            for w in pu[k0:k0 + 32]:
                v, i, r = int(w & 0xFFF), int(w >> 12) & 31, int(w >> 17)
                R, node = prow[r], rownode[r]
                per_row[r] += 1
                e = int(vis[v])
                assert Xh.l2g[e, i] // ncomp == node and (e, i) not in seen_units
                seen_units.add((e, i))
                if e not in local:
                    local[e] = orc.local_matrix(kind, order, ncomp, q, mesh.xyz[mesh.conn[e]], co)
                Ae = local[e]
                for c1 in range(ncomp):
                    for j in range(n):
                        for c2 in range(ncomp):
                            blk[int(R["off"]) + c1 * int(R["len"]) + pos[e, i, j] + c2] += Ae[n * c1 + i, n * c2 + j]
        assert np.array_equal(per_row, prow["nunits"])
        for R in prow:
            cnt = int(R["len"]) * ncomp
            a[R["base"]:R["base"] + cnt] = blk[R["off"]:R["off"] + cnt]
    assert len(seen_units) == mesh.numElements() * n and len(seen_nodes) == Xh.nDof() // ncomp
    assert not np.isnan(a).any(), "a CSR value was never written"
    ref = orc.form2(kind, order, ncomp, q, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, co)
    scale = np.abs(ref).max()
    np.testing.assert_allclose(a, ref, rtol=1e-13, atol=1e-14 * scale)


def test_a_row_that_does_not_fit_is_refused():
    mesh = fpp.unitCube(3, 0.0, False)
    Xh = fpp.FunctionSpace(mesh, 2, 1)
    ia, ja = Xh.graph()
    with pytest.raises(RuntimeError):
        build(Xh, ia, 8, 400)


def test_plan_cut_in_pieces_is_complete():
    """more than 65 536 nodes: the curve is cut into pieces whose patches are formed independently (by several threads).
    Structure of the whole plan, vectorised: every (element, local dof) unit once and in the row of its node, every node in
    one patch, row segments tile the row block, visits of a patch distinct, limits respected."""
    mesh = fpp.unitCube(22, 0.15, True)
    Xh = fpp.FunctionSpace(mesh, 2, 1)
    ia, ja = Xh.graph()
    nNode, n, nElt = Xh.nDof(), Xh.n, mesh.numElements()
    assert nNode > 65536
    maxslots, maxvisits = 12288, 384
    patches, visits, rows, units, sizes = build(Xh, ia, maxslots, maxvisits)
    assert (patches["nslots"] <= maxslots).all() and (patches["nvisit"] <= maxvisits).all()
    assert int(patches["nunits"].sum()) == len(units) == nElt * n and int(patches["nrows"].sum()) == len(rows) == nNode
    assert int(patches["nvisit"].sum()) == len(visits)
    # offsets are the running sums
    for off, cnt in (("visit_off", "nvisit"), ("row_off", "nrows"), ("unit_off", "nunits")):
        assert np.array_equal(patches[off].astype(np.int64), np.cumsum(patches[cnt].astype(np.int64)) - patches[cnt])
    pid_u = np.repeat(np.arange(len(patches)), patches["nunits"].astype(np.int64))
    pid_r = np.repeat(np.arange(len(patches)), patches["nrows"].astype(np.int64))
    pid_v = np.repeat(np.arange(len(patches)), patches["nvisit"].astype(np.int64))
    v, i, r = (units & 0xFFF).astype(np.int64), ((units >> 12) & 31).astype(np.int64), (units >> 17).astype(np.int64)
    assert (v < patches["nvisit"][pid_u]).all() and (r < patches["nrows"][pid_u]).all() and (i < n).all()
    e = visits[patches["visit_off"].astype(np.int64)[pid_u] + v].astype(np.int64)
    assert len(np.unique(e * 32 + i)) == nElt * n, "a unit occurs twice"
    row_node = np.searchsorted(ia, rows["base"])
    assert np.array_equal(ia[row_node], rows["base"]) and len(np.unique(row_node)) == nNode
    assert np.array_equal(rows["len"].astype(np.int64), ia[row_node + 1] - ia[row_node])
    grow = patches["row_off"].astype(np.int64)[pid_u] + r
    assert np.array_equal(Xh.l2g[e, i].astype(np.int64), row_node[grow]), "a unit sits in the row of another node"
    assert np.array_equal(np.bincount(grow, minlength=len(rows)), rows["nunits"].astype(np.int64))
    # row segments tile the row block of their patch in row order
    ln = rows["len"].astype(np.int64)
    excl = np.cumsum(ln) - ln
    first = patches["row_off"].astype(np.int64)[pid_r]
    assert np.array_equal(rows["off"].astype(np.int64), excl - excl[first])
    last = patches["row_off"].astype(np.int64) + patches["nrows"].astype(np.int64) - 1
    assert np.array_equal(patches["nslots"].astype(np.int64), rows["off"].astype(np.int64)[last] + ln[last])
    assert len(np.unique(pid_v * nElt + visits.astype(np.int64))) == len(visits), "an element is visited twice by a patch"


def test_plan_does_not_depend_on_the_thread_count(monkeypatch):
    mesh = fpp.unitCube(22, 0.15, True)
    Xh = fpp.FunctionSpace(mesh, 2, 1)
    ia, ja = Xh.graph()
    plans = []
    for nt in ("1", "7"):
        monkeypatch.setenv("FEELGPU_BUILD_THREADS", nt)
        plans.append(build(Xh, ia, 12288, 384))
    for a, b in zip(plans[0][:4], plans[1][:4]):
        assert a.tobytes() == b.tobytes()
