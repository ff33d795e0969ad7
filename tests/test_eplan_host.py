"""The entity-task plan (feelpp_b200/csrc/eplan_build.cpp) interpreted on the host, without a GPU.

The plan is built through the C ABI's host view (feelgpu_eplan_host_*), then executed here exactly as the kernel does
(phase A: S per visit; phase B: per-task sums over the unit words, coefficients, stores into the CTA's row block; phase C:
row block -> CSR values) and compared with the CPU oracle.  This pins the plan builder and the coefficient projection;
the GPU parity tests then only have to show that the kernel is a faithful interpreter.
"""
import ctypes as C

import numpy as np
import pytest

import feelpp_b200 as fpp
from feelpp_b200 import _capi as capi
from feelpp_b200._capi import Tables, c_dblp, c_i32p, c_i64p, ptr
from feelpp_b200 import hostmirror as hm
from oracle import oracle as orc
from oracle import parity

NLD = {1: [1, 1], 2: [1, 2, 2, 3, 5, 4]}
NOUT = {1: [1, 1], 2: [1, 2, 1, 3, 3, 1]}
CTA = np.dtype([("vert_off", "<u4"), ("visit_off", "<u4"), ("chunk_off", "<u4"), ("row_off", "<u4"), ("nverts", "<u2"),
                ("nvisit", "<u2"), ("nchunk", "<u2"), ("nrows", "<u2"), ("nout", "<u2"), ("blob_nq", "<u2"),
                ("blob_q0", "<u4")])
CHUNK = np.dtype([("q0", "<u4"), ("type", "u1"), ("niter", "u1"), ("pad", "<u2")])
ROW = np.dtype([("base", "<i8"), ("off", "<u4"), ("len", "<u4")])


def position_map(l2g, ia, ja, ncomp):
    """pos[e][i][j]: position of column l2g[e][j] in row l2g[e][i] (component 0), as k_build_pos computes it"""
    n = l2g.shape[1] // ncomp
    N = int(ia.shape[0])
    rows = np.repeat(np.arange(N - 1), np.diff(ia))
    keys = rows * N + ja
    r = l2g[:, :n, None].astype(np.int64)
    c = l2g[:, None, :n].astype(np.int64)
    idx = np.searchsorted(keys, r * N + c)
    assert (keys[idx] == r * N + c).all()
    return (idx - ia[r]).astype(np.uint16)


def build_plan(mesh, Xh, ia, ja):
    lib = capi.gpu()
    pos = np.ascontiguousarray(position_map(Xh.l2g, ia, ja, Xh.ncomp))
    h = C.c_void_p()
    sizes = np.zeros(8, dtype=np.int64)
    rc = lib.feelgpu_eplan_host_build(Xh.order, Xh.ncomp, mesh.numElements(), ptr(Xh.l2g, c_i32p), mesh.numPoints(),
                                      ptr(mesh.xyz, c_dblp), ptr(mesh.conn, c_i32p), Xh.nDof(), ptr(ia, c_i64p),
                                      pos.ctypes.data_as(C.c_void_p), C.byref(h), ptr(sizes, c_i64p))
    assert rc == 0, lib.feelgpu_eplan_host_error(h)
    P = {"ctas": np.zeros(sizes[0], dtype=CTA), "verts": np.zeros(sizes[1], dtype=np.int32),
         "visits": np.zeros(sizes[2], dtype=np.uint32), "chunks": np.zeros(sizes[3], dtype=CHUNK),
         "blob": np.zeros(sizes[4], dtype=np.uint32), "rows": np.zeros(sizes[5], dtype=ROW), "omax": int(sizes[6]),
         "vs": int(sizes[7])}
    stats = np.zeros(8)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    lib.feelgpu_eplan_host_copy(h, vp(P["ctas"]), ptr(P["verts"], c_i32p), vp(P["visits"]), vp(P["chunks"]),
                                vp(P["blob"]), vp(P["rows"]), ptr(stats, c_dblp))
    lib.feelgpu_eplan_host_free(h)
    P["stats"] = dict(zip(["tasks", "iters", "iters_padded", "visits", "outputs", "wavefronts", "wavefronts_ideal", "a1_ratio"], stats))
    return P


def coefficients(order, quad, form):
    xi, w = hm.im(quad)
    phi, dphi = hm.tabulate(order, xi)
    t = Tables()
    keep = [np.ascontiguousarray(x) for x in (w, xi, phi, dphi)]
    t.nq, t.n = len(w), phi.shape[0]
    t.w, t.xi, t.phi, t.dphi = (ptr(k, c_dblp) for k in keep)
    coef, mass = np.zeros((6, 3, 5)), np.zeros((6, 3))
    rc = capi.gpu().feelgpu_eplan_host_coefficients(order, C.byref(t), form, ptr(coef, c_dblp), ptr(mass, c_dblp))
    assert rc == 0
    return coef, mass


def visit_records(xyz):
    """S (10 per visit: diagonals, then (0,1),(0,2),(0,3),(1,2),(1,3),(2,3)) and J, as phase A1 of the kernel"""
    K = 0.5 * (xyz[:, 1:4, :] - xyz[:, 0:1, :]).transpose(0, 2, 1)    # K[r][b] = (x_{b+1}-x_0)_r / 2
    det = np.linalg.det(K)
    Binv = np.linalg.inv(K)                                          # rows b: grad xi_b
    h = np.zeros((xyz.shape[0], 4, 3))
    h[:, 1:4, :] = 0.5 * Binv
    h[:, 0, :] = -h[:, 1:4, :].sum(axis=1)
    J = np.abs(det)
    G = np.einsum("vkc,vlc->vkl", h, h) * J[:, None, None]
    pairs = [(0, 0), (1, 1), (2, 2), (3, 3), (0, 1), (0, 2), (0, 3), (1, 2), (1, 3), (2, 3)]
    return np.stack([G[:, k, l] for k, l in pairs], axis=1), J


def interpret(P, mesh, order, nnz, form, coef, mass):
    a = np.full(nnz, np.nan)
    vs = P["vs"]
    written = np.zeros(nnz, dtype=np.int32)
    for ct in P["ctas"]:
        verts = P["verts"][ct["vert_off"]:ct["vert_off"] + ct["nverts"]]
        vst = 4 if order == 2 else 1            # words per visit
        vrec = P["visits"][ct["visit_off"]:ct["visit_off"] + vst * int(ct["nvisit"])].reshape(-1, vst)
        vw = vrec[:, 0]
        lv = np.stack([(vw >> (8 * k)) & 255 for k in range(4)], axis=1)
        assert lv.max(initial=0) < ct["nverts"]
        S, J = visit_records(mesh.xyz[verts[lv]])
        rec = np.zeros((vs, 10))
        rec[:ct["nvisit"]] = S if form == 1 else J[:, None]
        nout, needzero = int(ct["nout"]) & 0x7FFF, int(ct["nout"]) >> 15
        out = np.zeros(P["omax"]) if needzero else np.full(P["omax"], np.nan)   # the kernel skips the zero fill
        hit = np.zeros(P["omax"], dtype=np.int32)
        if order == 2 and ct["nvisit"]:
            # phase A: (edge, opposite edge) entries, one element each; slot per original local edge of the row, 0 = trash
            if form == 1:
                s01, s02, s03, s12, s13, s23 = (S[:, k] for k in range(4, 10))
                x5 = coef[5][0][0]
                t = [x5 * ((s01 + s13) + (s02 + s23)), x5 * ((s12 + s23) + (s01 + s03)), x5 * ((s02 + s03) + (s12 + s13))]
            else:
                t = [mass[5][0] * J] * 3
            for le in range(6):
                pos = (vrec[:, 1 + le // 2] >> (16 * (le & 1))) & 0xFFFF
                live = pos != 0
                out[pos[live]] = t[le % 3][live]
                np.add.at(hit, pos[live], 1)
        for ch in P["chunks"][ct["chunk_off"]:ct["chunk_off"] + ct["nchunk"]]:
            ty, niter, nsub = int(ch["type"]), int(ch["niter"]), int(ch["pad"])
            nld, no = (NLD[order][ty] if form == 1 else 1), NOUT[order][ty]
            w0 = int(ch["q0"]) * 4
            npair = (niter + 1) // 2            # words of two consecutive iterations side by side per lane
            raw = P["blob"][w0:w0 + nsub * npair * 64].reshape(nsub, npair, 32, 2)
            allw = raw.transpose(0, 1, 3, 2).reshape(nsub, 2 * npair, 32)
            p0 = w0 + nsub * npair * 64
            allp = P["blob"][p0:p0 + nsub * no * 16].view(np.uint16).reshape(nsub, no, 32)
            for j in range(nsub):
                words, opos = allw[j], allp[j]
                slot = (words >> 3) & 0x1FF      # stored times 8
                sums = np.zeros((nld, 32))
                for t in range(niter):           # plan order
                    for k in range(nld):
                        idx = (words[t] >> (12 + 4 * k)) & 15 if form == 1 else 0
                        sums[k] += rec[slot[t], idx]
                for o in range(no):
                    v = mass[ty][o] * sums[0] if form == 0 else sum(coef[ty][o][k] * sums[k] for k in range(nld))
                    live = opos[o] != 0
                    out[opos[o][live]] = v[live]
                    np.add.at(hit, opos[o][live], 1)
        # slots of the row block: row segments (each written exactly once unless the block is zero-filled first) and the
        # one-slot pads that keep a segment on the 16-byte phase of its CSR position (never written, never copied)
        inrow = np.zeros(P["omax"], dtype=bool)
        for rw in P["rows"][ct["row_off"]:ct["row_off"] + ct["nrows"]]:
            assert not inrow[int(rw["off"]):int(rw["off"]) + int(rw["len"])].any(), "row segments overlap"
            inrow[int(rw["off"]):int(rw["off"]) + int(rw["len"])] = True
        assert not inrow[0] and not inrow[nout:].any() and hit[~inrow][1:].max(initial=0) == 0
        assert hit[inrow].max(initial=1) == 1 and (needzero or hit[inrow].min(initial=1) == 1), "every slot once"
        for rw in P["rows"][ct["row_off"]:ct["row_off"] + ct["nrows"]]:
            b, o, ln = int(rw["base"]), int(rw["off"]), int(rw["len"])
            a[b:b + ln] = out[o:o + ln]
            written[b:b + ln] += 1
    assert (written == 1).all(), "every CSR value is written exactly once"
    return a


@pytest.mark.parametrize("order", [1, 2])
@pytest.mark.parametrize("mk", ["reg", "jit", "scr"])
def test_plan_interpreted_on_host_matches_oracle(order, mk):
    mesh = fpp.unitCube(4) if mk == "reg" else fpp.unitCube(5, 0.15, True)
    if mk == "scr":
        from conftest import scrambled
        mesh = scrambled(mesh)
    Xh = fpp.Pch(mesh, order)
    ia, ja = Xh.graph()
    P = build_plan(mesh, Xh, ia, ja)
    nnz = int(ia[-1])
    for form, kind in ((1, orc.STIFFNESS), (0, orc.MASS)):
        quad = orc.quad_order_form2(kind, order)
        coef, mass = coefficients(order, quad, form)
        a = interpret(P, mesh, order, nnz, form, coef, mass)
        ref = orc.form2(kind, order, 1, quad, mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
        scale = orc.form2(kind | orc.ABS_FLAG, order, 1, quad, mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
        parity.check(a, ref, scale, ia, ja, "eplan host P%d form %d" % (order, form))
    st = P["stats"]
    assert st["outputs"] == nnz
    assert st["wavefronts"] >= st["wavefronts_ideal"] > 0


def test_plan_limits_split_ctas():
    """a mesh larger than one CTA's limits is split; every row belongs to exactly one CTA"""
    mesh = fpp.unitCube(8, 0.1, False)
    Xh = fpp.Pch(mesh, 2)
    ia, ja = Xh.graph()
    P = build_plan(mesh, Xh, ia, ja)
    assert len(P["ctas"]) > 1
    assert int(P["rows"]["len"].sum()) == int(ia[-1]) and int(P["ctas"]["nrows"].sum()) <= Xh.nDof()
    assert (P["chunks"]["type"] != 5).all()
    assert (P["ctas"]["nvisit"] < P["vs"]).all() and (P["ctas"]["nverts"] <= 256).all()
    assert ((P["ctas"]["nout"] & 0x7FFF) <= P["omax"]).all() and (P["ctas"]["nrows"] <= 384).all()


def test_plan_packed_in_pieces_matches_oracle():
    """more than 65 536 nodes: the curve is cut into pieces that are packed into CTAs independently (by several threads);
    the plan interpreted on the host still gives the oracle matrix and writes every CSR value once"""
    mesh = fpp.unitCube(22, 0.15, True)
    Xh = fpp.Pch(mesh, 2)
    assert Xh.nDof() > 65536
    ia, ja = Xh.graph()
    P = build_plan(mesh, Xh, ia, ja)
    nnz = int(ia[-1])
    quad = orc.quad_order_form2(orc.STIFFNESS, 2)
    coef, mass = coefficients(2, quad, 1)
    a = interpret(P, mesh, 2, nnz, 1, coef, mass)
    ref = orc.form2(orc.STIFFNESS, 2, 1, quad, mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    scale = orc.form2(orc.STIFFNESS | orc.ABS_FLAG, 2, 1, quad, mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    parity.check(a, ref, scale, ia, ja, "eplan host, packed in pieces")
    assert P["stats"]["outputs"] == nnz


def test_plan_does_not_depend_on_the_thread_count(monkeypatch):
    """the builder's stages run on all threads, the pieces of the packing pass depend on the mesh only: the plan of a
    mesh of several pieces is the same byte for byte with one builder thread and with seven"""
    mesh = fpp.unitCube(22, 0.15, True)
    Xh = fpp.Pch(mesh, 2)
    ia, ja = Xh.graph()
    plans = []
    for nt in ("1", "7"):
        monkeypatch.setenv("FEELGPU_BUILD_THREADS", nt)
        plans.append(build_plan(mesh, Xh, ia, ja))
    for k in ("ctas", "verts", "visits", "chunks", "blob", "rows"):
        assert plans[0][k].tobytes() == plans[1][k].tobytes(), k
