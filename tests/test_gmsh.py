"""loadMesh / markedfaces: the .msh reader on the caller side of the path (feelfilters/importergmsh.hpp:983-1420).
The files are written here in the four encodings the reader takes (MSH 2.2 and MSH 4.1, ASCII and binary) from
the synthetic cube, with non-contiguous node numbers listed out of order; what comes back must be the same mesh, the
same DofTable / graph (bit-exact against the oracle) and the marked sides of the cube."""
import os
import struct

import numpy as np
import pytest

import feelpp_b200 as fpp
from oracle import oracle as orc

F2P = np.array([[1, 2, 3], [0, 2, 3], [0, 1, 3], [0, 1, 2]])
SIDES = {"x0": 1, "x1": 2, "y0": 3, "y1": 4, "z0": 5, "z1": 6}


def cube_with_sides(m=3):
    mesh = fpp.unitCube(m, 0.0, True)
    e, f = fpp.boundaryfaces(mesh).faces()
    tri = mesh.conn[e[:, None], F2P[f]]
    c = mesh.xyz[tri].mean(1)
    marker = np.zeros(len(tri), dtype=np.int32)
    for ax in range(3):
        marker[np.abs(c[:, ax]) < 1e-12] = 1 + 2 * ax
        marker[np.abs(c[:, ax] - 1) < 1e-12] = 2 + 2 * ax
    assert marker.min() >= 1
    rng = np.random.default_rng(3)
    ids = np.cumsum(rng.integers(1, 4, mesh.numPoints())) + 10        # strictly increasing, with holes
    return mesh, tri, marker, ids, rng.permutation(mesh.numPoints())


def write_v2(path, mesh, tri, marker, ids, perm, binary, part=None):
    with open(path, "wb") as fh:
        fh.write(b"$MeshFormat\n2.2 %d 8\n" % (1 if binary else 0))
        if binary:
            fh.write(struct.pack("i", 1) + b"\n")
        fh.write(b"$EndMeshFormat\n$PhysicalNames\n%d\n" % (len(SIDES) + 1))
        for name, tag in SIDES.items():
            fh.write(b'2 %d "%s"\n' % (tag, name.encode()))
        fh.write(b'3 7 "Omega"\n$EndPhysicalNames\n$Nodes\n%d\n' % mesh.numPoints())
        for i in perm:
            if binary:
                fh.write(struct.pack("=i3d", int(ids[i]), *mesh.xyz[i]))
            else:
                fh.write(b"%d %.17g %.17g %.17g\n" % (ids[i], *mesh.xyz[i]))
        if binary:
            fh.write(b"\n")
        fh.write(b"$EndNodes\n$Elements\n%d\n" % (len(tri) + mesh.numElements() + 1))
        num = 1
        if binary:
            fh.write(struct.pack("3i", 15, 1, 2) + struct.pack("4i", num, 0, 1, int(ids[0])))
            num += 1
            fh.write(struct.pack("3i", 2, len(tri), 2))
            for t, mk in zip(tri, marker):
                fh.write(struct.pack("6i", num, int(mk), int(mk) + 100, *[int(ids[v]) for v in t]))
                num += 1
            fh.write(struct.pack("3i", 4, mesh.numElements(), 2 if part is None else 4))
            for k, t in enumerate(mesh.conn):
                tags = [7, 1] if part is None else [7, 1, 1, int(part[k]) + 1]
                fh.write(struct.pack("%di" % (5 + len(tags)), num, *tags, *[int(ids[v]) for v in t]))
                num += 1
            fh.write(b"\n")
        else:
            fh.write(b"%d 15 2 0 1 %d\n" % (num, ids[0]))
            num += 1
            for t, mk in zip(tri, marker):
                fh.write(b"%d 2 2 %d %d %d %d %d\n" % (num, mk, mk + 100, *ids[t]))
                num += 1
            for k, t in enumerate(mesh.conn):
                if part is None:
                    fh.write(b"%d 4 2 7 1 %d %d %d %d\n" % (num, *ids[t]))
                else:      # physical, elementary, number of partitions, owner (1-based), one ghost partition (negative)
                    fh.write(b"%d 4 5 7 1 2 %d %d %d %d %d %d\n" % (num, part[k] + 1, -(2 - part[k]), *ids[t]))
                num += 1
        fh.write(b"$EndElements\n")


def write_v4(path, mesh, tri, marker, ids, perm):
    with open(path, "w") as fh:
        fh.write("$MeshFormat\n4.1 0 8\n$EndMeshFormat\n$PhysicalNames\n%d\n" % (len(SIDES) + 1))
        for name, tag in SIDES.items():
            fh.write('2 %d "%s"\n' % (tag, name))
        fh.write('3 7 "Omega"\n$EndPhysicalNames\n')
        # entities: six surfaces 101..106 carrying physical 1..6, one volume 1 carrying physical 7
        fh.write("$Entities\n0 0 6 1\n")
        for s in range(6):
            fh.write("%d 0 0 0 1 1 1 1 %d 0\n" % (101 + s, 1 + s))
        fh.write("1 0 0 0 1 1 1 1 7 6 101 102 103 104 105 106\n$EndEntities\n")
        half = len(perm) // 2
        fh.write("$Nodes\n2 %d %d %d\n" % (len(perm), ids.min(), ids.max()))
        for blk, (dim, tag) in zip((perm[:half], perm[half:]), ((2, 101), (3, 1))):
            fh.write("%d %d 0 %d\n" % (dim, tag, len(blk)))
            for i in blk:
                fh.write("%d\n" % ids[i])
            for i in blk:
                fh.write("%.17g %.17g %.17g\n" % tuple(mesh.xyz[i]))
        fh.write("$EndNodes\n$Elements\n7 %d 1 %d\n" % (len(tri) + mesh.numElements(), len(tri) + mesh.numElements()))
        num = 1
        for s in range(6):
            sel = np.nonzero(marker == 1 + s)[0]
            fh.write("2 %d 2 %d\n" % (101 + s, len(sel)))
            for k in sel:
                fh.write("%d %d %d %d\n" % (num, *ids[tri[k]]))
                num += 1
        fh.write("3 1 4 %d\n" % mesh.numElements())
        for t in mesh.conn:
            fh.write("%d %d %d %d %d\n" % (num, *ids[t]))
            num += 1
        fh.write("$EndElements\n")


def write_v4_binary(path, mesh, tri, marker, ids, perm):
    Q, I = "=Q", "=i"
    with open(path, "wb") as fh:
        fh.write(b"$MeshFormat\n4.1 1 8\n" + struct.pack(I, 1) + b"\n$EndMeshFormat\n$PhysicalNames\n%d\n" % (len(SIDES) + 1))
        for name, tag in SIDES.items():
            fh.write(b'2 %d "%s"\n' % (tag, name.encode()))
        fh.write(b'3 7 "Omega"\n$EndPhysicalNames\n$Entities\n')
        fh.write(struct.pack("=4Q", 1, 0, 6, 1))
        fh.write(struct.pack("=i3dQ", 900, 0.0, 0.0, 0.0, 0))                       # a point without physical tags
        for sd in range(6):
            fh.write(struct.pack("=i6dQiQ", 101 + sd, 0, 0, 0, 1, 1, 1, 1, 1 + sd, 0))
        fh.write(struct.pack("=i6dQiQ6i", 1, 0, 0, 0, 1, 1, 1, 1, 7, 6, 101, 102, 103, 104, 105, 106))
        fh.write(b"\n$EndEntities\n$Unused\n" + struct.pack("=3d", 1.0, 2.0, 3.0) + b"\n$EndUnused\n$Nodes\n")
        half = len(perm) // 2
        fh.write(struct.pack("=4Q", 2, len(perm), int(ids.min()), int(ids.max())))
        for blk, (dim, tag) in zip((perm[:half], perm[half:]), ((2, 101), (3, 1))):
            fh.write(struct.pack("=3iQ", dim, tag, 0, len(blk)))
            fh.write(np.asarray(ids[blk], dtype="<u8").tobytes())
            fh.write(np.ascontiguousarray(mesh.xyz[blk], dtype="<f8").tobytes())
        ntot = len(tri) + mesh.numElements()
        fh.write(b"\n$EndNodes\n$Elements\n" + struct.pack("=4Q", 7, ntot, 1, ntot))
        num = 1
        for sd in range(6):
            sel = np.nonzero(marker == 1 + sd)[0]
            fh.write(struct.pack("=3iQ", 2, 101 + sd, 2, len(sel)))
            for k in sel:
                fh.write(struct.pack("=4Q", num, *[int(x) for x in ids[tri[k]]]))
                num += 1
        fh.write(struct.pack("=3iQ", 3, 1, 4, mesh.numElements()))
        for t in mesh.conn:
            fh.write(struct.pack("=5Q", num, *[int(x) for x in ids[t]]))
            num += 1
        fh.write(b"\n$EndElements\n")


@pytest.fixture(scope="module")
def cube():
    return cube_with_sides(3)


@pytest.mark.parametrize("enc", ["v2-ascii", "v2-binary", "v4-ascii", "v4-binary"])
def test_reader_returns_the_mesh_dofs_and_sides(tmp_path, cube, enc):
    mesh, tri, marker, ids, perm = cube
    path = str(tmp_path / "cube.msh")
    if enc == "v4-ascii":
        write_v4(path, mesh, tri, marker, ids, perm)
    elif enc == "v4-binary":
        write_v4_binary(path, mesh, tri, marker, ids, perm)
    else:
        write_v2(path, mesh, tri, marker, ids, perm, enc == "v2-binary")
    got = fpp.loadMesh(path)
    assert got.binary == enc.endswith("binary") and got.version == (4.1 if enc.startswith("v4") else 2.2)
    assert np.array_equal(got.pointId, ids)                     # ascending node numbers -> compact indices
    assert np.array_equal(got.xyz, mesh.xyz)                    # %.17g round-trips doubles
    assert np.array_equal(got.conn, mesh.conn) and (got.eltMarker == 7).all()
    assert got.markerNames["x0"] == (2, 1) and got.markerNames["Omega"] == (3, 7)
    assert sorted(map(tuple, np.sort(got.tri, 1))) == sorted(map(tuple, np.sort(tri, 1)))
    # DofTable and graph of the loaded mesh: bit-exact against the oracle (and so against the synthetic path)
    for order in (1, 2, 3):
        Xh = fpp.Pch(got, order)
        l2g, nd = orc.doftable(got.numPoints(), got.conn, order)
        assert nd == Xh.nDof() and np.array_equal(l2g, Xh.l2g)
    ia, ja = fpp.Pch(got, 2).graph()
    oia, oja = orc.graph(fpp.Pch(got, 2).l2g, fpp.Pch(got, 2).nDof())
    assert np.array_equal(ia, oia) and np.array_equal(ja, oja)
    # markedfaces: each side of the cube has area 1 and 2*m*m faces; by name, by number, by list
    allf = set(zip(*fpp.boundaryfaces(got).faces()))
    seen = set()
    for name, tag in SIDES.items():
        e, f = fpp.markedfaces(got, name).faces()
        assert len(e) == 2 * 3 * 3 and np.array_equal(e, np.sort(e))
        p = got.xyz[got.conn[e[:, None], F2P[f]]]
        area = 0.5 * np.linalg.norm(np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]), axis=1).sum()
        assert abs(area - 1) < 1e-13
        ax, side = (tag - 1) // 2, (tag - 1) % 2
        assert np.abs(p[..., ax] - side).max() < 1e-14
        e2, f2 = fpp.markedfaces(got, tag).faces()
        assert np.array_equal(e, e2) and np.array_equal(f, f2)
        seen |= set(zip(e, f))
    assert seen == allf
    e, f = fpp.markedfaces(got, ["x0", 2]).faces()
    assert len(e) == 4 * 3 * 3
    # the oracle's face form over a marked range: int_{x=0} 1 * phi_i sums to the area of the side
    Xh = fpp.Pch(got, 2)
    e, f = fpp.markedfaces(got, "x0").faces()
    b = orc.form1_faces(2, 1, 2, got.xyz, got.conn, Xh.l2g, Xh.nDof(), e, f)
    assert abs(b.sum() - 1) < 1e-13
    on_side = np.abs(Xh.dofPoints()[:, 0]) < 1e-14
    assert np.abs(b[~on_side]).max() == 0
    # on(markedfaces): the Dirichlet dofs are the nodes of that side only
    assert np.array_equal(np.sort(Xh.boundaryDofs((e, f))), np.nonzero(on_side)[0])


@pytest.mark.parametrize("binary", [False, True])
def test_partition_tags_of_the_file(tmp_path, cube, binary):
    mesh, tri, marker, ids, perm = cube
    part = (mesh.xyz[mesh.conn].mean(1)[:, 0] > 0.4).astype(np.int32)     # two unequal parts, not the bisection
    path = str(tmp_path / "cube.msh")
    write_v2(path, mesh, tri, marker, ids, perm, binary, part)
    got = fpp.loadMesh(path)
    assert got.numPartitions == 2 and np.array_equal(got.eltPartition, part) and np.array_equal(got.conn, mesh.conn)
    write_v2(path, mesh, tri, marker, ids, perm, binary)
    got = fpp.loadMesh(path)
    assert got.numPartitions == 1 and not got.eltPartition.any()


def test_elementary_tag_as_marker(tmp_path, cube):
    mesh, tri, marker, ids, perm = cube
    path = str(tmp_path / "cube.msh")
    write_v2(path, mesh, tri, marker, ids, perm, False)
    got = fpp.loadMesh(path, elementaryAsPhysical=True)     # importergmsh.hpp:1268-1271
    assert (got.eltMarker == 1).all() and set(got.triMarker) == {101, 102, 103, 104, 105, 106}


def test_reader_refuses_what_the_path_cannot_take(tmp_path):
    p = tmp_path / "bad.msh"
    with pytest.raises(RuntimeError, match="cannot open"):
        fpp.loadMesh(str(tmp_path / "missing.msh"))
    p.write_text("$MeshFormat\n1.0 0 8\n$EndMeshFormat\n")
    with pytest.raises(RuntimeError, match="version >= 2"):
        fpp.loadMesh(str(p))
    head = "$MeshFormat\n2.2 0 8\n$EndMeshFormat\n$Nodes\n4\n1 0 0 0\n2 1 0 0\n3 0 1 0\n4 0 0 1\n$EndNodes\n$Elements\n1\n"
    p.write_text(head + "1 4 2 1 1 1 2 3 9\n$EndElements\n")
    with pytest.raises(RuntimeError, match="unknown node 9"):
        fpp.loadMesh(str(p))
    p.write_text(head + "1 11 2 1 1 1 2 3 4 1 2 3 4 1 2\n$EndElements\n")
    with pytest.raises(RuntimeError, match="element type 11"):
        fpp.loadMesh(str(p))
    p.write_text(head + "1 2 2 1 1 1 2 3\n$EndElements\n")
    with pytest.raises(RuntimeError, match="no tetrahedra"):
        fpp.loadMesh(str(p))
    p.write_text(head + "1 4 2 1 1 1 2 3 4\n")
    with pytest.raises(RuntimeError, match="EndElements"):
        fpp.loadMesh(str(p))
    p.write_text(head + "1 4 2 1 1 1 2 3 4\n$EndElements\n")
    one = fpp.loadMesh(str(p))
    assert one.numElements() == 1 and one.numPoints() == 4 and len(one.tri) == 0
    with pytest.raises(KeyError):
        fpp.markedfaces(one, "nope").faces()


@pytest.mark.parametrize("v4", [False, True])
def test_truncated_binary_file_is_refused(tmp_path, cube, v4):
    mesh, tri, marker, ids, perm = cube
    path = str(tmp_path / "cube.msh")
    if v4:
        write_v4_binary(path, mesh, tri, marker, ids, perm)
    else:
        write_v2(path, mesh, tri, marker, ids, perm, True)
    data = open(path, "rb").read()
    open(path, "wb").write(data[:len(data) // 2])
    with pytest.raises(RuntimeError, match="truncated|expected|bad"):
        fpp.loadMesh(path)
