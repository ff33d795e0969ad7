"""gloo tests (world 2, 3, 4, 8) of the multi-GPU host path on CPU: partition, ownership, the reference's local numbering
(active dofs first, ghosts after, first-touch order), rank-contiguous cluster ids, local graphs and the ghost-row exchange
of matrices (MatAssemblyBegin/End) and vectors (VecAssemblyBegin/End + VecGhostUpdate).  Local assembly is done by the
oracle (this test checks the distributed bookkeeping, not the kernels); the gathered distributed matrix must equal the
sequential oracle matrix ENTRY BY ENTRY after the cluster -> sequential permutation: identical pattern, values within the
parity bound (tests/mp_util.py)."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _local_oracle_matrix(lp, kind, order, ncomp, co):
    """the rank's local value array (owned + ghost rows, cluster columns) assembled by the oracle over its own elements"""
    from oracle import oracle as orc
    ia, ja = lp.ia, lp.ja
    # the oracle wants ascending LOCAL columns: map cluster -> local (absent: -1), sort every row, remember the order
    c2l = np.full(lp.nGlobalDof, -1, dtype=np.int64)
    c2l[lp.p2c] = np.arange(lp.nDofWithGhost)
    ja_loc = c2l[ja]
    rows = np.repeat(np.arange(len(ia) - 1), np.diff(ia))
    order_in_row = np.lexsort((ja_loc, rows))
    vals = orc.form2(kind, order, ncomp, orc.quad_order_form2(kind, order), lp.mesh.xyz, lp.mesh.conn, lp.l2g, ia,
                     ja_loc[order_in_row], co)
    a = np.zeros(len(ja))
    a[order_in_row] = vals
    return a


def _worker(rank, world, port, m, order, ncomp, kind, how, out, msh=None):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    import feelpp_b200 as fpp
    from feelpp_b200 import parallel as par
    from oracle import oracle as orc
    import mp_util
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    if msh is None:
        mesh = fpp.unitCube(m, 0.15, True)
        part = mp_util.make_partition(mesh, world, how)
    else:       # a partitioned .msh file: every rank reads it and keeps the elements tagged with its own partition
        mesh = fpp.loadMesh(msh)
        part = mesh.eltPartition
        assert mesh.numPartitions == world
    lp = par.LocalProblem(mesh, part, order, ncomp, par.Comm(rank, world, dist))
    assert lp.mesh.numElements() == int((part == rank).sum())
    # the reference's order: ghosts behind the owned dofs, both groups in first-touch order of the local elements
    first_seen = np.full(lp.nDofWithGhost // ncomp, -1, dtype=np.int64)
    nodes = lp.l2g[:, :lp.l2g.shape[1] // ncomp] // ncomp
    flat = nodes.ravel()
    uniq, idx = np.unique(flat, return_index=True)
    first_seen[uniq] = idx
    nOn = lp.nOwnedNodes
    assert (np.diff(first_seen[:nOn]) > 0).all() and (np.diff(first_seen[nOn:]) > 0).all(), "not in first-touch order"
    co = orc.Coeff(lam=2.0, mu=1.5)
    a = _local_oracle_matrix(lp, kind, order, ncomp, co)
    par.exchange_numpy(lp, a, dist)
    # vector: f*id(v) / inner(f, id(v)) assembled over the local elements, closed, ghost copies refreshed
    vec = [1.0, -2.0, 0.5] if ncomp == 3 else None
    b = orc.form1(order, ncomp, orc.quad_order_form1(order, 0), lp.mesh.xyz, lp.mesh.conn, lp.l2g, lp.nDofWithGhost,
                  orc.Coeff(value=3.0, vector=vec))
    par.exchange_vector_numpy(lp, b, dist)
    payload = mp_util.rank_payload(lp, a[:lp.ia[lp.nDofOwned]], b)
    payload["ghost_check"] = (lp.p2c[lp.nDofOwned:], b[lp.nDofOwned:])      # ghost copies after VecGhostUpdate
    payload["neighbours"] = lp.num_neighbours()
    gathered = [None] * world
    dist.all_gather_object(gathered, payload)
    if rank == 0:
        out.put(gathered)
    dist.barrier()
    dist.destroy_process_group()


CASES = [(2, 8, 1, 1, 1, "rcb"), (3, 12, 2, 1, 0, "bfs1"), (4, 8, 2, 3, 2, "metis"), (8, 12, 2, 1, 1, "metis"),
         (4, 5, 3, 1, 1, "bfs2"), (8, 8, 1, 1, 1, "bfs3")]


@pytest.mark.parametrize("world,m,order,ncomp,kind,how", CASES)
def test_distributed_assembly_matches_sequential(world, m, order, ncomp, kind, how, msh=None):
    import feelpp_b200 as fpp
    from oracle import oracle as orc
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import mp_util
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + (os.getpid() * 7 + 13 * world + order + m) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, m, order, ncomp, kind, how, out, msh)) for r in range(world)]
    for p in procs:
        p.start()
    gathered = out.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    mesh = fpp.loadMesh(msh) if msh else fpp.unitCube(m, 0.15, True)
    what = "world %d P%d x%d %s" % (world, order, ncomp, how)
    mp_util.check_against_sequential(gathered, mesh, order, ncomp, kind, orc.Coeff(lam=2.0, mu=1.5), what)
    vec = [1.0, -2.0, 0.5] if ncomp == 3 else None
    ref_of = lambda Xh: orc.form1(order, ncomp, orc.quad_order_form1(order, 0), mesh.xyz, mesh.conn, Xh.l2g, Xh.nDof(),
                                  orc.Coeff(value=3.0, vector=vec))
    mp_util.check_vector_against_sequential(gathered, mesh, order, ncomp, ref_of, what + " form1")
    # VecGhostUpdate: every ghost copy holds its owner's final value
    Xh = fpp.FunctionSpace(mesh, order, ncomp)
    perm = mp_util.cluster_to_sequential(gathered, mesh, order, ncomp, Xh)
    ref = ref_of(Xh)
    for g in gathered:
        gc, gb = g["ghost_check"]
        np.testing.assert_allclose(gb, ref[perm[gc]], rtol=1e-12, atol=1e-13 * np.abs(ref).max())
    if how.startswith("bfs") and world >= 4:
        assert max(g["neighbours"] for g in gathered) >= 3, "the irregular partition should give a rank 3 neighbours"


def test_distributed_assembly_from_a_partitioned_msh_file(tmp_path):
    """the partition is an input (Feel++ reads it from the tags of the .msh file, importergmsh.hpp:1239-1242,1272-1281):
    two unequal parts written into the file, both ranks load it, the gathered matrix is the sequential one"""
    import feelpp_b200 as fpp
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_gmsh as G
    mesh = fpp.unitCube(4, 0.15, True)
    part = (mesh.xyz[mesh.conn].mean(1)[:, 0] > 0.4).astype(np.int32)
    ids = np.arange(mesh.numPoints()) * 2 + 1
    path = str(tmp_path / "cube_part.msh")
    G.write_v2(path, mesh, np.zeros((0, 3), dtype=np.int32), np.zeros(0, dtype=np.int32), ids,
               np.random.default_rng(1).permutation(mesh.numPoints()), True, part)
    test_distributed_assembly_matches_sequential(2, 4, 2, 1, 1, "file", msh=path)
