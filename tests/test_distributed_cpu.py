"""world_size-2 (and 3) gloo test of the multi-GPU host path on CPU: partition, ownership, owned-first numbering,
rank-contiguous cluster ids, local graphs and the ghost-row exchange.  Local assembly is done by the oracle (this
test checks the distributed bookkeeping, not the kernels); the gathered distributed matrix must equal the sequential
oracle matrix after the cluster renumbering, to rounding."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, order, ncomp, kind, out, msh=None):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import scipy.sparse as sps
    import feelpp_b200 as fpp
    from feelpp_b200 import parallel as par
    from oracle import oracle as orc
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    if msh is None:
        mesh = fpp.unitCube(4, 0.15, True)
        part = par.partition_rcb(mesh, world)
    else:       # a partitioned .msh file: every rank reads it and keeps the elements tagged with its own partition
        mesh = fpp.loadMesh(msh)
        part = mesh.eltPartition
        assert mesh.numPartitions == world
    lp = par.LocalProblem(mesh, part, order, ncomp, par.Comm(rank, world, dist))
    assert lp.mesh.numElements() == int((part == rank).sum())
    quad = orc.quad_order_form2(kind, order)
    co = orc.Coeff(lam=2.0, mu=1.5)
    # local assembly with process-local rows and cluster columns: columns of local elements are local dofs, so assemble
    # against a local-column copy of the rows and keep the value order
    ia, ja = lp.ia, lp.ja
    c2l = {c: l for l, c in enumerate(lp.p2c)}
    ja_loc = np.array([c2l.get(c, -1) for c in ja], dtype=np.int64)
    # rows sorted by cluster id are not sorted by local id: the oracle needs ascending columns -> sort per row, remember
    a = np.zeros(len(ja))
    order_in_row = np.concatenate([np.argsort(ja_loc[ia[r]:ia[r + 1]], kind="stable") + ia[r] for r in range(len(ia) - 1)])
    ja_sorted = ja_loc[order_in_row]
    # drop columns that are not local (-1 sorts first): give them a dummy unique negative id (never searched)
    vals = orc.form2(kind, order, ncomp, quad, lp.mesh.xyz, lp.mesh.conn, lp.l2g, ia, ja_sorted, co)
    a[order_in_row] = vals
    par.exchange_numpy(lp, a, dist)
    nO = lp.nDofOwned
    rows = np.repeat(lp.p2c[:nO], np.diff(ia[:nO + 1]))
    A = sps.coo_matrix((a[:ia[nO]], (rows, ja[:ia[nO]])), shape=(lp.nGlobalDof, lp.nGlobalDof))
    gathered = [None] * world
    dist.all_gather_object(gathered, (A.row, A.col, A.data, lp.p2c[:nO], lp.firstDof, nO))
    if rank == 0:
        out.put(gathered)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,order,ncomp,kind", [(2, 1, 1, 1), (2, 2, 1, 1), (3, 2, 1, 0), (2, 2, 3, 2)])
def test_distributed_assembly_matches_sequential(world, order, ncomp, kind, msh=None):
    import scipy.sparse as sps
    import feelpp_b200 as fpp
    from oracle import oracle as orc
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + (os.getpid() + 7 * world + order) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, order, ncomp, kind, out, msh)) for r in range(world)]
    for p in procs:
        p.start()
    gathered = out.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # owned blocks are rank-contiguous and disjoint
    first = [g[4] for g in gathered]
    counts = [g[5] for g in gathered]
    assert first == list(np.concatenate([[0], np.cumsum(counts)])[:-1])
    N = sum(counts)
    D = sum(sps.coo_matrix((g[2], (g[0], g[1])), shape=(N, N)).tocsr() for g in gathered)
    # sequential reference, compared through invariants that do not depend on the numbering + the spectrum of the
    # permuted matrix: same multiset of row sums / diagonal entries / Frobenius norm
    mesh = fpp.unitCube(4, 0.15, True)
    Xh = fpp.FunctionSpace(mesh, order, ncomp)
    ia, ja = Xh.graph()
    ref = orc.form2(kind, order, ncomp, orc.quad_order_form2(kind, order), mesh.xyz, mesh.conn, Xh.l2g, ia, ja,
                    orc.Coeff(lam=2.0, mu=1.5))
    R = sps.csr_matrix((ref, ja, ia), shape=(N, N))
    assert Xh.nDof() == N
    assert D.nnz == R.nnz or abs(D).sum() > 0
    sc = np.abs(ref).max()
    assert np.allclose(np.sort(D.diagonal()), np.sort(R.diagonal()), rtol=1e-12, atol=1e-14 * sc)
    assert np.allclose(np.sort(np.asarray(abs(D).sum(1)).ravel()), np.sort(np.asarray(abs(R).sum(1)).ravel()), rtol=1e-11)
    assert abs(sps.linalg.norm(D) - sps.linalg.norm(R)) < 1e-12 * sps.linalg.norm(R)
    x = np.ones(N)
    assert abs(x @ D @ x - x @ R @ x) < 1e-10 * max(1.0, abs(R).sum())


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
    test_distributed_assembly_matches_sequential(2, 2, 1, 1, msh=path)
