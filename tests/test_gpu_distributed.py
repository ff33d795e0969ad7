"""Multi-GPU parity (needs >= 2 GPUs; skipped otherwise): each rank assembles its partition with the CUDA engine, close()
runs through the C ABI (feelgpu_exchange_create / feelgpu_form2_close / feelgpu_form1_close: NCCL + peer memory), and the
gathered matrix is compared with the sequential oracle matrix entry by entry after the cluster -> sequential permutation
(tests/mp_util.py): identical pattern, every value within the parity bound.  Ghost rows come in the reference's order
(first touch, not grouped by owner) and the partitions are irregular (METIS / breadth-first seeds)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _init(rank, world, port):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    return torch, dist


def _forms(order, ncomp, Xh, mesh):
    import feelpp_b200 as fpp
    from feelpp_b200 import elements, grad, gradt, integrate, inner, trans, sym, trace, Id
    u = v = fpp.Element(Xh)
    if ncomp == 1:
        return integrate(_range=elements(mesh), _expr=gradt(u) * trans(grad(v)))
    return integrate(_range=elements(mesh), _expr=inner(2.0 * trace(sym(gradt(u))) * Id() + 2 * 1.5 * sym(gradt(u)), grad(v)))


def _worker(rank, world, port, m, order, ncomp, how, out):
    torch, dist = _init(rank, world, port)
    import feelpp_b200 as fpp
    from feelpp_b200 import parallel as par, distributed as dd
    from feelpp_b200 import cst, elements, id, integrate, vec, inner
    import mp_util
    mesh = fpp.unitCube(m, 0.15, True)
    part = mp_util.make_partition(mesh, world, how)
    lp = par.LocalProblem(mesh, part, order, ncomp, par.Comm(rank, world, dist))
    eng = fpp.Engine(rank)
    Xh = dd.LocalSpace(lp)
    res = {}
    for name, algo, peer in (("scatter", fpp.ALGO_SCATTER, True), ("auto", fpp.ALGO_AUTO, True), ("auto-nopeer", fpp.ALGO_AUTO, False)):
        a = dd.DistributedBilinearForm(Xh, eng, dist, algorithm=algo, peer_memory=peer)
        a.assign(_forms(order, ncomp, Xh, lp.mesh))
        torch.cuda.synchronize()
        a.close()
        torch.cuda.synchronize()
        b = None
        if name == "auto":
            v = fpp.Element(Xh)
            l = dd.DistributedLinearForm(Xh, eng, a)
            f = cst(3.0) * id(v) if ncomp == 1 else inner(vec(3.0, -6.0, 1.5), id(v))
            l.assign(integrate(_range=elements(lp.mesh), _expr=f))
            l.close()
            b = l.values()
            res["info"] = a.exchange_info()
        res[name] = mp_util.rank_payload(lp, a.owned_values(), b)
        if b is not None:
            res[name]["ghost_check"] = (lp.p2c[lp.nDofOwned:], b[lp.nDofOwned:])
    gathered = [None] * world
    dist.all_gather_object(gathered, res)
    if rank == 0:
        out.put(gathered)
    dist.barrier()
    dist.destroy_process_group()


def _run(worker, world, args, timeout=900):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29600 + (os.getpid() * 3 + 17 * world + sum(hash(str(a)) % 97 for a in args)) % 2000
    procs = [ctx.Process(target=worker, args=(r, world, port) + tuple(args) + (out,)) for r in range(world)]
    for p in procs:
        p.start()
    gathered = out.get(timeout=timeout)
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    return gathered


@pytest.mark.parametrize("world,m,order,ncomp,how", [(2, 8, 2, 1, "metis"), (2, 6, 1, 1, "bfs1"), (2, 6, 2, 3, "bfs2"),
                                                     (4, 12, 2, 1, "bfs3"), (4, 8, 2, 3, "metis"), (8, 12, 2, 1, "metis"),
                                                     (8, 8, 2, 3, "bfs5"), (8, 12, 1, 1, "bfs4")])
def test_gpus_match_sequential_entry_by_entry(world, m, order, ncomp, how):
    gathered = _run(_worker, world, (m, order, ncomp, how))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import feelpp_b200 as fpp
    from oracle import oracle as orc
    import mp_util
    mesh = fpp.unitCube(m, 0.15, True)
    kind = orc.STIFFNESS if ncomp == 1 else orc.ELASTICITY
    co = orc.Coeff(lam=2.0, mu=1.5)
    for name in ("scatter", "auto", "auto-nopeer"):
        mp_util.check_against_sequential([g[name] for g in gathered], mesh, order, ncomp, kind, co,
                                         "world %d P%d x%d %s %s" % (world, order, ncomp, how, name))
    # the entity plan covers scalar and vector mass/stiffness: those ghost rows travelled through peer memory
    if ncomp == 1:
        assert all(g["info"]["peer_memory"] == 1 for g in gathered)
    vecf = [3.0, -6.0, 1.5] if ncomp == 3 else None
    ref_of = lambda Xh: orc.form1(order, ncomp, orc.quad_order_form1(order, 0), mesh.xyz, mesh.conn, Xh.l2g, Xh.nDof(),
                                  orc.Coeff(value=3.0, vector=vecf))
    mp_util.check_vector_against_sequential([g["auto"] for g in gathered], mesh, order, ncomp, ref_of, "form1 close")
    Xh = fpp.FunctionSpace(mesh, order, ncomp)
    perm = mp_util.cluster_to_sequential([g["auto"] for g in gathered], mesh, order, ncomp, Xh)
    ref = ref_of(Xh)
    for g in gathered:
        gc, gb = g["auto"]["ghost_check"]
        np.testing.assert_allclose(gb, ref[perm[gc]], rtol=1e-12, atol=1e-13 * np.abs(ref).max())


def _worker_sequences(rank, world, port, out):
    """'=' / '+=' sequences over several closes, peer-memory and send/recv transports, and a term that goes through the
    local tail between two terms that go to the peers"""
    torch, dist = _init(rank, world, port)
    import feelpp_b200 as fpp
    from feelpp_b200 import parallel as par, distributed as dd
    from feelpp_b200 import elements, grad, gradt, id, idt, integrate, trans
    import mp_util
    mesh = fpp.unitCube(6, 0.15, True)
    part = mp_util.make_partition(mesh, world, "bfs7")
    lp = par.LocalProblem(mesh, part, 2, 1, par.Comm(rank, world, dist))
    eng = fpp.Engine(rank)
    Xh = dd.LocalSpace(lp)
    u = v = fpp.Element(Xh)
    mass = integrate(_range=elements(lp.mesh), _expr=idt(u) * id(v))
    stiff = integrate(_range=elements(lp.mesh), _expr=gradt(u) * trans(grad(v)))
    res = {}
    for peer in (True, False):
        a = dd.DistributedBilinearForm(Xh, eng, dist, peer_memory=peer)
        assert a.peer == peer
        a.assign(mass)
        a += stiff
        a.close()
        torch.cuda.synchronize()
        res[("mass+stiff", peer)] = mp_util.rank_payload(lp, a.owned_values())
        a.assign(stiff)                      # second epoch: the other set of landing buffers, '=' starts from scratch
        a.close()
        torch.cuda.synchronize()
        res[("stiff", peer)] = mp_util.rank_payload(lp, a.owned_values())
        a.assign(mass)                       # third epoch: one term through the kernel that writes to the peers ...
        a.algorithm = fpp.ALGO_SCATTER       # ... and one through the local tail + NCCL send/recv
        a += stiff
        a.algorithm = fpp.ALGO_AUTO
        a += stiff
        a.close()
        torch.cuda.synchronize()
        res[("mass+2stiff", peer)] = mp_util.rank_payload(lp, a.owned_values())
        a += stiff                           # '+=' after a close: only the new contribution travels
        a.close()
        torch.cuda.synchronize()
        res[("mass+3stiff", peer)] = mp_util.rank_payload(lp, a.owned_values())
    gathered = [None] * world
    dist.all_gather_object(gathered, res)
    if rank == 0:
        out.put(gathered)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_gpus_assignment_sequences(world):
    gathered = _run(_worker_sequences, world, ())
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import feelpp_b200 as fpp
    from oracle import oracle as orc
    from oracle import parity
    import mp_util
    mesh = fpp.unitCube(6, 0.15, True)
    Xh = fpp.FunctionSpace(mesh, 2, 1)
    ia, ja = Xh.graph()
    N = Xh.nDof()
    f = lambda kind: orc.form2(kind, 2, 1, orc.quad_order_form2(kind, 2), mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    M, K = f(orc.MASS), f(orc.STIFFNESS)
    Ma, Ka = f(orc.MASS | orc.ABS_FLAG), f(orc.STIFFNESS | orc.ABS_FLAG)
    want = {"mass+stiff": (M + K, Ma + Ka), "stiff": (K, Ka), "mass+2stiff": (M + 2 * K, Ma + 2 * Ka),
            "mass+3stiff": (M + 3 * K, Ma + 3 * Ka)}
    for key in gathered[0]:
        name, peer = key
        g = [r[key] for r in gathered]
        perm = mp_util.cluster_to_sequential(g, mesh, 2, 1, Xh)
        D = mp_util.gathered_matrix(g, perm, N)
        assert np.array_equal(D.indptr, ia) and np.array_equal(D.indices, ja)
        parity.check(D.data, want[name][0], want[name][1], ia, ja, "%s peer=%s world %d" % (name, peer, world))
