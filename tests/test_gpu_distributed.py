"""Multi-GPU parity (needs >= 2 GPUs; skipped otherwise): each rank assembles its partition with the CUDA engine,
ghost rows are exchanged with NCCL, and the gathered matrix is compared with the sequential oracle matrix."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, order, ncomp, out):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import feelpp_b200 as fpp
    from feelpp_b200 import parallel as par, distributed as dd
    from feelpp_b200 import elements, grad, gradt, integrate, inner, trans, sym, trace, Id
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    mesh = fpp.unitCube(6, 0.15, True)
    part = par.partition_rcb(mesh, world)
    lp = par.LocalProblem(mesh, part, order, ncomp, par.Comm(rank, world, dist))
    eng = fpp.Engine(rank)
    Xh = dd.LocalSpace(lp)
    u = v = fpp.Element(Xh)
    res = {}
    for algo in (fpp.ALGO_SCATTER, fpp.ALGO_AUTO):
        a = dd.DistributedBilinearForm(Xh, eng, dist, algorithm=algo)
        if ncomp == 1:
            a.assign(integrate(_range=elements(lp.mesh), _expr=gradt(u) * trans(grad(v))))
        else:
            a.assign(integrate(_range=elements(lp.mesh),
                               _expr=inner(2.0 * trace(sym(gradt(u))) * Id() + 2 * 1.5 * sym(gradt(u)), grad(v))))
        torch.cuda.synchronize()
        a.close()
        torch.cuda.synchronize()
        nO = lp.nDofOwned
        rows = np.repeat(lp.p2c[:nO], np.diff(lp.ia[:nO + 1]))
        res[algo] = (rows, lp.ja[:lp.ia[nO]], a.owned_values())
    gathered = [None] * world
    dist.all_gather_object(gathered, (res, lp.firstDof, lp.nDofOwned))
    if rank == 0:
        out.put(gathered)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("order,ncomp", [(2, 1), (1, 1), (2, 3)])
def test_two_gpus_match_sequential(order, ncomp):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import scipy.sparse as sps
    import torch.multiprocessing as mp
    import feelpp_b200 as fpp
    from oracle import oracle as orc
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29600 + (os.getpid() + order + ncomp) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, order, ncomp, out)) for r in range(world)]
    for p in procs:
        p.start()
    gathered = out.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    N = sum(g[2] for g in gathered)
    mesh = fpp.unitCube(6, 0.15, True)
    Xh = fpp.FunctionSpace(mesh, order, ncomp)
    ia, ja = Xh.graph()
    kind = orc.STIFFNESS if ncomp == 1 else orc.ELASTICITY
    ref = orc.form2(kind, order, ncomp, orc.quad_order_form2(kind, order), mesh.xyz, mesh.conn, Xh.l2g, ia, ja,
                    orc.Coeff(lam=2.0, mu=1.5))
    R = sps.csr_matrix((ref, ja, ia), shape=(N, N))
    for algo in gathered[0][0]:
        D = sum(sps.coo_matrix((g[0][algo][2], (g[0][algo][0], g[0][algo][1])), shape=(N, N)).tocsr() for g in gathered)
        sc = np.abs(ref).max()
        assert np.allclose(np.sort(D.diagonal()), np.sort(R.diagonal()), rtol=1e-12, atol=1e-14 * sc)
        assert np.allclose(np.sort(np.asarray(abs(D).sum(1)).ravel()), np.sort(np.asarray(abs(R).sum(1)).ravel()), rtol=1e-11)
        assert abs(sps.linalg.norm(D) - sps.linalg.norm(R)) < 1e-12 * sps.linalg.norm(R)


def _worker_sequences(rank, world, port, out):
    """peer-memory transport: '=' / '+=' sequences, two epochs, and a term that goes through the local tail"""
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import feelpp_b200 as fpp
    from feelpp_b200 import parallel as par, distributed as dd
    from feelpp_b200 import elements, grad, gradt, id, idt, integrate, trans
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    mesh = fpp.unitCube(6, 0.15, True)
    part = par.partition_rcb(mesh, world)
    lp = par.LocalProblem(mesh, part, 2, 1, par.Comm(rank, world, dist))
    eng = fpp.Engine(rank)
    Xh = dd.LocalSpace(lp)
    u = v = fpp.Element(Xh)
    mass = integrate(_range=elements(lp.mesh), _expr=idt(u) * id(v))
    stiff = integrate(_range=elements(lp.mesh), _expr=gradt(u) * trans(grad(v)))
    nO = lp.nDofOwned
    rows = np.repeat(lp.p2c[:nO], np.diff(lp.ia[:nO + 1]))
    res = {}
    for peer in (True, False):
        a = dd.DistributedBilinearForm(Xh, eng, dist, peer_memory=peer)
        assert a.peer == peer
        a.assign(mass)
        a += stiff
        a.close()
        torch.cuda.synchronize()
        res[("mass+stiff", peer)] = a.owned_values()
        a.assign(stiff)                      # second epoch: the other set of landing buffers, '=' starts from scratch
        a.close()
        torch.cuda.synchronize()
        res[("stiff", peer)] = a.owned_values()
        a.assign(mass)                       # third epoch: one term through the kernel that writes to the peers ...
        a.algorithm = fpp.ALGO_SCATTER       # ... and one through the local tail + NCCL send/recv
        a += stiff
        a.algorithm = fpp.ALGO_AUTO
        a += stiff
        a.close()
        torch.cuda.synchronize()
        res[("mass+2stiff", peer)] = a.owned_values()
    gathered = [None] * world
    dist.all_gather_object(gathered, (res, rows, lp.ja[:lp.ia[nO]], lp.nDofOwned))
    if rank == 0:
        out.put(gathered)
    dist.barrier()
    dist.destroy_process_group()


def test_two_gpus_peer_memory_sequences():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import scipy.sparse as sps
    import torch.multiprocessing as mp
    import feelpp_b200 as fpp
    from oracle import oracle as orc
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29600 + (os.getpid() + 77) % 2000
    procs = [ctx.Process(target=_worker_sequences, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    gathered = out.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    N = sum(g[3] for g in gathered)
    mesh = fpp.unitCube(6, 0.15, True)
    Xh = fpp.FunctionSpace(mesh, 2, 1)
    ia, ja = Xh.graph()
    M = orc.form2(orc.MASS, 2, 1, orc.quad_order_form2(orc.MASS, 2), mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    K = orc.form2(orc.STIFFNESS, 2, 1, orc.quad_order_form2(orc.STIFFNESS, 2), mesh.xyz, mesh.conn, Xh.l2g, ia, ja)
    want = {"mass+stiff": M + K, "stiff": K, "mass+2stiff": M + 2 * K}
    for (name, peer) in gathered[0][0]:
        ref = want[name]
        R = sps.csr_matrix((ref, ja, ia), shape=(N, N))
        D = sum(sps.coo_matrix((g[0][(name, peer)], (g[1], g[2])), shape=(N, N)).tocsr() for g in gathered)
        sc = np.abs(ref).max()
        assert np.allclose(np.sort(D.diagonal()), np.sort(R.diagonal()), rtol=1e-12, atol=1e-14 * sc), (name, peer)
        assert np.allclose(np.sort(np.asarray(abs(D).sum(1)).ravel()), np.sort(np.asarray(abs(R).sum(1)).ravel()),
                           rtol=1e-11), (name, peer)
        assert abs(sps.linalg.norm(D) - sps.linalg.norm(R)) < 1e-12 * sps.linalg.norm(R), (name, peer)
    # both transports give the same owned values bit for bit where the summation order is the same
    for g in gathered:
        assert np.array_equal(g[0][("stiff", True)], g[0][("stiff", False)])
