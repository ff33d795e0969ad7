"""The partitioned DofTable / DataMap / graph of the harness (feelpp_b200/parallel.py: LocalProblem) against the oracle's
restatement of the reference in MPI mode (oracle_doftable_mpi / oracle_graph_mpi: feeldiscr/doftablempi.hpp:138-200,548,
593-636, feeldiscr/doftable.hpp:1975-1992, feelalg/graphcsr.cpp:800-951): local-to-global table, owned count, process ->
cluster map, ghost owners and the local CSR pattern, BIT-EXACT, for every rank of METIS / breadth-first / bisection
partitions.  The ranks run as threads of this process over an in-memory all-gather."""
import threading

import numpy as np
import pytest

import feelpp_b200 as fpp
from feelpp_b200 import parallel as par
from oracle import oracle as orc


class ThreadComm:
    """all-gather between threads: what par.Comm does over torch.distributed"""

    def __init__(self, size):
        self.size, self.slots, self.bar = size, [None] * size, threading.Barrier(size)

    def view(self, rank):
        outer = self

        class View:
            size = outer.size

            def __init__(self):
                self.rank = rank

            def allgather(self, obj):
                outer.slots[rank] = obj
                outer.bar.wait()
                out = list(outer.slots)
                outer.bar.wait()
                return out
        return View()


def local_problems(mesh, part, world, order, ncomp):
    comm = ThreadComm(world)
    out, err = [None] * world, []

    def run(r):
        try:
            out[r] = par.LocalProblem(mesh, part, order, ncomp, comm.view(r))
        except Exception as e:      # a dead rank must not leave the others at the barrier
            err.append(e)
            comm.bar.abort()
    th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    if err:
        raise err[0]
    return out


@pytest.mark.parametrize("world,m,order,ncomp,how", [(2, 4, 1, 1, "rcb"), (3, 4, 2, 1, "bfs1"), (4, 4, 2, 3, "metis"),
                                                     (5, 3, 3, 1, "bfs2"), (8, 5, 2, 1, "metis")])
def test_local_problem_is_the_reference_numbering_bit_exact(world, m, order, ncomp, how):
    mesh = fpp.unitCube(m, 0.15, True)
    part = {"rcb": lambda: par.partition_rcb(mesh, world), "metis": lambda: par.partition_metis(mesh, world)}.get(
        how, lambda: par.partition_bfs(mesh, world, int(how[3:] or 0)))()
    lps = local_problems(mesh, part, world, order, ncomp)
    ghosts_interleaved = False
    for r, lp in enumerate(lps):
        l2g, nown, p2c, gho, first = orc.doftable_mpi(mesh.conn, part, world, r, order, ncomp)
        assert lp.nDofOwned == nown and lp.nDofWithGhost == len(p2c) and lp.firstDof == first[r] and lp.nGlobalDof == first[-1]
        assert np.array_equal(lp.l2g, l2g), "rank %d: local-to-global table" % r
        assert np.array_equal(lp.p2c, p2c), "rank %d: process -> cluster map" % r
        assert np.array_equal(lp.ghost_owner_dof, gho), "rank %d: ghost owners" % r
        ia, ja = orc.graph_mpi(mesh.conn, part, world, r, order, ncomp, l2g, len(p2c))
        assert np.array_equal(lp.ia, ia) and np.array_equal(lp.ja, ja), "rank %d: local CSR pattern" % r
        g = gho[::ncomp]
        ghosts_interleaved |= len(g) > 2 and (np.diff(g) < 0).any()
    if world >= 3:
        assert ghosts_interleaved, "with three and more ranks the ghosts of one owner should not be contiguous"
