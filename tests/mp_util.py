"""Helpers of the multi-rank tests: gather a distributed matrix / vector into the SEQUENTIAL numbering and compare it with
the sequential oracle entry by entry.

The permutation cluster dof -> sequential dof goes through the geometric identity of the Lagrange nodes
(parallel.node_keys: point id / edge end points + rank along the edge / face point triple), which both the sequential
DofTable and every rank's local DofTable can state.  After the permutation the pattern must be the sequential graph bit for
bit and every value within oracle/parity.py's bound -- a lost, doubled or misplaced ghost contribution cannot hide."""
import numpy as np
import scipy.sparse as sps

import feelpp_b200 as fpp
from feelpp_b200 import parallel as par
from oracle import oracle as orc
from oracle import parity


def make_partition(mesh, world, how):
    if how == "rcb":
        return par.partition_rcb(mesh, world)
    if how == "metis":
        return par.partition_metis(mesh, world)
    if how.startswith("bfs"):
        return par.partition_bfs(mesh, world, seed=int(how[3:] or 0))
    raise ValueError(how)


def rank_payload(lp, owned_values, b=None):
    """what a rank ships to the comparing process: its owned rows in cluster numbering + the keys of its owned nodes"""
    nO = lp.nDofOwned
    rows = np.repeat(lp.p2c[:nO], np.diff(lp.ia[:nO + 1]))
    return {"rows": rows, "cols": lp.ja[:lp.ia[nO]], "vals": np.asarray(owned_values),
            "keys": lp.node_keys_local[:lp.nOwnedNodes], "first": lp.firstNode, "nown": lp.nOwnedNodes,
            "b": None if b is None else np.asarray(b)[:nO]}


def cluster_to_sequential(gathered, mesh, order, ncomp, Xh):
    """dof permutation cluster -> sequential"""
    nn = Xh.nDof() // ncomp
    first = [g["first"] for g in gathered]
    counts = [g["nown"] for g in gathered]
    assert first == list(np.concatenate([[0], np.cumsum(counts)])[:-1]), "owned blocks are not rank-contiguous"
    assert sum(counts) == nn, "owned nodes of all ranks != nodes of the sequential space"
    key_of_cluster = np.concatenate([g["keys"] for g in gathered])
    kk, seq = par.sequential_numbering_of_keys(mesh, order, Xh)
    pos = np.searchsorted(kk, key_of_cluster)
    assert (pos < len(kk)).all() and (kk[np.minimum(pos, len(kk) - 1)] == key_of_cluster).all(), "unknown node key"
    node_perm = seq[pos]
    assert len(np.unique(node_perm)) == nn, "two cluster nodes map to one sequential node"
    return (node_perm[:, None] * ncomp + np.arange(ncomp)[None, :]).ravel()


def gathered_matrix(gathered, perm, N):
    r = np.concatenate([perm[g["rows"]] for g in gathered])
    c = np.concatenate([perm[g["cols"]] for g in gathered])
    v = np.concatenate([g["vals"] for g in gathered])
    D = sps.coo_matrix((v, (r, c)), shape=(N, N)).tocsr()
    D.sort_indices()
    return D


def check_against_sequential(gathered, mesh, order, ncomp, kind, coeff=None, what="", factor=1.0):
    """pattern bit-exact + every value within the parity bound; returns worst error / tolerance"""
    Xh = fpp.FunctionSpace(mesh, order, ncomp)
    ia, ja = Xh.graph()
    N = Xh.nDof()
    perm = cluster_to_sequential(gathered, mesh, order, ncomp, Xh)
    D = gathered_matrix(gathered, perm, N)
    assert sum(len(g["vals"]) for g in gathered) == len(ja), "%s: owned nonzeros of all ranks != sequential nnz" % what
    assert np.array_equal(D.indptr, ia) and np.array_equal(D.indices, ja), "%s: pattern differs from the sequential graph" % what
    q = orc.quad_order_form2(kind, order)
    co = coeff or orc.Coeff()
    ref = factor * orc.form2(kind, order, ncomp, q, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, co)
    scale = abs(factor) * orc.form2(kind | orc.ABS_FLAG, order, ncomp, q, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, co)
    return parity.check(D.data, ref, scale, ia, ja, what)


def check_vector_against_sequential(gathered, mesh, order, ncomp, ref_of, what=""):
    Xh = fpp.FunctionSpace(mesh, order, ncomp)
    perm = cluster_to_sequential(gathered, mesh, order, ncomp, Xh)
    b = np.zeros(Xh.nDof())
    b[perm] = np.concatenate([g["b"] for g in gathered])
    ref = ref_of(Xh)
    np.testing.assert_allclose(b, ref, rtol=1e-12, atol=1e-13 * np.abs(ref).max(), err_msg=what)
