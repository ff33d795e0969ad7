"""Domain decomposition of the assembly: one process per GPU, element partition + owned row block + ghost rows.

Mirrors what the reference does in MPI mode:
  * every rank assembles its active elements only (feelmesh/filters.hpp:201);
  * a dof is owned by the lowest rank that shares it (feeldiscr/doftablempi.hpp:138-200); local numbering = owned dofs
    first, ghosts after (feeldiscr/doftable.hpp:1975-2030); global-cluster numbering = rank-contiguous blocks
    (feelalg/datamap.hpp:305-347);
  * the graph of the owned rows is complete (ghost rows are shipped to their owner in GraphCSR::close,
    feelalg/graphcsr.cpp:800-951); the values assembled into ghost rows travel once per assembly
    (MatAssemblyBegin/End, feelalg/matrixpetsc.cpp:750-752) -- here: values only, NCCL send/recv, then a scatter-add
    through slot lists precomputed from the graph exchange.

The partition itself is an input (Feel++ gets it from METIS, feelpartition/partitionermetis_impl.hpp:42-316); the
harness uses recursive coordinate bisection of the element centroids.  Setup communication goes through
torch.distributed object collectives (any backend); the per-assembly exchange is P2P on tensors (NCCL on GPUs).
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from . import hostmirror as hm
from ._capi import c_i32p, c_i64p, hcheck, ptr

E2P = np.array([[1, 2], [2, 0], [0, 1], [0, 3], [1, 3], [2, 3]])


def partition_rcb(mesh, nparts):
    """recursive coordinate bisection of the element centroids into nparts (power of two or not) balanced parts"""
    cent = [sum(mesh.xyz[mesh.conn[:, v], c].astype(np.float32) for v in range(4)) for c in range(3)]
    part = np.zeros(mesh.numElements(), dtype=np.int32)

    def split(idx, p0, np_):
        if np_ == 1:
            part[idx] = p0
            return
        left = np_ // 2
        ext = [float(cent[c][idx].max() - cent[c][idx].min()) for c in range(3)]
        axis = int(np.argmax(ext))
        k = int(len(idx) * left / np_)
        order = np.argpartition(cent[axis][idx], k)
        split(idx[order[:k]], p0, left)
        split(idx[order[k:]], p0 + left, np_ - left)

    split(np.arange(mesh.numElements()), 0, nparts)
    return part


def node_keys(conn, order, npts):
    """global identity of every local node of every element: vertices by point id, edge nodes by their end points"""
    if order > 2:
        raise NotImplementedError("the distributed harness covers Lagrange orders 1 and 2")
    conn = conn.astype(np.int64)
    keys = [conn]
    if order == 2:
        a, b = conn[:, E2P[:, 0]], conn[:, E2P[:, 1]]
        keys.append(npts + np.minimum(a, b) * npts + np.maximum(a, b))
    return np.concatenate(keys, axis=1)


class Comm:
    """thin wrapper over torch.distributed (or a single process)"""

    def __init__(self, rank=0, size=1, dist=None):
        self.rank, self.size, self.dist = rank, size, dist

    def allgather(self, obj):
        if self.size == 1:
            return [obj]
        out = [None] * self.size
        self.dist.all_gather_object(out, obj)
        return out


class LocalProblem:
    """everything rank `comm.rank` needs: local mesh, renumbered DofTable, local CSR rows, exchange plan"""

    def __init__(self, mesh, part, order, ncomp, comm):
        h = capi.host()
        rank, P = comm.rank, comm.size
        self.comm, self.order, self.ncomp = comm, order, ncomp
        n = hm.nLocalDof(order)
        npts = mesh.numPoints()
        conn = mesh.conn
        part = np.asarray(part, dtype=np.int32)
        mine = np.flatnonzero(part == rank)
        # ---- extended element set: my elements + every element sharing a vertex with them ----
        vmask = np.zeros(npts, dtype=bool)
        vmask[conn[mine]] = True
        ext = np.flatnonzero(vmask[conn].any(axis=1))
        pts_ext = np.unique(conn[ext])                       # ascending: edge orientation (by point id) is preserved
        conn_ext = np.searchsorted(pts_ext, conn[ext]).astype(np.int32)
        mesh_ext = hm.Mesh(mesh.xyz[pts_ext], conn_ext)
        X = hm.FunctionSpace(mesh_ext, order, 1)            # node-level table on the extended set
        L = X.l2g.astype(np.int64)                           # [nExt][n] ext node ids
        nnode = X.nDof()
        keys = np.zeros(nnode, dtype=np.int64)
        keys[L] = node_keys(conn[ext], order, npts)
        part_ext = part[ext]
        is_mine = part_ext == rank
        owner = np.full(nnode, P, dtype=np.int32)
        for q in range(P - 1, -1, -1):                       # descending: the lowest rank sharing a node wins
            owner[L[part_ext == q]] = q
        local = np.zeros(nnode, dtype=bool)
        local[L[is_mine]] = True
        owned = local & (owner == rank)
        ghost = local & (owner != rank)
        owned_ids = np.flatnonzero(owned)
        ghost_ids = np.flatnonzero(ghost)
        ghost_ids = ghost_ids[np.lexsort((keys[ghost_ids], owner[ghost_ids]))]   # grouped by owner: contiguous sends
        self.nOwnedNodes, self.nGhostNodes = len(owned_ids), len(ghost_ids)
        counts = comm.allgather(self.nOwnedNodes)
        first = np.concatenate([[0], np.cumsum(counts)])
        self.firstNode, self.nGlobalNodes = int(first[rank]), int(first[-1])
        # ---- cluster ids: mine are known, the others come from their owners ----
        cluster = np.full(nnode, -1, dtype=np.int64)
        cluster[owned_ids] = first[rank] + np.arange(len(owned_ids))
        # owners publish the nodes of their interface elements (elements with a vertex shared with another rank)
        vbits = np.zeros(npts, dtype=np.uint64)
        for q in range(P):
            seen = np.zeros(npts, dtype=bool)
            seen[conn[part == q]] = True
            vbits[seen] |= np.uint64(1) << np.uint64(q % 64)
        shared_v = (vbits & (vbits - np.uint64(1))) != 0      # touched by elements of more than one rank
        iface = is_mine & shared_v[conn[ext]].any(axis=1)
        pub = np.unique(L[iface])
        pub = pub[owned[pub]]
        got = comm.allgather((keys[pub], cluster[pub]))
        allk = np.concatenate([g[0] for g in got])
        allc = np.concatenate([g[1] for g in got])
        srt = np.argsort(allk)
        allk, allc = allk[srt], allc[srt]
        need = np.flatnonzero(cluster < 0)
        # only nodes that can appear in my rows matter; the others keep -1 and must never be referenced
        pos = np.searchsorted(allk, keys[need])
        pos[pos >= len(allk)] = 0
        hit = allk[pos] == keys[need] if len(allk) else np.zeros(len(need), dtype=bool)
        cluster[need[hit]] = allc[pos[hit]]
        self.cluster_of_ext = cluster
        # ---- local numbering: owned first, ghosts after ----
        loc_nodes = np.concatenate([owned_ids, ghost_ids])
        newidx = np.full(nnode, -1, dtype=np.int64)
        newidx[loc_nodes] = np.arange(len(loc_nodes))
        nloc = len(loc_nodes)
        if (cluster[loc_nodes] < 0).any():
            raise RuntimeError("distributed setup: a local node has no global-cluster id")
        Lm = newidx[L[is_mine]]                              # [nMine][n] local node ids
        self.l2g = (Lm[:, None, :] * ncomp + np.arange(ncomp)[None, :, None]).reshape(len(mine), ncomp * n).astype(np.int32)
        self.nDofOwned, self.nDofWithGhost = len(owned_ids) * ncomp, nloc * ncomp
        self.p2c = (cluster[loc_nodes][:, None] * ncomp + np.arange(ncomp)[None, :]).reshape(-1).astype(np.int64)
        self.firstDof, self.nGlobalDof = self.firstNode * ncomp, self.nGlobalNodes * ncomp
        self.elements = mine
        self.mesh = hm.Mesh(mesh.xyz[pts_ext], conn_ext[is_mine])
        self.ghost_owner = owner[ghost_ids]
        # ---- local CSR rows: complete pattern for owned rows, local-elements pattern for ghost rows ----
        ia, ja = self._local_graph(h, L, is_mine, nnode, loc_nodes, cluster, n)
        self.ia, self.ja = ia, ja
        self._exchange_plan()

    def _graph(self, h, l2g_nodes, nnode):
        """dof-level CSR (ext numbering) of the elements given by their node table"""
        ncomp, n = self.ncomp, l2g_nodes.shape[1]
        l2g = (l2g_nodes[:, None, :] * ncomp + np.arange(ncomp)[None, :, None]).reshape(len(l2g_nodes), ncomp * n)
        l2g = np.ascontiguousarray(l2g, dtype=np.int32)
        hd, nnz = C.c_void_p(), C.c_int64()
        hcheck(h.feelhost_graph_build(len(l2g), self.order, ncomp, ptr(l2g, c_i32p), nnode * ncomp, 0, C.byref(hd),
                                      C.byref(nnz)))
        ia = np.empty(nnode * ncomp + 1, dtype=np.int64)
        ja = np.empty(nnz.value, dtype=np.int64)
        h.feelhost_graph_get(hd, ptr(ia, c_i64p), ptr(ja, c_i64p))
        h.feelhost_graph_free(hd)
        return ia, ja

    def _local_graph(self, h, L, is_mine, nnode, loc_nodes, cluster, n):
        ncomp = self.ncomp
        iaA, jaA = self._graph(h, L, nnode)                  # all extended elements: complete rows of owned dofs
        iaB, jaB = self._graph(h, L[is_mine], nnode)         # my elements only: what I contribute to ghost rows
        rid = (loc_nodes[:, None] * ncomp + np.arange(ncomp)[None, :]).reshape(-1).astype(np.int64)
        colmap = (cluster[:, None] * ncomp + np.arange(ncomp)[None, :]).reshape(-1).astype(np.int64)
        nnew = len(rid)
        ia = np.empty(nnew + 1, dtype=np.int64)
        args = [nnew, self.nDofOwned, ptr(rid, c_i64p), ptr(iaA, c_i64p), ptr(jaA, c_i64p), ptr(iaB, c_i64p),
                ptr(jaB, c_i64p), ptr(colmap, c_i64p), ptr(ia, c_i64p)]
        hcheck(h.feelhost_csr_select_remap_sort(*args, None))
        ja = np.empty(int(ia[-1]), dtype=np.int64)
        hcheck(h.feelhost_csr_select_remap_sort(*args, ptr(ja, c_i64p)))
        if len(ja) and ja.min() < 0:
            raise RuntimeError("distributed setup: a column of a local row has no global-cluster id")
        return ia, ja

    # ------------------------------------------------------------------------------------------------------------
    def _exchange_plan(self):
        """graph exchange (GraphCSR::close analogue): tell every owner which (row, columns) my ghost rows carry, so
        that it can precompute the CSR slots of the values it will receive at every assembly"""
        comm, ncomp = self.comm, self.ncomp
        rank = comm.rank
        nO = self.nDofOwned
        ia, ja = self.ia, self.ja
        gowner = np.repeat(self.ghost_owner, ncomp)          # owner of every ghost dof row, grouped by owner
        self.send = []                                        # (dest, value offset, count)
        msg = {}
        for q in np.unique(gowner):
            rows = nO + np.flatnonzero(gowner == q)
            r0, r1 = rows[0], rows[-1] + 1                    # contiguous by construction
            off, cnt = int(ia[r0]), int(ia[r1] - ia[r0])
            self.send.append((int(q), off, cnt))
            msg[int(q)] = (self.p2c[r0:r1].copy(), np.diff(ia[r0:r1 + 1]), ja[ia[r0]:ia[r1]].copy())
        allmsg = comm.allgather(msg)
        self.recv = []                                        # (source, slots int64[count])
        for p, m in enumerate(allmsg):
            if p == rank or rank not in m:
                continue
            rowc, lens, cols = m[rank]
            lrow = rowc - self.firstDof                       # my local owned row of every incoming ghost row
            if len(lrow) and (lrow.min() < 0 or lrow.max() >= nO):
                raise RuntimeError("distributed setup: received a ghost row I do not own")
            slots = np.empty(len(cols), dtype=np.int64)
            k = 0
            for r, ln in zip(lrow, lens):
                seg = ja[ia[r]:ia[r + 1]]
                pos = np.searchsorted(seg, cols[k:k + ln])
                if (pos >= len(seg)).any() or (seg[np.minimum(pos, len(seg) - 1)] != cols[k:k + ln]).any():
                    raise RuntimeError("distributed setup: a ghost-row entry is not in the owner's graph")
                slots[k:k + ln] = ia[r] + pos
                k += ln
            self.recv.append((p, slots))

    def nnz_local(self):
        return int(self.ia[-1])

    def nnz_owned(self):
        return int(self.ia[self.nDofOwned])


def exchange_numpy(lp, a, dist):
    """CPU/gloo form of close(): used by the world_size-2 tests; a is the rank's full local value array (numpy)"""
    import torch
    ops, bufs = [], []
    for q, off, cnt in lp.send:
        t = torch.from_numpy(a[off:off + cnt].copy())
        ops.append(dist.P2POp(dist.isend, t, q))
    for p, slots in lp.recv:
        t = torch.empty(len(slots), dtype=torch.float64)
        bufs.append((t, slots))
        ops.append(dist.P2POp(dist.irecv, t, p))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    for t, slots in bufs:
        np.add.at(a, slots, t.numpy())
    return a
