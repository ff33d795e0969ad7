"""Domain decomposition of the assembly: one process per GPU, element partition + owned row block + ghost rows.

Host-side mirror of what the reference's DofTable / DataMap / GraphCSR hold in MPI mode -- the
inputs a Feel++ adapter flattens into the C ABI:
  * every rank assembles its active elements only (feelmesh/filters.hpp:201);
  * the local DofTable is built by first touch over the ACTIVE elements in local element order (feeldiscr/doftable.hpp
    :1802-1940); a dof is a ghost when a lower rank shares its entity (feeldiscr/doftablempi.hpp:138-200, :548); the
    local numbering is then reordered "active dofs before, ghost dofs after", each group keeping its first-touch order
    (feeldiscr/doftable.hpp:1975-1992) -- so the ghosts of one owner are NOT contiguous;
  * global-cluster numbering = rank-contiguous blocks, active dofs numbered in their local order
    (feeldiscr/doftablempi.hpp:593-636, feelalg/datamap.hpp:305-347);
  * the graph of the owned rows is complete (ghost rows are shipped to their owner in GraphCSR::close,
    feelalg/graphcsr.cpp:800-951); the values assembled into ghost rows travel once per assembly
    (MatAssemblyBegin/End, feelalg/matrixpetsc.cpp:750-752).

The partition itself is an input: METIS on the face-adjacency graph as Feel++ does it
(feelpartition/partitionermetis_impl.hpp:118-253; vendored sources compiled into libfeelhost.so), or, for tests and
comparison, coordinate bisection / seeds grown breadth-first.  Setup communication of this harness goes through
torch.distributed object collectives (any backend); the engine's own exchange (feelgpu_exchange_create, NCCL) takes the
ghost owners and cluster ids computed here, exactly what DataMap gives an adapter.
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from . import hostmirror as hm
from ._capi import c_i32p, c_i64p, hcheck, ptr

E2P = np.array([[1, 2], [2, 0], [0, 1], [0, 3], [1, 3], [2, 3]])      # feelmesh/simplexordering.hpp:283-291
F2P = np.array([[1, 2, 3], [0, 2, 3], [0, 1, 3], [0, 1, 2]])          # :341-347


# ---- partitioners ----------------------------------------------------------------------------------------------------
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


def face_adjacency(mesh):
    """dual graph of the mesh by faces (the graph Feel++ hands to METIS, partitionermetis_impl.hpp:118-186):
    (xadj[nElt+1], adjncy) int32"""
    h = capi.host()
    ne = mesh.numElements()
    xadj = np.empty(ne + 1, dtype=np.int32)
    adj = np.empty(4 * ne, dtype=np.int32)
    hcheck(h.feelhost_dual_graph(mesh.numPoints(), ne, ptr(mesh.conn, c_i32p), ptr(xadj, c_i32p), ptr(adj, c_i32p)))
    return xadj, adj[:xadj[-1]]


def partition_metis(mesh, nparts, seed=None):
    """Feel++'s partitioner (feelpartition/partitionermetis_impl.hpp:118-253): METIS 5.1.0 on the face-adjacency graph,
    vertex weight = points per element (:133), recursive bisection for nparts <= 8, k-way above (:227-239)"""
    h = capi.host()
    ne = mesh.numElements()
    part = np.zeros(ne, dtype=np.int32)
    if nparts > 1:
        hcheck(h.feelhost_partition_metis(mesh.numPoints(), ne, ptr(mesh.conn, c_i32p), int(nparts),
                                          -1 if seed is None else int(seed), ptr(part, c_i32p)))
    return part


def partition_bfs(mesh, nparts, seed=0):
    """irregular partition for tests: nparts random seed elements grown breadth-first over the face-adjacency graph,
    unequal sizes, ranks with three and more neighbours"""
    xadj, adj = face_adjacency(mesh)
    ne = mesh.numElements()
    rng = np.random.default_rng(seed)
    part = np.full(ne, -1, dtype=np.int32)
    fronts = [[int(s)] for s in rng.choice(ne, nparts, replace=False)]
    for q, f in enumerate(fronts):
        part[f[0]] = q
    speed = 1 + rng.integers(0, 3, nparts)          # unequal growth -> unequal sizes
    left = ne - nparts
    while left > 0:
        moved = False
        for q in range(nparts):
            for _ in range(int(speed[q])):
                nxt = []
                for e in fronts[q]:
                    for nb in adj[xadj[e]:xadj[e + 1]]:
                        if part[nb] < 0:
                            part[nb] = q
                            nxt.append(int(nb))
                            left -= 1
                if nxt:
                    moved = True
                    fronts[q] = nxt
        if not moved:                               # disconnected leftovers
            part[part < 0] = 0
            break
    return part


def partition(mesh, nparts, method="auto"):
    """(part, description): METIS when libfeelhost.so carries it (the reference's choice), else coordinate bisection"""
    if method in ("auto", "metis"):
        if hasattr(capi.host(), "feelhost_partition_metis"):
            return partition_metis(mesh, nparts), ("METIS %s element partition of the face-adjacency graph "
                                                   "(partitionermetis_impl.hpp:118-253)"
                                                   % ("recursive-bisection" if nparts <= 8 else "k-way"))
        if method == "metis":
            raise RuntimeError("libfeelhost.so was built without METIS")
    return partition_rcb(mesh, nparts), "recursive coordinate bisection element partition"


# ---- geometric identity of the nodes ---------------------------------------------------------------------------------
def node_keys(conn, order, npts):
    """global identity of every local node of every element, [nElt][n] int64: vertices by point id, edge nodes by their end
    points and their rank along the edge counted from the lower point id, face nodes by the sorted point triple.  Equal
    keys <=> the same Lagrange node, whatever the rank, the element or the local numbering."""
    conn = np.asarray(conn).astype(np.int64)
    keys = [conn]
    npe = order - 1
    if order >= 2:
        a, b = conn[:, E2P[:, 0]], conn[:, E2P[:, 1]]
        base = npts + (np.minimum(a, b) * npts + np.maximum(a, b)) * npe
        if npe == 1:
            keys.append(base)
        else:
            # local dof 4 + npe*e + s is the s-th node from the edge's first local vertex (simplexordering.hpp:306-314)
            fwd = (a < b)[:, :, None]
            s = np.arange(npe)[None, None, :]
            keys.append((base[:, :, None] + np.where(fwd, s, npe - 1 - s)).reshape(conn.shape[0], 6 * npe))
    if order >= 3:
        if order > 3:
            raise NotImplementedError("node keys: Lagrange orders 1 to 3")
        if npts >= (1 << 20):
            raise NotImplementedError("node keys of face dofs: more than 2^20 points")
        f = np.sort(conn[:, F2P], axis=2)
        keys.append(npts + 2 * npts * npts * npe + (f[:, :, 0] * npts + f[:, :, 1]) * npts + f[:, :, 2])
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
        # ---- extended element set: my elements + every element sharing a vertex with them (the ghost layer) ----
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
        # ---- the local DofTable: first touch over MY elements in local element order (doftable.hpp:1802-1940) ----
        self.elements = mine
        self.mesh = hm.Mesh(mesh.xyz[pts_ext], conn_ext[is_mine])
        self.point_ids = pts_ext                             # global point id of every local point
        Xl = hm.FunctionSpace(self.mesh, order, 1)
        Ll = Xl.l2g.astype(np.int64)                         # [nMine][n] first-touch ids
        nloc = Xl.nDof()
        ext_of_ft = np.empty(nloc, dtype=np.int64)
        ext_of_ft[Ll] = L[is_mine]
        ghost_ft = owner[ext_of_ft] != rank                  # a lower rank shares the entity (doftablempi.hpp:548)
        # active dofs before, ghost dofs after, each group in first-touch order (doftable.hpp:1975-1992)
        order_ft = np.concatenate([np.flatnonzero(~ghost_ft), np.flatnonzero(ghost_ft)])
        loc_nodes = ext_of_ft[order_ft]                      # ext node id of every local node, new local order
        self.nOwnedNodes = int((~ghost_ft).sum())
        self.nGhostNodes = nloc - self.nOwnedNodes
        owned_ids, ghost_ids = loc_nodes[:self.nOwnedNodes], loc_nodes[self.nOwnedNodes:]
        owned = np.zeros(nnode, dtype=bool)
        owned[owned_ids] = True
        counts = comm.allgather(self.nOwnedNodes)
        first = np.concatenate([[0], np.cumsum(counts)])
        self.firstNode, self.nGlobalNodes = int(first[rank]), int(first[-1])
        self.first_nodes = first.astype(np.int64)            # firstDofGlobalClusterWorld / ncomp
        # ---- cluster ids: the active dofs in their local order (doftablempi.hpp:623-635), ghosts from their owners ----
        cluster = np.full(nnode, -1, dtype=np.int64)
        cluster[owned_ids] = first[rank] + np.arange(len(owned_ids))
        # owners publish the nodes of their interface elements (elements with a vertex shared with another rank)
        touched = np.zeros(npts, dtype=bool)
        shared_v = np.zeros(npts, dtype=bool)
        for q in range(P):
            seen = np.zeros(npts, dtype=bool)
            seen[conn[part == q]] = True
            shared_v |= seen & touched                       # touched by elements of more than one rank
            touched |= seen
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
        self.cluster_of_ext, self.key_of_ext = cluster, keys
        self.ext_elements, self.ext_l2g_nodes, self.ext_mesh = ext, L, mesh_ext
        # ---- local numbering ----
        newidx = np.full(nnode, -1, dtype=np.int64)
        newidx[loc_nodes] = np.arange(nloc)
        if (cluster[loc_nodes] < 0).any():
            raise RuntimeError("distributed setup: a local node has no global-cluster id")
        Lm = newidx[L[is_mine]]                              # [nMine][n] local node ids
        self.l2g = (Lm[:, None, :] * ncomp + np.arange(ncomp)[None, :, None]).reshape(len(mine), ncomp * n).astype(np.int32)
        self.nDofOwned, self.nDofWithGhost = len(owned_ids) * ncomp, nloc * ncomp
        self.p2c = (cluster[loc_nodes][:, None] * ncomp + np.arange(ncomp)[None, :]).reshape(-1).astype(np.int64)
        self.node_keys_local = keys[loc_nodes]               # geometric identity of every local node (tests, parity)
        self.firstDof, self.nGlobalDof = self.firstNode * ncomp, self.nGlobalNodes * ncomp
        self.ghost_owner = owner[ghost_ids].astype(np.int32)  # per ghost node, local (first-touch) order
        self.ghost_owner_dof = np.repeat(self.ghost_owner, ncomp).astype(np.int32)
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
        # my elements only: what I contribute to ghost rows.  Only elements that hold a ghost node have a say in those rows,
        # so the second graph is built over them alone (a rank without ghost nodes -- rank 0 always -- builds none)
        ghost_node = np.zeros(nnode, dtype=bool)
        ghost_node[loc_nodes[self.nOwnedNodes:]] = True
        selB = is_mine & ghost_node[L].any(axis=1)
        if selB.any():
            iaB, jaB = self._graph(h, L[selB], nnode)
        else:
            iaB, jaB = np.zeros(nnode * ncomp + 1, dtype=np.int64), np.zeros(1, dtype=np.int64)
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
        """graph exchange (GraphCSR::close analogue, feelalg/graphcsr.cpp:800-951): tell every owner which (row, columns)
        my ghost rows carry, so that it can precompute the CSR slots of the values it will receive at every assembly.
        Host-side restatement used by the CPU (gloo) tests and the parity check; on GPUs the engine does the same
        exchange itself over NCCL (feelgpu_exchange_create)."""
        comm = self.comm
        rank = comm.rank
        nO = self.nDofOwned
        ia, ja = self.ia, self.ja
        gowner = self.ghost_owner_dof                         # owner of every ghost dof row, local order (interleaved)
        self.send = []                                        # (dest, ghost rows (local ids), value indices)
        msg = {}
        for q in np.unique(gowner):
            rows = nO + np.flatnonzero(gowner == q)
            lens = ia[rows + 1] - ia[rows]
            idx = np.repeat(ia[rows] - np.concatenate([[0], np.cumsum(lens)[:-1]]), lens) + np.arange(int(lens.sum()))
            self.send.append((int(q), rows, idx))
            msg[int(q)] = (self.p2c[rows].copy(), lens, ja[idx].copy())
        allmsg = comm.allgather(msg)
        self.recv = []                                        # (source, owned rows (local ids), slots int64[count])
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
            self.recv.append((p, lrow.astype(np.int64), slots))

    def nnz_local(self):
        return int(self.ia[-1])

    def nnz_owned(self):
        return int(self.ia[self.nDofOwned])

    def send_values(self):
        return int(self.ia[-1] - self.ia[self.nDofOwned])

    def num_neighbours(self):
        return len({q for q, _, _ in self.send} | {p for p, _, _ in self.recv})


def exchange_numpy(lp, a, dist):
    """CPU/gloo form of MatAssemblyBegin/End: a is the rank's full local value array (numpy)"""
    import torch
    ops, bufs = [], []
    for q, _rows, idx in lp.send:
        t = torch.from_numpy(a[idx].copy())
        ops.append(dist.P2POp(dist.isend, t, q))
    for p, _rows, slots in lp.recv:
        t = torch.empty(len(slots), dtype=torch.float64)
        bufs.append((t, slots))
        ops.append(dist.P2POp(dist.irecv, t, p))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    for t, slots in bufs:
        np.add.at(a, slots, t.numpy())
    return a


def exchange_vector_numpy(lp, b, dist, ghost_update=True):
    """CPU/gloo form of VectorPetscMPI::close (feelalg/vectorpetsc.cpp:1683-1693): VecAssemblyBegin/End adds the ghost
    entries to their owners, then VecGhostUpdate (:1667-1676) refreshes the ghost copies from the owners; b is the rank's
    local vector [nDofWithGhost]"""
    import torch
    ops, bufs = [], []
    for q, rows, _ in lp.send:
        ops.append(dist.P2POp(dist.isend, torch.from_numpy(b[rows].copy()), q))
    for p, lrow, _ in lp.recv:
        t = torch.empty(len(lrow), dtype=torch.float64)
        bufs.append((t, lrow))
        ops.append(dist.P2POp(dist.irecv, t, p))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    for t, lrow in bufs:
        np.add.at(b, lrow, t.numpy())
    if ghost_update:
        ops, bufs = [], []
        for p, lrow, _ in lp.recv:
            ops.append(dist.P2POp(dist.isend, torch.from_numpy(b[lrow].copy()), p))
        for q, rows, _ in lp.send:
            t = torch.empty(len(rows), dtype=torch.float64)
            bufs.append((t, rows))
            ops.append(dist.P2POp(dist.irecv, t, q))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        for t, rows in bufs:
            b[rows] = t.numpy()
    return b


# ---- gathering a distributed matrix in the sequential numbering (tests, bench parity) ----------------------------------
def sequential_numbering_of_keys(mesh, order, Xh_seq):
    """sorted (keys, sequential node id) of the sequential space on the global mesh: the permutation cluster -> sequential
    goes through the geometric identity of the nodes"""
    k = node_keys(mesh.conn, order, mesh.numPoints())
    nodes = Xh_seq.l2g[:, :Xh_seq.n].astype(np.int64) // Xh_seq.ncomp
    seq = np.empty(Xh_seq.nDof() // Xh_seq.ncomp, dtype=np.int64)
    kk = np.empty_like(seq)
    kk[nodes] = k
    seq[:] = np.arange(len(seq))
    srt = np.argsort(kk)
    return kk[srt], seq[srt]
