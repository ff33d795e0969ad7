"""Multi-GPU forms: one process per GPU, partition-local assembly + NCCL exchange of the ghost-row values.

torch is used for what it is here for: device memory (the value array is a torch tensor so that torch.distributed can
send slices of it), streams and the NCCL communicator.  All kernels (assembly, owner-side scatter-add) are the engine's.
"""
import ctypes as C

import numpy as np

from . import hostmirror as hm
from . import parallel as par
from ._capi import check
from .engine import ALGO_AUTO, BilinearForm


class LocalSpace:
    """the rank-local view of a distributed Pch/Pchv space: owned dofs first, ghosts after (doftable.hpp:1975-2030)"""

    def __init__(self, lp):
        self.lp = lp
        self.mesh, self.order, self.ncomp = lp.mesh, lp.order, lp.ncomp
        self.n = hm.nLocalDof(lp.order)
        self.nd = self.n * lp.ncomp
        self.l2g = lp.l2g
        self.p2c = lp.p2c
        self.nDofOwned = lp.nDofOwned

    def nDof(self):
        return self.lp.nDofWithGhost

    nLocalDofWithGhost = nDof

    def nLocalDofWithoutGhost(self):
        return self.lp.nDofOwned

    def graph(self):
        return self.lp.ia, self.lp.ja


class DistributedBilinearForm(BilinearForm):
    """form2 on a LocalSpace; close() hands the ghost rows to their owners (MatAssemblyBegin/End analogue).

    Two transports.  peer memory (default when every rank's pattern carries the entity plan): the owners export one
    landing buffer per neighbour (CUDA IPC over NVLink/NVSwitch), the assembly kernel stores the ghost rows straight into
    them while it runs, and close() is an NCCL barrier + the owner-side scatter-add; the buffers alternate between two
    sets so that one barrier per close orders everything.  Otherwise, and for forms the entity plan does not cover, the
    ghost rows are the tail of the local value array and close() ships it with NCCL send/recv."""

    def __init__(self, Xh, engine, dist, algorithm=ALGO_AUTO, peer_memory=None):
        import os
        import torch
        self.torch, self.dist = torch, dist
        super().__init__(Xh, Xh, engine, 0, algorithm)
        lp = Xh.lp
        dev = self.t_values.device
        self.recv = [(p, torch.from_numpy(slots).to(dev), torch.empty(len(slots), dtype=torch.float64, device=dev))
                     for p, slots in lp.recv]
        self.send = list(lp.send)
        self.tail0 = int(self.ia[lp.nDofOwned])
        self._tail_used = self._direct_used = False
        self.peer = False
        if peer_memory is None:
            peer_memory = os.environ.get("FEELGPU_PEER_MEMORY", "1") != "0"
        if peer_memory and dev.type == "cuda" and dist.get_world_size() > 1:
            self._setup_peer()

    def _alloc_values(self):
        self.t_values = self.torch.zeros(max(self.nnz, 1), dtype=self.torch.float64, device="cuda")
        return C.c_void_p(self.t_values.data_ptr())

    # ---- peer-memory transport ----
    def _setup_peer(self):
        torch, dist, eng = self.torch, self.dist, self.eng
        rank, world = dist.get_rank(), dist.get_world_size()
        ok = int(self.plan_stats()["eplan"]["has_plan"] == 1 and len(self.send) <= 8)
        t = torch.tensor([ok], device="cuda", dtype=torch.int32)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        if int(t.item()) == 0:
            return
        # landing buffers for the rows my neighbours hold as ghosts: two sets (epoch parity), exported to the neighbours
        self.land = []
        mine = {}
        for p, slots, _ in self.recv:
            bufs, hs = [], []
            for _e in range(2):
                d = eng.malloc(8 * max(len(slots), 1))
                h = C.create_string_buffer(64)
                check(eng.lib.feelgpu_ipc_export(eng.ctx, d, h), eng.ctx)
                bufs.append(d)
                hs.append(h.raw)
            self.land.append((p, slots, bufs))
            mine[p] = hs
        allh = [None] * world
        dist.all_gather_object(allh, mine)
        # the buffers of the owners of my ghost rows, in slot order
        self.remote = []
        for q, off, cnt in sorted(self.send, key=lambda x: x[1]):
            ptrs = []
            for e in range(2):
                p = C.c_void_p()
                check(eng.lib.feelgpu_ipc_open(eng.ctx, allh[q][rank][e], C.byref(p)), eng.ctx)
                ptrs.append(p)
            self.remote.append((off, off + cnt, ptrs))
        self.flag = torch.zeros(1, device="cuda", dtype=torch.int32)
        self.epoch = 0
        self.peer = True
        self._set_remote()

    def _set_remote(self):
        n = len(self.remote)
        lo = np.array([r[0] for r in self.remote], dtype=np.int64)
        hi = np.array([r[1] for r in self.remote], dtype=np.int64)
        tg = (C.c_void_p * max(n, 1))(*[r[2][self.epoch & 1] for r in self.remote])
        check(self.eng.lib.feelgpu_pattern_set_remote(self.pat, n, lo.ctypes.data_as(C.POINTER(C.c_int64)),
                                                      hi.ctypes.data_as(C.POINTER(C.c_int64)), tg), self.eng.ctx)

    def assemble_desc(self, desc, accumulate, stream=None):
        """feelgpu_form2_assemble with the bookkeeping of where the ghost rows went (asynchronous on `stream`)"""
        direct = C.c_int(0)
        if self.peer:
            check(self.eng.lib.feelgpu_form2_uses_remote(self.pat, C.byref(desc), C.byref(direct)), self.eng.ctx)
        if direct.value:
            if not accumulate:
                self._tail_used = False          # '=' voids what the local tail held
            self._direct_used = True
        else:
            if self.peer and not accumulate:
                self._direct_used = False        # '=' voids the landing buffers: they count as empty again
                self._set_remote()
            if accumulate and not self._tail_used and self.nnz > self.tail0:
                self.t_values[self.tail0:].zero_()
            self._tail_used = True
        check(self.eng.lib.feelgpu_form2_assemble(self.pat, C.byref(desc), accumulate, self.d_values, stream), self.eng.ctx)

    def _assemble(self, integ, accumulate):
        from .engine import _Keep
        keep = _Keep()
        d = self._desc(integ, keep)
        sptr = C.c_void_p(self.torch.cuda.current_stream().cuda_stream)
        self.assemble_desc(d, accumulate, sptr)
        self.torch.cuda.current_stream().synchronize()
        for r in keep.refs:
            if isinstance(r, tuple):
                self.eng.free(r[1])
        return self

    def close(self, stream=None):
        """stream-ordered on torch's current stream: owners add the ghost-row contributions of their neighbours"""
        torch, dist = self.torch, self.dist
        sptr = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        if self.peer and self._direct_used:
            dist.all_reduce(self.flag)     # every rank's kernel is complete: its stores into my landing buffers are visible
            for p, slots, bufs in self.land:
                check(self.eng.lib.feelgpu_scatter_add(self.eng.ctx, len(slots), C.c_void_p(slots.data_ptr()),
                                                       bufs[self.epoch & 1], self.d_values, sptr), self.eng.ctx)
            self.epoch += 1
            self._set_remote()
            self._direct_used = False
        if self._tail_used or not self.peer:
            ops = [dist.P2POp(dist.isend, self.t_values[off:off + cnt], q) for q, off, cnt in self.send]
            ops += [dist.P2POp(dist.irecv, buf, p) for p, _, buf in self.recv]
            if ops:
                for w in dist.batch_isend_irecv(ops):
                    w.wait()
            for p, slots, buf in self.recv:
                check(self.eng.lib.feelgpu_scatter_add(self.eng.ctx, len(buf), C.c_void_p(slots.data_ptr()),
                                                       C.c_void_p(buf.data_ptr()), self.d_values, sptr), self.eng.ctx)
            self._tail_used = False
        return self

    def owned_values(self):
        nO = self.Xh.lp.nDofOwned
        return self.t_values[:int(self.ia[nO])].cpu().numpy()
