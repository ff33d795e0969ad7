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
    """form2 on a LocalSpace; close() ships the ghost rows to their owners (MatAssemblyBegin/End analogue)"""

    def __init__(self, Xh, engine, dist, algorithm=ALGO_AUTO):
        import torch
        self.torch, self.dist = torch, dist
        super().__init__(Xh, Xh, engine, 0, algorithm)
        lp = Xh.lp
        dev = self.t_values.device
        self.recv = [(p, torch.from_numpy(slots).to(dev), torch.empty(len(slots), dtype=torch.float64, device=dev))
                     for p, slots in lp.recv]
        self.send = list(lp.send)

    def _alloc_values(self):
        self.t_values = self.torch.zeros(max(self.nnz, 1), dtype=self.torch.float64, device="cuda")
        return C.c_void_p(self.t_values.data_ptr())

    def close(self, stream=None):
        """stream-ordered on torch's current stream: sends read the ghost-row tail, owners add what they receive"""
        torch, dist = self.torch, self.dist
        ops = [dist.P2POp(dist.isend, self.t_values[off:off + cnt], q) for q, off, cnt in self.send]
        ops += [dist.P2POp(dist.irecv, buf, p) for p, _, buf in self.recv]
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        sptr = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        for p, slots, buf in self.recv:
            check(self.eng.lib.feelgpu_scatter_add(self.eng.ctx, len(buf), C.c_void_p(slots.data_ptr()),
                                                   C.c_void_p(buf.data_ptr()), self.d_values, sptr), self.eng.ctx)
        return self

    def owned_values(self):
        nO = self.Xh.lp.nDofOwned
        return self.t_values[:int(self.ia[nO])].cpu().numpy()
