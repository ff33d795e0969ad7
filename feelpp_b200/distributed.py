"""Multi-GPU forms: one process per GPU, partition-local assembly, close() through the engine's C ABI.

Everything that happens per assembly -- where the ghost rows go, the NCCL exchange, the owner-side add -- lives in
libfeelgpu.so (feelpp_b200/csrc/exchange.cu: feelgpu_comm_init, feelgpu_exchange_create, feelgpu_form2_close,
feelgpu_form1_close); this module only flattens the host-side DofTable / DataMap mirror (parallel.LocalProblem) into those
calls, as a Feel++ adapter would.  torch.distributed is used for what it is here for: shipping the 128-byte NCCL id at
start-up (Feel++ would use its WorldComm) and the harness' own bookkeeping collectives.
"""
import ctypes as C

import numpy as np

from . import hostmirror as hm
from ._capi import c_i32p, c_i64p, check, ptr
from .engine import ALGO_AUTO, BilinearForm, LinearForm


def comm_init(engine, dist):
    """feelgpu_comm_init on this rank's context (idempotent): rank 0 creates the NCCL id, everybody receives it"""
    if getattr(engine, "_comm", False):
        return
    rank, world = dist.get_rank(), dist.get_world_size()
    box = [None]
    if rank == 0:
        uid = C.create_string_buffer(128)
        check(engine.lib.feelgpu_comm_unique_id(uid), None)
        box[0] = uid.raw
    dist.broadcast_object_list(box, src=0)
    check(engine.lib.feelgpu_comm_init(engine.ctx, box[0], rank, world), engine.ctx)
    engine._comm = True


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
    """form2 on a LocalSpace; close() hands the ghost rows to their owners (MatrixPetsc::close -> MatAssemblyBegin/End,
    feelalg/matrixpetsc.cpp:726-760).  peer_memory=False keeps the ghost rows in the local tail and ships them with NCCL
    send/recv (what forms outside the entity plan do anyway)."""

    def __init__(self, Xh, engine, dist, algorithm=ALGO_AUTO, peer_memory=None, pattern_flags=0):
        import os
        import torch
        self.torch, self.dist = torch, dist
        comm_init(engine, dist)
        super().__init__(Xh, Xh, engine, pattern_flags, algorithm)
        lp = Xh.lp
        if peer_memory is None:
            peer_memory = os.environ.get("FEELGPU_PEER_MEMORY", "1") != "0"
        self.ex = C.c_void_p()
        go = np.ascontiguousarray(lp.ghost_owner_dof, dtype=np.int32)
        check(engine.lib.feelgpu_exchange_create(self.pat, ptr(go, c_i32p) if len(go) else None, lp.firstDof,
                                                 0 if peer_memory else 1, C.byref(self.ex)), engine.ctx)
        self.peer = bool(self.exchange_info()["peer_memory"])

    def exchange_info(self):
        st = np.zeros(8, dtype=np.int64)
        check(self.eng.lib.feelgpu_exchange_info(self.ex, ptr(st, c_i64p)), self.eng.ctx)
        keys = ["owners", "ghost_rows", "send_values", "holders", "recv_rows", "recv_values", "peer_memory", "epoch"]
        return dict(zip(keys, (int(x) for x in st)))

    def _alloc_values(self):
        self.t_values = self.torch.zeros(max(self.nnz, 1), dtype=self.torch.float64, device="cuda")
        return C.c_void_p(self.t_values.data_ptr())

    def assemble_desc(self, desc, accumulate, stream=None):
        """feelgpu_form2_assemble (asynchronous on `stream`); the library notes where the ghost rows went"""
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
        sptr = C.c_void_p(self.torch.cuda.current_stream().cuda_stream)
        check(self.eng.lib.feelgpu_form2_close(self.pat, self.d_values, sptr), self.eng.ctx)
        return self

    def owned_values(self):
        nO = self.Xh.lp.nDofOwned
        return self.t_values[:int(self.ia[nO])].cpu().numpy()


class DistributedLinearForm(LinearForm):
    """form1 on a LocalSpace; close() = VectorPetscMPI::close (VecAssemblyBegin/End + VecGhostUpdate,
    feelalg/vectorpetsc.cpp:1667-1693) through the exchange of a bilinear form on the same space"""

    def __init__(self, Xh, engine, form2):
        super().__init__(Xh, engine)
        self.ex = form2.ex

    def close(self, ghost_update=True):
        check(self.eng.lib.feelgpu_form1_close(self.ex, self.d_values, 1 if ghost_update else 0, None), self.eng.ctx)
        self.eng.sync()
        return self

    def owned_values(self):
        return self.values()[:self.Xh.lp.nDofOwned]
