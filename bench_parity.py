"""bench.py's checker (harness code next to bench.py, never imported by feelpp_b200/): parity block of the multi-GPU line: is the distributed matrix after close() the reference's matrix?

Two checks per rank, on the values the timed step produced (this module is harness code: it may call the oracle, which
the product path never does):
  (1) an invariant over ALL owned rows: constants are in the kernel of the stiffness operator and rigid translations in the
      kernel of the elasticity operator (testsuite/feelvf/test_laplaciant.cpp:108-126), i.e. for every owned row and every
      column component the entries sum to zero, |sum_j a_ij| <= 1e-12 sum_j |a_ij|.  A ghost contribution that was lost,
      doubled or landed in the wrong row breaks it.
  (2) entry by entry against the oracle on a sample of owned INTERFACE rows (rows that received ghost contributions,
      the only ones close() touches): the oracle assembles the elements of the global mesh around those rows; rows and
      columns are matched through the geometric identity of the nodes (parallel.node_keys), values within
      oracle/parity.py's bound.
"""
import numpy as np

from feelpp_b200 import hostmirror as hm
from feelpp_b200 import parallel as par  # noqa: F401

SAMPLE_ROWS = 12000


def check(args, lp, gmesh, a, dist, oracle_kind_coeff, quad, sample_rows=SAMPLE_ROWS):
    import torch
    from oracle import oracle as orc
    from oracle import parity
    kind, co = oracle_kind_coeff
    ncomp, order, n = lp.ncomp, lp.order, hm.nLocalDof(lp.order)
    nO = lp.nDofOwned
    nnzO = lp.nnz_owned()
    ia = lp.ia
    dev = a.t_values.device
    # ---- (1) row sums per column component over every owned row, on the device ----
    vals = a.t_values[:nnzO]
    rows = torch.repeat_interleave(torch.arange(nO, device=dev), torch.from_numpy(np.diff(ia[:nO + 1])).to(dev))
    cols = torch.from_numpy(lp.ja[:nnzO]).to(dev)
    key = rows * ncomp + (cols % ncomp)
    s = torch.zeros(nO * ncomp, dtype=torch.float64, device=dev).index_add_(0, key, vals)
    sabs = torch.zeros(nO * ncomp, dtype=torch.float64, device=dev).index_add_(0, key, vals.abs())
    rabs = sabs.view(nO, ncomp).sum(1, keepdim=True).expand(nO, ncomp).reshape(-1)
    ratio = (s.abs() / rabs.clamp_min(1e-300))
    worst_inv = float(ratio.max().item()) if nO else 0.0
    inv_ok = worst_inv <= 1e-12
    n_inv_bad = int((ratio > 1e-12).sum().item()) if nO else 0
    n_zero_rows = int((rabs == 0).sum().item()) if nO else 0        # an assembled row always has a positive diagonal
    if n_inv_bad:
        # where: row, its sum and absolute sum, its largest entry -- on stderr, for the record
        import sys
        badk = torch.nonzero(ratio > 1e-12).flatten()[:8].cpu().numpy()
        recv_rows_all = np.unique(np.concatenate([r for _, r, _ in lp.recv])) if lp.recv else np.zeros(0, dtype=np.int64)
        for kk in badk:
            r = int(kk) // ncomp
            lo, hi = int(ia[r]), int(ia[r + 1])
            rv = a.t_values[lo:hi].cpu().numpy()
            print("[parity] rank %d: row %d (of %d owned) comp %d: sum %.3e abs-sum %.3e, %d entries, max |a| %.3e at col %d (diag col %d), "
                  "receives ghost rows: %s" % (lp.comm.rank, r, nO, int(kk) % ncomp, float(s[kk]), float(rabs[kk]), hi - lo,
                                              float(np.abs(rv).max()), int(lp.ja[lo + int(np.abs(rv).argmax())]), int(lp.p2c[r]),
                                              bool(r in set(recv_rows_all.tolist()))), file=sys.stderr)
    del rows, cols, key, s, sabs, rabs, ratio
    # ---- (2) a sample of owned interface rows against the oracle ----
    recv_rows = np.unique(np.concatenate([r for _, r, _ in lp.recv])) if lp.recv else np.zeros(0, dtype=np.int64)
    nodes = np.unique(recv_rows // ncomp)
    if len(nodes) > sample_rows // ncomp:
        rng = np.random.default_rng(12345 + lp.comm.rank)
        nodes = np.sort(rng.choice(nodes, sample_rows // ncomp, replace=False))
    rows_checked, worst, bad = 0, 0.0, 0
    if len(nodes):
        # ext node ids of the sampled local nodes, the elements of the extended (= global around my partition) mesh touching them
        keys_loc = lp.node_keys_local
        ext_sorted = np.argsort(lp.key_of_ext)
        ext_of = ext_sorted[np.searchsorted(lp.key_of_ext[ext_sorted], keys_loc[nodes])]
        mark = np.zeros(len(lp.key_of_ext), dtype=bool)
        mark[ext_of] = True
        L = lp.ext_l2g_nodes
        esel = np.flatnonzero(mark[L].any(axis=1))
        sub = hm.Mesh(lp.ext_mesh.xyz, lp.ext_mesh.conn[esel])
        Xs = hm.FunctionSpace(sub, order, ncomp)
        sia, sja = Xs.graph()
        c2 = co
        if kind == orc.STIFFNESS and getattr(args, "form", "") == "kstiffness" and args.coeff == "idv":
            Kh = hm.FunctionSpace(sub, 1, 1)
            c2 = orc.Coeff(kind=orc.COEFF_FIELD, field_order=1, field_l2g=Kh.l2g,
                           field_values=1.0 + 0.5 * np.sin(2 * np.pi * Kh.dofPoints()[:, 0]))
        ref = orc.form2(kind, order, ncomp, quad, sub.xyz, sub.conn, Xs.l2g, sia, sja, c2)
        scale = orc.form2(kind | orc.ABS_FLAG, order, ncomp, quad, sub.xyz, sub.conn, Xs.l2g, sia, sja, c2)
        # sub-space node -> ext node -> cluster id
        sub_nodes = Xs.l2g[:, :n].astype(np.int64) // ncomp
        ext_of_sub = np.empty(Xs.nDof() // ncomp, dtype=np.int64)
        ext_of_sub[sub_nodes] = L[esel]
        clus_of_sub = lp.cluster_of_ext[ext_of_sub]
        sdiag = np.zeros(Xs.nDof())
        srows = np.repeat(np.arange(Xs.nDof()), np.diff(sia))
        on = srows == sja
        sdiag[srows[on]] = scale[on]
        sub_of_ext = np.full(len(lp.key_of_ext), -1, dtype=np.int64)
        sub_of_ext[ext_of_sub] = np.arange(len(ext_of_sub))
        hv = a.t_values[:nnzO].cpu().numpy() if nnzO < (1 << 27) else None
        for k, nd in enumerate(nodes):
            sn = sub_of_ext[ext_of[k]]
            for c in range(ncomp):
                r = int(nd) * ncomp + c
                lo, hi = int(ia[r]), int(ia[r + 1])
                got_cols = lp.ja[lo:hi]
                got = hv[lo:hi] if hv is not None else a.t_values[lo:hi].cpu().numpy()
                sr = sn * ncomp + c
                slo, shi = int(sia[sr]), int(sia[sr + 1])
                want_cols = clus_of_sub[sja[slo:shi] // ncomp] * ncomp + sja[slo:shi] % ncomp
                o = np.argsort(want_cols)
                if len(got_cols) != len(want_cols) or not np.array_equal(want_cols[o], got_cols):
                    bad += 1
                    continue
                tol = parity.RTOL * np.abs(ref[slo:shi][o]) + parity.FLOOR * np.sqrt(sdiag[sr] * sdiag[sja[slo:shi][o]])
                err = np.abs(got - ref[slo:shi][o])
                worst = max(worst, float((err / np.maximum(tol, 1e-300)).max()))
                bad += int((err > tol).any())
                rows_checked += 1
    t = torch.tensor([worst_inv, worst, float(bad), float(rows_checked), float(nO), float(n_inv_bad), float(n_zero_rows)],
                     device=dev, dtype=torch.float64)
    every = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(every, t)
    E = np.array([e.cpu().numpy() for e in every])
    ok = bool((E[:, 2] == 0).all() and (E[:, 5] == 0).all() and (E[:, 6] == 0).all() and (E[:, 1] <= 1.0).all())
    return {"ok": ok,
            "invariant": "per owned row and column component |sum_j a_ij| <= 1e-12 sum_j |a_ij| (constants / rigid translations in the kernel)",
            "invariant_rows": int(E[:, 4].sum()), "invariant_rows_failed": int(E[:, 5].sum()), "all_zero_rows": int(E[:, 6].sum()), "invariant_worst": float(E[:, 0].max()),
            "oracle_rows_checked": int(E[:, 3].sum()), "oracle_rows_failed": int(E[:, 2].sum()),
            "worst_err_over_tol": float(E[:, 1].max()),
            "oracle_rows": "owned interface rows (rows that received ghost contributions), <= %d per rank, entry by entry "
                           "against the oracle element loop over the elements around them; pattern of each row identical" % sample_rows,
            "per_rank_rows_checked": [int(x) for x in E[:, 3]]}
