#!/usr/bin/env python3
"""bench.py -- form2 assembly on B200 for BASELINE.json's configs, one JSON line on stdout.

A "step" is one `a = integrate(elements(mesh), <form>)` over the whole mesh (zero + assemble [+ ghost-row exchange]).
Default workload (N=1): BASELINE.json configs[1] -- P2 scalar Laplacian on the ~8M-element unit-cube tet mesh (m=110,
7 986 000 tets), jittered vertices, generated (cell-major) element order: the configuration the metric is quoted on.
`--config` selects the other configurations of BASELINE.json:

  1  P1 scalar  gradt(u)*trans(grad(v))                                   m=55   (configs[0], with the single-core CPU time)
  2  P2 scalar  gradt(u)*trans(grad(v))                                   m=110  (configs[1], the headline line)
  3  P3 scalar  k*inner(gradt(u),grad(v)), k = expr("1+x*y+z^2") (Q=24)   m=87   (configs[2]; --coeff idv: k = idv(P1 field), Q=14)
  4  P2 vector  inner(lambda*trace(sym(gradt u))*Id + 2 mu sym(gradt u), grad v), E=1e8, nu=0.25
                                                                          m=139 on 8 GPUs (configs[3]); at N GPUs its
                                                                          per-GPU share m = 139 (N/8)^(1/3): 70 / 88 / 110 / 139

Device-resident inputs (mesh, DofTable, CSR graph and the element->CSR maps / plans) are prepared untimed: they are
Feel++'s setup (Pch, newMatrix); their cost is reported in run_info.setup together with the break-even count.

  value  : elements/s, inputs resident in HBM, CUDA events on the launching stream
  e2e    : the same through feelgpu_form2_assemble_host: per step H2D of the vertex coordinates from pinned host memory
           (padded to one sector per vertex on the device) and D2H of the CSR values into pinned host memory
  --impl reference : the CPU element loop (oracle restatement of the reference; Feel++ itself cannot be built here)
  --scaling strong : the global mesh is fixed (--m) and split over the GPUs; weak (default): m grows with N^(1/3)
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "P2 tet Laplacian form2 assembly throughput"
UNIT = "elements/s"

# per config: Lagrange order, components, form, cells per edge at ONE GPU (weak scaling multiplies by N^(1/3))
CONFIGS = {
    1: {"order": 1, "ncomp": 1, "form": "stiffness", "m1": 55.0, "name": "configs[0]"},
    2: {"order": 2, "ncomp": 1, "form": "stiffness", "m1": 110.0, "name": "configs[1]"},
    3: {"order": 3, "ncomp": 1, "form": "kstiffness", "m1": 87.0, "name": "configs[2]"},
    4: {"order": 2, "ncomp": 3, "form": "elasticity", "m1": 69.5, "name": "configs[3]"},
}
E_MOD, NU = 1e8, 0.25          # benchmarks/ATTIC/perf/assembly.cpp:86-89
LAME_LAMBDA = E_MOD * NU / ((1 + NU) * (1 - 2 * NU))
LAME_MU = E_MOD / (2 * (1 + NU))
K_EXPR = "1+x*y+z^2"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS))
    ap.add_argument("--coeff", default="expr", choices=["expr", "idv"], help="config 3: closed-form k or a P1 field idv(k)")
    ap.add_argument("--m", type=int, default=None, help="cells per edge of the unit cube (overrides the config's size)")
    ap.add_argument("--order", type=int, default=None, help="Lagrange order (overrides the config's)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--partition", default="auto", choices=["auto", "metis", "rcb"],
                    help="element partition at N>1: METIS on the face-adjacency graph as the reference does, or coordinate bisection")
    ap.add_argument("--algo", default="auto", choices=["auto", "scatter", "rowplan", "eplan", "rplan"])
    ap.add_argument("--shuffle", action="store_true", help="Fisher-Yates shuffled element order")
    ap.add_argument("--mesh", default=None, help="a Gmsh .msh file (tetrahedra) used instead of the synthetic cube; "
                    "one GPU only, reported as its own workload (never a BASELINE config line)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the multi-GPU parity block (N>1)")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    if args.order is None:
        args.order = cfg["order"]
    args.ncomp, args.form = cfg["ncomp"], cfg["form"]
    # cells per edge of the GLOBAL mesh: weak scaling grows the one-GPU size with N^(1/3), strong scaling keeps it
    m1 = float(args.m) if args.m is not None else cfg["m1"]
    world = max(args.gpus, 1)
    args.m_global = int(round(m1 * world ** (1.0 / 3.0))) if args.scaling == "weak" else int(round(m1))
    return args


def physical_cores():
    try:
        out = subprocess.check_output(["lscpu", "-p=core,socket"], text=True)
        cores = {l for l in out.splitlines() if l and not l.startswith("#")}
        n = len(cores)
    except Exception:
        n = os.cpu_count() or 1
    try:
        n = min(n, len(os.sched_getaffinity(0)))
    except Exception:
        pass
    return max(1, n)


# ---- the workload: space, expression, oracle arguments, algorithmic bytes -----------------------------------------------
def quad_order(args):
    k = args.order
    if args.form == "kstiffness":
        return 2 * (k - 1) + (2 if args.coeff == "expr" else 1)     # degree-2 polynomial / P1 field (expr.hpp:1075)
    return max(1, 2 * (k - 1))


def make_space(args, mesh):
    import feelpp_b200 as fpp
    return fpp.Pchv(mesh, args.order) if args.ncomp == 3 else fpp.Pch(mesh, args.order)


def field_values(mesh_xyz):
    """nodal values of the P1 conductivity field of config 3: 1 + 0.5 sin(2 pi x)"""
    return 1.0 + 0.5 * np.sin(2 * np.pi * mesh_xyz[:, 0])


def make_expr(args, Xh, mesh):
    """the expression of the step in the harness' Feel++ language (feelpp_b200/vf.py)"""
    import feelpp_b200 as fpp
    from feelpp_b200 import Id, expr, grad, gradt, idv, inner, sym, trace, trans
    u = v = fpp.Element(Xh)
    if args.form == "stiffness":
        return gradt(u) * trans(grad(v))
    if args.form == "kstiffness":
        if args.coeff == "expr":
            return expr(K_EXPR) * inner(gradt(u), grad(v))
        Kh = fpp.Pch(mesh, 1)
        kf = fpp.Element(Kh, Kh.interpolate(lambda x, y, z: 1.0 + 0.5 * np.sin(2 * np.pi * x)))
        return idv(kf) * inner(gradt(u), grad(v))
    return inner(LAME_LAMBDA * trace(sym(gradt(u))) * Id() + 2 * LAME_MU * sym(gradt(u)), grad(v))


def oracle_args(args, mesh):
    """(kind, coefficient) of the same form for the oracle element loop"""
    import feelpp_b200 as fpp
    from oracle import oracle as orc
    if args.form == "stiffness":
        return orc.STIFFNESS, orc.Coeff()
    if args.form == "kstiffness":
        if args.coeff == "expr":
            prog = fpp.GinacProgram(K_EXPR)
            return orc.STIFFNESS, orc.Coeff(kind=orc.COEFF_EXPR, ops=prog.ops, consts=prog.consts)
        Kh = fpp.Pch(mesh, 1)
        return orc.STIFFNESS, orc.Coeff(kind=orc.COEFF_FIELD, field_order=1, field_l2g=Kh.l2g,
                                        field_values=Kh.interpolate(lambda x, y, z: 1.0 + 0.5 * np.sin(2 * np.pi * x)))
    return orc.ELASTICITY, orc.Coeff(lam=LAME_LAMBDA, mu=LAME_MU)


def form_text(args):
    if args.form == "stiffness":
        return "gradt(u)*trans(grad(v))"
    if args.form == "kstiffness":
        return ("expr(\"%s\")" % K_EXPR if args.coeff == "expr" else "idv(k), k in P1 = 1+0.5 sin(2 pi x)") + "*inner(gradt(u),grad(v))"
    return "inner(lambda*trace(sym(gradt(u)))*Id + 2 mu sym(gradt(u)), grad(v)), E=1e8, nu=0.25"


def bytes_per_element(args, npts, nelt, nnz):
    """algorithmic bytes per element: connectivity + coordinates (each point once) + int32
    element->slot map (node level: the 3x3 blocks of a product space are contiguous) + every CSR value written once
    [+ coefficient: the field's l2g and nodal values]"""
    n = (args.order + 1) * (args.order + 2) * (args.order + 3) // 6
    nd = n * args.ncomp
    slotmap = 4 * n * n if args.ncomp == 3 else 4 * nd * nd
    be = 16 + 24.0 * npts / nelt + slotmap + 8.0 * nnz / nelt
    if args.form == "kstiffness" and args.coeff == "idv":
        be += 16 + 8.0 * npts / nelt
    return be


def fp64_instr_per_element(args):
    """DFMA-class instructions per element of the row-thread kernel for the forms it serves (a count of the kernel's own
    arithmetic): these forms are bound by the FP64 pipe, not by HBM.  None for the entity-plan forms."""
    n = (args.order + 1) * (args.order + 2) * (args.order + 3) // 6
    nq = {1: 1, 2: 4, 3: 5, 4: 11, 5: 14, 6: 24}.get(quad_order(args))
    if nq is None:
        return None
    if args.form == "kstiffness":
        return n * (3 * n * nq + 12 * nq)          # per row: s = w J k G grad(phi_i) (12 per point), then 3 DFMA per (column, point)
    if args.form == "elasticity" or args.ncomp == 3 and args.form not in ("stiffness",):
        return n * (9 * nq + n * (18 * nq + 21))   # per row: gradients (9 per point); per column 18 per point + the 3x3 combination
    return None


def workload(args, want_graph=True, eng=None):
    """builds the host-side objects (untimed setup); with an engine the COUPLED graph is built on the device
    (feelgpu_graph_create, bit-exact with the host graph: tests/test_gpu_parity.py)"""
    import feelpp_b200 as fpp
    t0 = time.time()
    mesh = fpp.loadMesh(args.mesh) if args.mesh else fpp.unitCube(args.m_global, 0.15, args.shuffle)
    Xh = make_space(args, mesh)
    t1 = time.time()
    if not want_graph:
        ia, ja = None, None
    elif eng is not None:
        dg = eng.device_graph(Xh)  # kept on the device: form2() builds its pattern from it (feelgpu_pattern_create_from_graph)
        ia, ja = dg.ia, dg.ja
    else:
        ia, ja = Xh.graph()
    t2 = time.time()
    info = {"mesh_doftable_s": round(t1 - t0, 2), "graph_s": round(t2 - t1, 2),
            "graph": "device (feelgpu_graph_create + copy to the host)" if eng is not None and want_graph else "host"}
    if eng is not None and want_graph:
        info["_device_graph"] = dg
    return mesh, Xh, ia, ja, info


def config_dict(args, npts, nelt, ndof, nnz):
    """identical in both arms (the driver compares the dicts): what is computed, not how"""
    sp = "P%d %s" % (args.order, "vector (3 components)" if args.ncomp == 3 else "scalar")
    if args.mesh:
        return {"workload": "%s form2 %s on the mesh file %s (not a BASELINE config)" % (sp, form_text(args), os.path.basename(args.mesh)),
                "elements": int(nelt), "points": int(npts), "dofs": int(ndof), "nnz": int(nnz),
                "quadrature": "IMTetrahedra<%d>" % quad_order(args), "vertices": "file", "element_order": "file",
                "l2_policy": "inputs larger than L2 only if the mesh is (CSR values %.0f MB)" % (8e-6 * nnz)}
    return {"workload": "%s: %s form2 %s, unit cube m=%d" % (CONFIGS[args.config]["name"], sp, form_text(args), args.m_global),
            "elements": int(nelt), "points": int(npts), "dofs": int(ndof), "nnz": int(nnz),
            "quadrature": "IMTetrahedra<%d>" % quad_order(args),
            "vertices": "jittered 0.15h", "element_order": "shuffled" if args.shuffle else "generated",
            "l2_policy": "inputs larger than L2 (CSR values + element maps >> 126 MB)"}


def metric_name(args):
    if args.config == 2 and args.order == 2:
        return METRIC
    return "P%d%s tet form2 assembly throughput (%s)" % (args.order, "v" if args.ncomp == 3 else "", args.form)


def cpu_run(args, mesh, Xh, ia, ja, threads, sample_elts):
    """times the oracle element loop (restatement of integrator.hpp:1093-1207 + MatSetValues) on `threads` host threads"""
    from oracle import oracle as orc
    n = min(sample_elts, mesh.numElements())
    kind, co = oracle_args(args, mesh)
    a = np.zeros(int(ia[-1]))
    _, secs = orc.form2(kind, args.order, args.ncomp, quad_order(args), mesh.xyz, mesh.conn, Xh.l2g, ia, ja, co, a=a,
                        elt_range=(0, n), threads=threads)
    return n / secs, n, secs


def cpu_sample_size(args, cores, nelt):
    """elements per timed CPU step: about 10-30 s of host work in total, scaled by the cost of an element of the form"""
    per_core = {"stiffness": 400_000 if args.order <= 2 else 60_000, "kstiffness": 50_000, "elasticity": 50_000}[args.form]
    if args.order == 1:
        per_core = 1_000_000
    return int(min(nelt, per_core * cores))


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # the same global mesh as the GPU arm at this N; a step is a bounded sample of it
    if args.mesh and args.gpus > 1:
        raise SystemExit("--mesh runs on one GPU (the multi-GPU line is the cube)")
    mesh, Xh, ia, ja, info = workload(args)
    cores = physical_cores()
    sample = cpu_sample_size(args, cores, mesh.numElements())
    for _ in range(args.warmup):
        cpu_run(args, mesh, Xh, ia, ja, cores, max(1, sample // 8))
    t = 0.0
    n = 0
    for _ in range(args.steps):
        _, ni, si = cpu_run(args, mesh, Xh, ia, ja, cores, sample)
        t += si
        n += ni
    value = n / t
    line = {"impl": "reference", "metric": metric_name(args), "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
            "data": "mesh file" if args.mesh else "synthetic",
            "config": config_dict(args, mesh.numPoints(), mesh.numElements(), Xh.nDof(), ia[-1]),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "first %d elements of the mesh per step, %d threads, contiguous element blocks "
                                       "with owner/stash merge (oracle_form2_threads)" % (sample, cores)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def pin_to_gpu_numa(local):
    """bind this rank's threads to the CPUs of its GPU's NUMA node BEFORE any pinned host buffer is allocated: page-locked
    memory is then placed on that node (first touch) and the D2H/H2D copies of the end-to-end step do not cross sockets.
    Returns a short description for run_info."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local)
        bus = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
        if node < 0:
            return "gpu %s: no NUMA node reported" % bus
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if not allowed:
            return "gpu %s on NUMA node %d: none of its CPUs is in this process' affinity mask" % (bus, node)
        os.sched_setaffinity(0, allowed)
        return "gpu %s on NUMA node %d: bound to %d of its CPUs" % (bus, node, len(allowed))
    except Exception as ex:      # best effort: containers may hide sysfs
        return "not pinned (%s)" % type(ex).__name__


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def wait_first(self, timeout=8.0):
        """nvidia-smi needs a moment to come up: do not start the timed region before it prints"""
        t0 = time.time()
        while self.proc and not self.rows and time.time() - t0 < timeout:
            time.sleep(0.02)

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1] or [r for _, r in self.rows[-3:]]
        sm = sorted(float(r[0]) for r in rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 3 + k and r[3 + k] == "Active" for r in rows)]
        smax = float(rows[0][1]) if rows and rows[0][1].replace(".", "").isdigit() else None
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": smax, "reasons": reasons, "samples": len(rows),
                "window": "timed region and the same step repeated back to back for %.2f s after it" % (t1 - t0)}


def gpu_arm(args):
    import torch
    import torch.distributed as dist
    import feelpp_b200 as fpp
    from feelpp_b200 import elements, integrate
    from feelpp_b200._capi import check, c_dblp
    from feelpp_b200 import engine as E

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    numa = pin_to_gpu_numa(local)
    if world > 1 and args.mesh:
        raise SystemExit("--mesh runs on one GPU (the multi-GPU line is the cube)")
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    algo = {"auto": fpp.ALGO_AUTO, "scatter": fpp.ALGO_SCATTER, "rowplan": fpp.ALGO_ROWPLAN, "eplan": fpp.ALGO_EPLAN,
            "rplan": fpp.ALGO_RPLAN}[args.algo]
    eng = fpp.Engine(local)
    lp = None
    if world == 1:
        mesh, Xh, ia, ja, info = workload(args, eng=eng)
        nElt, nnz = mesh.numElements(), int(ia[-1])
        t0 = time.time()
        # FEELGPU_PATTERN_ASYNC: pattern_create returns once the row-thread plan stands (assembly can start); the entity plan
        # of P1/P2 spaces is built by a background thread and takes over when it is ready.  The timed steps wait for it.
        dg = info.pop("_device_graph")
        a = fpp.form2(_test=Xh, _trial=Xh, engine=eng, algorithm=algo, device_graph=dg,
                      pattern_flags={"rowplan": 3, "scatter": 1, "rplan": 9}.get(args.algo, fpp.PATTERN_ASYNC))
        dg.destroy()
        info["pattern_create_s"] = round(time.time() - t0, 2)
        _, eplan_s = a.plans_ready(wait=True)
        info["plans_ready_s"] = round(time.time() - t0, 2)
        info["eplan_build_s"] = round(eplan_s, 2)
        nElt_total, nnz_total, gmesh, ndof_total = nElt, nnz, mesh, Xh.nDof()
        part_name = "1 GPU"
    else:
        # one partition + owned row block per GPU (the reference's MPI layout), ghost rows handed to their owners per step
        from feelpp_b200 import parallel as par, distributed as dd
        t0 = time.time()
        gmesh = fpp.unitCube(args.m_global, 0.15, args.shuffle)
        # the partitioner runs on the master rank only and the result is broadcast, as in the reference
        # (partitionermetis_impl.hpp:114-116 "Invoke METIS, but only on processor 0. Then broadcast")
        tp = torch.zeros(gmesh.numElements(), dtype=torch.int32, device="cuda")
        name_box = [None]
        if rank == 0:
            part, name_box[0] = par.partition(gmesh, world, args.partition)
            tp.copy_(torch.from_numpy(part))
        dist.broadcast(tp, src=0)
        dist.broadcast_object_list(name_box, src=0)
        part, part_name = tp.cpu().numpy(), name_box[0]
        del tp
        t1 = time.time()
        lp = par.LocalProblem(gmesh, part, args.order, args.ncomp, par.Comm(rank, world, dist))
        t2 = time.time()
        Xh = dd.LocalSpace(lp)
        mesh = lp.mesh
        a = dd.DistributedBilinearForm(Xh, eng, dist, algorithm=algo)
        info = {"mesh_partition_s": round(t1 - t0, 2), "local_problem_s": round(t2 - t1, 2),
                "pattern_create_s": round(time.time() - t2, 2), "local_elements": int(mesh.numElements()),
                "owned_dofs": int(lp.nDofOwned), "ghost_dofs": int(lp.nDofWithGhost - lp.nDofOwned),
                "send_values": int(lp.send_values()), "neighbours": int(lp.num_neighbours())}
        ia = lp.ia
        nElt, nnz = mesh.numElements(), int(ia[-1])
        nElt_total = gmesh.numElements()
        tn = torch.tensor([lp.nnz_owned()], device="cuda", dtype=torch.int64)
        dist.all_reduce(tn)
        nnz_total = int(tn.item())
        ndof_total = int(lp.nGlobalDof)
    info["plan"] = a.plan_stats()
    integ = integrate(_range=elements(mesh), _expr=make_expr(args, Xh, mesh))
    keep = E._Keep()
    desc = a._desc(integ, keep)
    stream = torch.cuda.current_stream()
    sptr = C.c_void_p(stream.cuda_stream)
    lib = eng.lib

    def step():
        if world > 1:
            a.assemble_desc(desc, 0, sptr)   # ghost rows land in their owners' memory while the kernel runs (entity plan)
            a.close()                        # completion barrier / NCCL exchange + owner-side add
        else:
            check(lib.feelgpu_form2_assemble(a.pat, C.byref(desc), 0, a.d_values, sptr), eng.ctx)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    sampler = ClockSampler(local) if rank == 0 else None
    for _ in range(max(args.warmup, 3)):
        step()
    if sampler:
        sampler.wait_first()
    barrier()
    check(lib.feelgpu_set_timing(eng.ctx, 1), eng.ctx)
    l0 = eng.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kern_ms = []
    tw0 = time.time()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    barrier()
    tw1 = time.time()
    ms = ev0.elapsed_time(ev1)
    launches = eng.launch_count() - l0
    # dominant-kernel duration, measured live with CUDA events on the launching stream (a few extra steps)
    tm = (C.c_double * 3)()
    for _ in range(min(args.steps, 10)):
        step()
        check(lib.feelgpu_last_timing(eng.ctx, tm), eng.ctx)
        kern_ms.append((tm[0], tm[1]))
    check(lib.feelgpu_set_timing(eng.ctx, 0), eng.ctx)
    tmax = torch.tensor([ms], device="cuda", dtype=torch.float64)
    if world > 1:
        # per rank: elements, ghost-row values sent, step and kernel milliseconds (the step time reported is the maximum)
        mine = torch.tensor([float(nElt), float(lp.send_values()), ms / args.steps,
                             sorted(k[1] for k in kern_ms)[len(kern_ms) // 2], float(lp.num_neighbours())],
                            device="cuda", dtype=torch.float64)
        every = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        info["per_rank"] = [{"elements": int(t[0].item()), "send_values": int(t[1].item()),
                             "step_ms": round(t[2].item(), 4), "kernel_ms": round(t[3].item(), 4),
                             "neighbours": int(t[4].item())} for t in every]
        ne = [p["elements"] for p in info["per_rank"]]
        info["imbalance"] = round(max(ne) * world / float(sum(ne)), 4)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())
    value = nElt_total * args.steps / (ms * 1e-3)
    # clocks under load: the timed region may last only ~26 ms and nvidia-smi samples every 100 ms, so the same step keeps
    # running for ~0.6 s more (the count is derived from the all-reduced time: every rank runs the same number of collectives)
    for _ in range(int(min(2000, max(10, 0.6 / (ms / args.steps * 1e-3))))):
        step()
    barrier()
    clocks = sampler.stop(tw0, time.time()) if sampler else None

    # ---- multi-GPU parity: every owned row after close(), against the invariant and the oracle ----
    parity = None
    if world > 1 and not args.no_parity:
        import bench_parity
        step()
        barrier()
        parity = bench_parity.check(args, lp, gmesh, a, dist, oracle_args(args, gmesh), quad_order(args))

    # ---- end to end: host buffers, H2D of the coordinates + D2H of the CSR values inside the timed region ----
    e2e = None
    if not args.no_e2e:
        # host buffers in, host buffers out: per step H2D of the (local) vertex coordinates from host memory, assembly
        # [+ ghost-row exchange], D2H of the (owned) CSR values into pinned host memory
        nown = nnz if world == 1 else lp.nnz_owned()
        hp = C.c_void_p()
        check(lib.feelgpu_host_alloc_pinned(eng.ctx, 8 * max(nown, 1), C.byref(hp)), eng.ctx)
        hostvals = C.cast(hp, c_dblp)
        mh = eng.mesh(mesh)
        # the step's input, the vertex coordinates, waits in pinned host memory ([nPts][3] doubles, the caller's layout)
        hx = C.c_void_p()
        check(lib.feelgpu_host_alloc_pinned(eng.ctx, mesh.xyz.nbytes, C.byref(hx)), eng.ctx)
        C.memmove(hx, mesh.xyz.ctypes.data, mesh.xyz.nbytes)
        hostxyz = C.cast(hx, c_dblp)
        ksteps = max(2, min(args.steps, 5))
        for it in range(1 + ksteps):
            if it == 1:
                barrier()
                te0 = time.perf_counter()
            check(lib.feelgpu_mesh_update_points(mh, hostxyz, None if world == 1 else sptr), eng.ctx)
            if world == 1:
                check(lib.feelgpu_form2_assemble_host(a.pat, C.byref(desc), 0, hostvals), eng.ctx)
            else:
                step()
                torch.cuda.synchronize()
                check(lib.feelgpu_memcpy_d2h(eng.ctx, hp, a.d_values, 8 * nown), eng.ctx)
        barrier()
        te = (time.perf_counter() - te0) / ksteps
        tt = torch.tensor([te], device="cuda", dtype=torch.float64)
        tb = torch.tensor([mesh.xyz.nbytes, 8 * nown], device="cuda", dtype=torch.int64)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(tb)
        te = float(tt.item())
        e2e = {"value": nElt_total / te, "unit": UNIT, "h2d_bytes_per_step": int(tb[0].item()),
               "d2h_bytes_per_step": int(tb[1].item()), "ms_per_step": 1e3 * te, "steps": ksteps,
               "timer": "host perf_counter around synchronous host-buffer calls, max over ranks"}
        lib.feelgpu_host_free_pinned(eng.ctx, hp)
        lib.feelgpu_host_free_pinned(eng.ctx, hx)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        Be = bytes_per_element(args, gmesh.numPoints(), nElt_total, nnz_total)
        kms = sorted(k[1] for k in kern_ms)[len(kern_ms) // 2]
        zms = sorted(k[0] for k in kern_ms)[len(kern_ms) // 2]
        achieved = Be * nElt / (kms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "traffic_source": "not measured in this run (ncu captures: profiles/)",
                "kernel_ms": kms, "zero_fill_ms": zms, "algorithmic_bytes_per_element": Be,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "6650 GB/s (of fallback)",
                "step_frac_vs_peak": Be * nElt_total / world / (ms / args.steps * 1e-3) / 1e9 / peak}
        try:
            fi = fp64_instr_per_element(args)
            if fi and args.algo in ("auto", "rplan"):
                # informational: the pipe that bounds this kernel (ncu: DRAM at 5 %, profiles/r2_rplan_summary.md)
                mhz = float((clocks or {}).get("sm_max_mhz") or 1965.0)
                sms = torch.cuda.get_device_properties(local).multi_processor_count
                pk = sms * 64 * 2 * mhz * 1e6 / 1e12
                ach = 2.0 * fi * nElt / (kms * 1e-3) / 1e12
                roof["fp64"] = {"achieved": ach, "peak": pk, "unit": "TFLOP/s", "frac": ach / pk,
                                "instructions_per_element": fi,
                                "peak_source": "%d SMs x 64 DFMA/clk x 2 x %.0f MHz (nominal pipe rate, not measured)" % (sms, mhz)}
        except Exception as ex:      # never lose the line over the informational block
            roof["fp64"] = {"error": str(ex)}
        cpu = None
        # host setup until the steady-state plans stand: mesh + DofTable, graph [partition, local problem], pattern + plans
        setup_s = sum(info.get(k, 0.0) for k in ("mesh_doftable_s", "graph_s", "mesh_partition_s", "local_problem_s")) + \
            info.get("plans_ready_s", info.get("pattern_create_s", 0.0))
        if not args.no_cpu_baseline and world == 1:
            cores = physical_cores()
            sample = cpu_sample_size(args, cores, nElt)
            cpu_run(args, mesh, Xh, ia, ja, cores, max(1, sample // 8))
            v, n, s = cpu_run(args, mesh, Xh, ia, ja, cores, sample)
            cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": "first %d elements, %d threads, %.1f s (oracle_form2_threads)" % (n, cores, s)}
            if args.config == 1:
                # BASELINE.json configs[0]: "single process on CPU" -- the whole mesh on one core
                v1, n1, s1 = cpu_run(args, mesh, Xh, ia, ja, 1, nElt)
                cpu["single_core"] = {"value": v1, "unit": UNIT, "cores": 1, "elements": n1, "seconds": round(s1, 3)}
            # break-even: assemblies on one mesh+space after which engine setup + k GPU steps beat k CPU assemblies
            t_cpu = nElt / v
            t_gpu = (e2e["ms_per_step"] if e2e else ms / args.steps) * 1e-3
            info["engine_setup_s"] = round(info.get("pattern_create_s", 0.0), 2)
            info["breakeven_assemblies"] = (round(info["engine_setup_s"] / (t_cpu - t_gpu), 1) if t_cpu > t_gpu else None)
        info["setup_s"] = round(setup_s, 2)
        run_info = {"algorithm": args.algo,
                    "parallelism": part_name if world == 1 else (
                        "%d GPUs: %s, owned row blocks; every step hands the ghost rows to their owners (%s)"
                        % (world, part_name, "stored over NVLink peer memory by the assembly kernel, NCCL barrier, "
                           "owner-side add" if a.peer else "NCCL send/recv of the ghost rows, owner-side add")),
                    "host_affinity": numa, "setup": info}
        line = {"metric": metric_name(args), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
                "data": "mesh file" if args.mesh else "synthetic",
                "config": config_dict(args, gmesh.numPoints(), nElt_total, ndof_total, nnz_total),
                "nnz_per_s": nnz_total * args.steps / (ms * 1e-3), "roofline": roof, "cpu_baseline": cpu,
                "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "run_info": run_info}
        if parity is not None:
            line["parity"] = parity
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and not parity.get("ok", False):
        sys.exit(3)


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)
