#!/usr/bin/env python3
"""bench.py -- P2 tet Laplacian form2 assembly on B200 (BASELINE.json metric), one JSON line on stdout.

A "step" is one `a = integrate(elements(mesh), gradt(u)*trans(grad(v)))` over the whole mesh (zero + assemble).
Workload at N=1: BASELINE.json configs[1] -- P2 on the ~8M-element unit-cube tet mesh (m=110, 7 986 000 tets),
jittered vertices, generated (cell-major) element order.  Device-resident inputs (mesh, DofTable, CSR graph and the
element->CSR maps) are prepared untimed: they are Feel++'s setup (Pch, newMatrix).

  value  : elements/s, inputs resident in HBM, CUDA events on the launching stream
  e2e    : the same through feelgpu_form2_assemble_host: per step H2D of the vertex coordinates from pinned host memory
           (padded to one sector per vertex on the device) and D2H of the CSR values into pinned host memory
  --impl reference : the CPU element loop (oracle restatement of the reference; Feel++ itself cannot be built here)
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--m", type=int, default=110, help="cells per edge of the unit cube (110 -> 7 986 000 tets)")
    ap.add_argument("--order", type=int, default=2)
    ap.add_argument("--algo", default="auto", choices=["auto", "scatter", "rowplan", "eplan"])
    ap.add_argument("--shuffle", action="store_true", help="Fisher-Yates shuffled element order")
    ap.add_argument("--mesh", default=None, help="a Gmsh .msh file (tetrahedra) used instead of the synthetic cube; "
                    "one GPU only, reported as its own workload (never the configs[1] line)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


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


def workload(args, want_graph=True):
    """builds the host-side objects (untimed setup)"""
    import feelpp_b200 as fpp
    t0 = time.time()
    mesh = fpp.loadMesh(args.mesh) if args.mesh else fpp.unitCube(args.m, 0.15, args.shuffle)
    Xh = fpp.Pch(mesh, args.order)
    t1 = time.time()
    ia, ja = Xh.graph() if want_graph else (None, None)
    t2 = time.time()
    info = {"mesh_doftable_s": round(t1 - t0, 2), "graph_s": round(t2 - t1, 2)}
    return mesh, Xh, ia, ja, info


def config_dict(args, mesh, Xh, nnz):
    if args.mesh:
        return {"workload": "P%d scalar Laplacian form2 gradt(u)*trans(grad(v)) on the mesh file %s (not a BASELINE config)"
                            % (args.order, os.path.basename(args.mesh)),
                "elements": int(mesh.numElements()), "points": int(mesh.numPoints()), "dofs": int(Xh.nDof()),
                "nnz": int(nnz), "quadrature": "IMTetrahedra<%d>" % max(1, 2 * (args.order - 1)),
                "vertices": "file", "element_order": "file",
                "l2_policy": "inputs larger than L2 only if the mesh is (CSR values %.0f MB)" % (8e-6 * nnz)}
    return {"workload": "configs[1]: P%d scalar Laplacian form2 gradt(u)*trans(grad(v)), unit cube m=%d" % (args.order, args.m),
            "elements": int(mesh.numElements()), "points": int(mesh.numPoints()), "dofs": int(Xh.nDof()),
            "nnz": int(nnz), "quadrature": "IMTetrahedra<%d>" % max(1, 2 * (args.order - 1)),
            "vertices": "jittered 0.15h", "element_order": "shuffled" if args.shuffle else "generated",
            "l2_policy": "inputs larger than L2 (CSR values + element maps >> 126 MB)"}


def cpu_run(args, mesh, Xh, ia, ja, threads, sample_elts):
    """times the oracle element loop (restatement of integrator.hpp:1093-1207 + MatSetValues) on `threads` host threads"""
    from oracle import oracle as orc
    n = min(sample_elts, mesh.numElements())
    quad = max(1, 2 * (args.order - 1))
    a = np.zeros(int(ia[-1]))
    _, secs = orc.form2(orc.STIFFNESS, args.order, 1, quad, mesh.xyz, mesh.conn, Xh.l2g, ia, ja, a=a,
                        elt_range=(0, n), threads=threads)
    return n / secs, n, secs


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # the same global mesh as the GPU arm at this N (weak scaling: m grows with N^(1/3)); a step is a bounded sample of it
    if args.mesh and args.gpus > 1:
        raise SystemExit("--mesh runs on one GPU (the multi-GPU line is the weak-scaling cube)")
    args.m = int(round(args.m * max(args.gpus, 1) ** (1.0 / 3.0)))
    mesh, Xh, ia, ja, info = workload(args)
    cores = physical_cores()
    sample = min(mesh.numElements(), 400_000 * cores)
    for _ in range(args.warmup):
        cpu_run(args, mesh, Xh, ia, ja, cores, max(1, sample // 8))
    t = 0.0
    n = 0
    for _ in range(args.steps):
        _, ni, si = cpu_run(args, mesh, Xh, ia, ja, cores, sample)
        t += si
        n += ni
    value = n / t
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "mesh file" if args.mesh else "synthetic",
            "config": config_dict(args, mesh, Xh, ia[-1]),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "first %d elements of the mesh per step, %d threads, contiguous element blocks "
                                       "with owner/stash merge (oracle_form2_threads)" % (sample, cores)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


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
    from feelpp_b200 import elements, grad, gradt, integrate, trans
    from feelpp_b200._capi import Form2Desc, check, c_dblp, ptr
    from feelpp_b200 import engine as E

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1 and args.mesh:
        raise SystemExit("--mesh runs on one GPU (the multi-GPU line is the weak-scaling cube)")
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    algo = {"auto": fpp.ALGO_AUTO, "scatter": fpp.ALGO_SCATTER, "rowplan": fpp.ALGO_ROWPLAN, "eplan": fpp.ALGO_EPLAN}[args.algo]
    eng = fpp.Engine(local)
    if world == 1:
        mesh, Xh, ia, ja, info = workload(args)
        nElt, nnz = mesh.numElements(), int(ia[-1])
        t0 = time.time()
        a = fpp.form2(_test=Xh, _trial=Xh, engine=eng, algorithm=algo,
                      pattern_flags={"rowplan": 3, "scatter": 1}.get(args.algo, 0))
        info["pattern_create_s"] = round(time.time() - t0, 2)
        nElt_total, nnz_total, gmesh = nElt, nnz, mesh
    else:
        # weak scaling: the global mesh grows with the number of GPUs (m = 110 * N^(1/3): 16.1M / 32.2M / 63.9M tets),
        # recursive-bisection partition, one partition + owned row block per GPU, NCCL ghost-row exchange per step
        from feelpp_b200 import parallel as par, distributed as dd
        args.m = int(round(args.m * world ** (1.0 / 3.0)))
        t0 = time.time()
        gmesh = fpp.unitCube(args.m, 0.15, args.shuffle)
        part = par.partition_rcb(gmesh, world)
        t1 = time.time()
        lp = par.LocalProblem(gmesh, part, args.order, 1, par.Comm(rank, world, dist))
        t2 = time.time()
        Xh = dd.LocalSpace(lp)
        mesh = lp.mesh
        a = dd.DistributedBilinearForm(Xh, eng, dist, algorithm=algo)
        info = {"mesh_partition_s": round(t1 - t0, 2), "local_problem_s": round(t2 - t1, 2),
                "pattern_create_s": round(time.time() - t2, 2), "local_elements": int(mesh.numElements()),
                "owned_dofs": int(lp.nDofOwned), "ghost_dofs": int(lp.nDofWithGhost - lp.nDofOwned),
                "send_values": int(sum(c for _, _, c in lp.send))}
        ia = lp.ia
        nElt, nnz = mesh.numElements(), int(ia[-1])
        nElt_total = gmesh.numElements()
        tn = torch.tensor([lp.nnz_owned()], device="cuda", dtype=torch.int64)
        dist.all_reduce(tn)
        nnz_total = int(tn.item())
    info["plan"] = a.plan_stats()
    u = v = fpp.Element(Xh)
    integ = integrate(_range=elements(mesh), _expr=gradt(u) * trans(grad(v)))
    keep = E._Keep()
    desc = a._desc(integ, keep)
    stream = torch.cuda.current_stream()
    sptr = C.c_void_p(stream.cuda_stream)
    lib = eng.lib

    def step():
        if world > 1:
            a.assemble_desc(desc, 0, sptr)   # ghost rows land in their owners' memory while the kernel runs
            a.close()                        # NCCL barrier + owner-side add
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
        mine = torch.tensor([float(nElt), float(sum(c for _, _, c in lp.send)), ms / args.steps,
                             sorted(k[1] for k in kern_ms)[len(kern_ms) // 2]], device="cuda", dtype=torch.float64)
        every = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        info["per_rank"] = [{"elements": int(t[0].item()), "send_values": int(t[1].item()),
                             "step_ms": round(t[2].item(), 4), "kernel_ms": round(t[3].item(), 4)} for t in every]
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())
    value = nElt_total * args.steps / (ms * 1e-3)
    # clocks under load: 20 steps last ~26 ms and nvidia-smi samples every 100 ms, so the same step keeps running for
    # ~0.6 s more (the count is derived from the all-reduced time: every rank runs the same number of collectives)
    for _ in range(int(min(2000, max(50, 0.6 / (ms / args.steps * 1e-3))))):
        step()
    barrier()
    clocks = sampler.stop(tw0, time.time()) if sampler else None

    # ---- end to end: host buffers, H2D of the coordinates + D2H of the CSR values inside the timed region ----
    e2e = None
    if not args.no_e2e:
        # host buffers in, host buffers out: per step H2D of the (local) vertex coordinates from host memory, assembly
        # [+ NCCL ghost-row exchange], D2H of the (owned) CSR values into pinned host memory
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
        nd = Xh.nd
        # algorithmic bytes per element: conn + coords + int32 element->slot map + CSR values once
        Be = 16 + 24.0 * gmesh.numPoints() / nElt_total + 4 * nd * nd + 8.0 * nnz_total / nElt_total
        kms = sorted(k[1] for k in kern_ms)[len(kern_ms) // 2]
        zms = sorted(k[0] for k in kern_ms)[len(kern_ms) // 2]
        achieved = Be * nElt / (kms * 1e-3) / 1e9
        traffic = None     # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (same config only)
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r1_eplan_traffic.json")))
            if world == 1 and args.m == 110 and args.order == 2 and args.algo in ("auto", "eplan") and not args.shuffle and not args.mesh:
                traffic = tr["dram_bytes_per_launch"]
        except Exception:
            pass
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel_ms": kms, "zero_fill_ms": zms, "algorithmic_bytes_per_element": Be,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "6650 GB/s (of fallback)",
                "step_frac_vs_peak": Be * nElt_total / world / (ms / args.steps * 1e-3) / 1e9 / peak}
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            cores = physical_cores()
            sample = min(nElt, 400_000 * cores)
            cpu_run(args, mesh, Xh, ia, ja, cores, max(1, sample // 8))
            v, n, s = cpu_run(args, mesh, Xh, ia, ja, cores, sample)
            cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": "first %d elements, %d threads, %.1f s (oracle_form2_threads)" % (n, cores, s)}
        cfg = config_dict(args, gmesh, Xh, nnz_total)
        cfg["dofs"] = int(Xh.nDof()) if world == 1 else int(lp.nGlobalDof)
        cfg["algorithm"] = args.algo
        cfg["parallelism"] = "1 GPU" if world == 1 else ("%d GPUs: recursive-bisection element partition, owned row "
                                                         "blocks; every step hands the ghost rows to their owners (%s)"
                                                         % (world, "stored over NVLink peer memory by the assembly "
                                                            "kernel, NCCL barrier, owner-side add" if a.peer else
                                                            "NCCL send/recv of the ghost-row tail, owner-side add"))
        cfg["setup"] = info
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "mesh file" if args.mesh else "synthetic", "config": cfg,
                "nnz_per_s": nnz_total * args.steps / (ms * 1e-3), "roofline": roof, "cpu_baseline": cpu,
                "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)
