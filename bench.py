#!/usr/bin/env python
"""Headline benchmark: GDoF/s of the matrix-free Laplacian action (and CG iteration).

    python bench.py --gpus N --steps K --warmup W            # our arm (CUDA, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K ...  # reference CPU kernels on host cores

A "step" is one operator application y = A u (src/laplacian.hpp:355-403) over the whole local
mesh; the metric is the reference's own `ndofs_global * nreps / (1e9 * seconds)`
(src/main.cpp:129-130).  Default workload: BASELINE.json's headline configuration, Q3, qmode 1
(GLL), fp64, 300M dofs per GPU on the `compute_mesh_size` box (weak scaling: the global box is
re-derived from ndofs x ranks exactly as src/main.cpp:237-238,272-273 does).  The same run also
times `steps` CG iterations (src/cg.hpp:89-169) and reports them under "cg".

Rank 0 prints ONE JSON line.  Timing: CUDA events on the launching stream, barrier +
synchronize on both sides, max over ranks.  The inputs (G alone is 66 GB) are far larger than
the 126 MB L2, so no explicit L2 flush is needed between iterations.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


# ----------------------------------------------------------------------------- byte / flop model
def algorithmic_bytes_per_apply(ndofs_local, ncells, degree, nq, tsize):
    """SURVEY.md section 8(d), variant "G precomputed": read u + write y once (2T per dof),
    dofmap (4 nd^3 per cell), G (6 T nq^3 per cell), BC marker (1 per dof)."""
    nd3 = (degree + 1) ** 3
    return 2 * tsize * ndofs_local + 4 * nd3 * ncells + 6 * tsize * nq ** 3 * ncells + ndofs_local


def algorithmic_bytes_per_apply_onthefly(ndofs_local, ncells, nnodes, degree, tsize):
    """SURVEY.md section 8(d), variant "G on the fly": the G term is replaced by the vertex coordinates
    (3T per geometry node) and the geometry dofmap (32 bytes per cell)."""
    nd3 = (degree + 1) ** 3
    return 2 * tsize * ndofs_local + 4 * nd3 * ncells + 3 * tsize * nnodes + 32 * ncells + ndofs_local


def algorithmic_flops_per_cell(degree, nq, qmode):
    nd = degree + 1
    return 2 * (2 * (nq * nd ** 3 + nq ** 2 * nd ** 2 + nq ** 3 * nd)) * qmode + 12 * nq ** 4 + 18 * nq ** 3


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms during the timed region (20 steps of the headline
    workload last about 0.12 s now)."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.proc, self.lines = None, []
        sel = str(device_index)
        try:
            import torch
            uuid = getattr(torch.cuda.get_device_properties(device_index), "uuid", None)
            if uuid is not None:
                sel = "GPU-" + str(uuid)
        except Exception:
            pass
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if not sel.startswith("GPU-") and vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if device_index < len(ids):
                sel = ids[device_index]
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", sel, f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                smax.append(float(parts[1]))
                power.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smax), "power_w_max": max(power),
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------- CPU arm
def cpu_reference_run(degree, bits, steps, warmup, target_seconds, cells_per_dim=None):
    """The reference's own CPU implementation of the path (oracle/_ref: unmodified
    stiffness_operator_cpu, src/laplacian_cpu.hpp:57-146; qmode 0 is all it supports) on every
    host core: one independent instance per core on its own sub-box, which emulates the
    reference's `mpirun -n <cores>` (one single-threaded rank per core) without halo cost.
    Falls back to the plain-C port (oracle/port.c) when oracle/_ref is absent.
    Returns (GDoF/s, cores, kind, sample description, steps run, seconds)."""
    from oracle import native
    from oracle.boxmesh import create_box_mesh
    from oracle.tables import Tables
    dt = np.float64 if bits == 64 else np.float32
    ref = native.ref_lib(False)
    port = native.port_lib()
    kind = "reference" if ref is not None else "port"
    if ref is None and port is None:
        raise RuntimeError("neither oracle/_ref nor oracle/libport.so is built")
    cores = os.cpu_count() or 1
    qmode = 0
    t = Tables(degree, qmode)
    nc = cells_per_dim or max(2, int(round((24 ** 3 * 27 / degree ** 3) ** (1 / 3))))
    mesh = create_box_mesh([nc, nc, nc], degree)
    cells = np.concatenate([mesh.lcells, mesh.bcells]).astype(np.int32)
    xg = mesh.xgeom.astype(dt)
    dphi, w = t.dphi_geometry.astype(dt), t.weights3d.astype(dt)
    phi0, dphi1 = t.phi0.astype(dt), t.dphi1.astype(dt)
    if ref is not None:
        G = native.ref_geometry(ref, t.nq, xg, mesh.geometry_dofmap, dphi, w, cells)
    else:
        G = native.port_geometry(port, t.nq, xg, mesh.geometry_dofmap, dphi, w, cells)
    consts = np.full(mesh.ncells, 2.0, dtype=dt)
    rng = np.random.default_rng(1234)
    insts = []
    for _ in range(cores):   # private copies: every emulated rank streams its own arrays
        insts.append(dict(u=rng.uniform(-1, 1, mesh.ndofs_local).astype(dt),
                          y=np.zeros(mesh.ndofs_local, dtype=dt), G=G.copy(),
                          dm=mesh.cell_dofmap.copy()))

    def one_apply(inst):
        inst["y"][:] = 0
        if ref is not None:
            native.ref_stiffness(ref, degree, inst["u"], inst["y"], inst["G"], consts, phi0, dphi1,
                                 inst["dm"], cells, mesh.bc_marker)
        else:
            native.port_stiffness(port, t.nd, t.nq, True, inst["u"], inst["y"], inst["G"], consts,
                                  phi0, dphi1, inst["dm"], cells, mesh.bc_marker)

    def run_steps(nsteps):
        def work(inst):
            for _ in range(nsteps):
                one_apply(inst)
        th = [threading.Thread(target=work, args=(i,)) for i in insts]
        t0 = time.perf_counter()
        for x in th:
            x.start()
        for x in th:
            x.join()
        return time.perf_counter() - t0

    t1 = run_steps(max(1, warmup))           # warm-up, also calibrates the step count
    per_step = t1 / max(1, warmup)
    if target_seconds is not None:
        steps = max(1, min(steps, int(target_seconds / max(per_step, 1e-9))))
    secs = run_steps(steps)
    gdofs = mesh.ndofs_global * cores * steps / (1e9 * secs)
    cpu_reference_run.last = {"qmode": 0, "cells_per_instance": [nc, nc, nc], "ndofs_per_instance": int(mesh.ndofs_global),
                              "instances": cores}
    sample = (f"{cores} independent instances (one per core) of a {nc}x{nc}x{nc}-cell Q{degree} box, "
              f"{mesh.ndofs_global} dofs each, qmode 0 (the reference CPU kernel has no qmode 1), "
              f"fp{bits}, {steps} applies each, {kind} kernel built -O3 -ffast-math")
    return gdofs, cores, kind, sample, steps, secs


def run_reference_arm(args, rank):
    if rank != 0:
        return
    gdofs, cores, kind, sample, steps, secs = cpu_reference_run(
        args.degree, args.float, args.steps, args.warmup, target_seconds=120.0)
    line = {
        "impl": "reference", "metric": "gdof_per_second_laplacian_action", "value": gdofs,
        "unit": "GDoF/s", "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * secs / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": f"f{args.float}", "data": "synthetic",
        "config": reference_config(args),
        "cpu_baseline": {"value": gdofs, "unit": "GDoF/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": gdofs, "unit": "GDoF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def reference_config(args):
    """What the reference arm really ran: the reference CPU kernel supports qmode 0 only
    (src/laplacian_cpu.hpp:73-74) and its `int` offsets limit the mesh, so the arm runs one small box per core."""
    r = getattr(cpu_reference_run, "last", {})
    nc = r.get("cells_per_instance", [0, 0, 0])
    return {
        "workload": (f"Q{args.degree} matrix-free Laplacian action on the host CPU, qmode 0 (all the reference CPU kernel "
                     f"supports), fp{args.float}, {r.get('instances')} independent instances (one per core) of a "
                     f"{nc[0]}x{nc[1]}x{nc[2]}-cell box with {r.get('ndofs_per_instance')} dofs each; NOT the GPU arm's "
                     f"qmode {args.qmode} / {args.ndofs}-dofs-per-GPU mesh (the GPU arm reports its own qmode-0 number "
                     "under gpu_qmode0 for a like-for-like ratio)"),
        "degree": args.degree, "qmode": 0, "ndofs_per_instance": r.get("ndofs_per_instance"),
        "instances": r.get("instances"), "cells_per_instance": nc,
    }


def workload_config(args, world):
    return {
        "workload": (f"Q{args.degree} matrix-free Laplacian action, qmode {args.qmode} (GLL), fp{args.float}, "
                     f"{args.ndofs} dofs per GPU on the compute_mesh_size unit-cube hex box, kappa=2, "
                     "u = assembled RHS of src/main.cpp:81-92"),
        "degree": args.degree, "qmode": args.qmode, "ndofs_per_gpu": args.ndofs,
        "partition": {1: "1x1x1", 2: "2x1x1", 4: "2x2x1", 8: "2x2x2"}.get(world, f"{world} ranks"),
        "l2": "inputs (G, dofmap, vectors) are >100x larger than the 126 MB L2; no flush needed",
    }


def reference_gpu_kernel_leg(bd, degree, qmode, dtype, ndofs, steps):
    """GDoF/s of the reference's stiffness_operator_gpu<T,P,Q> (src/laplacian_gpu.hpp:91-426, launched as
    src/laplacian.hpp:298-306: one CTA of (Q,Q,Q) threads per cell) and of our kernels on the same mesh."""
    import ctypes as C
    import torch
    from benchmark_dolfinx_b200 import _lib
    from oracle import native
    ref = native.ref_gpu_lib()
    if ref is None:
        return {"value": None, "note": "oracle/_ref/libref_gpu.so not built"}
    bits = 64 if dtype == np.float64 else 32
    n = bd.compute_mesh_size(ndofs, degree)
    mesh = bd.BoxMesh(n, degree, dtype=dtype)
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype, geometry="precomputed")
    t = op.tables
    gp, gn = C.c_void_p(), C.c_int64()
    _lib.check(_lib.lib.bd_laplacian_geometry(op._h, C.byref(gp), C.byref(gn)))   # reference-layout copy of G
    tdt = torch.float64 if bits == 64 else torch.float32
    u = mesh.assemble_rhs(qmode)
    y = torch.zeros_like(u)
    consts = torch.full((mesh.ncells,), 2.0, dtype=tdt, device="cuda")
    phi0, dphi1 = torch.from_numpy(t.phi0).cuda(), torch.from_numpy(t.dphi1).cuda()
    i = mesh.info

    def ref_apply():
        y.zero_()
        native.ref_gpu_stiffness(ref, degree, t.nq, t.is_identity, u.data_ptr(), y.data_ptr(), gp.value,
                                 consts.data_ptr(), phi0.data_ptr(), dphi1.data_ptr(), i.cell_dofmap, i.lcells,
                                 int(mesh.n_lcells), i.bc_marker, bits)

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 1e3 / steps

    t_ref = timed(ref_apply)
    ynorm_ref = bd.norm(y)
    t_ours = timed(lambda: op.apply(u, y))
    ynorm = bd.norm(y)
    out = {"value": mesh.ndofs_global / (1e9 * t_ref), "unit": "GDoF/s", "ms_per_apply": 1e3 * t_ref,
           "ours_same_mesh_precomputed_G": mesh.ndofs_global / (1e9 * t_ours), "speedup_precomputed_G": t_ref / t_ours,
           "ndofs": mesh.ndofs_global, "y_norm_reference_kernel": ynorm_ref, "y_norm_ours": ynorm,
           "rel_diff_y_norm": abs(ynorm - ynorm_ref) / ynorm_ref,
           "kernel": f"benchdolfinx::stiffness_operator_gpu<{'double' if bits == 64 else 'float'},{degree},{t.nq}> "
                     "(reference source, unmodified, -arch sm_100a; one CTA per cell)"}
    del op
    op2 = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype)            # the default choice
    if op2.geometry_mode()[0] != "precomputed":
        t2 = timed(lambda: op2.apply(u, y))
        out.update(ours_same_mesh_default=mesh.ndofs_global / (1e9 * t2), speedup_default=t_ref / t2,
                   default_geometry=op2.geometry_mode()[0], y_norm_ours_default=bd.norm(y))
    return out


def bind_to_gpu_numa_node(device_index):
    """Run this rank on the host cores next to its GPU (the NUMA node the GPU's PCIe root hangs off), BEFORE any
    pinned host memory is allocated: pinned pages are placed on the node of the allocating thread, and with every
    rank on node 0 the H2D/D2H copies of 8 GPUs share one socket's memory controllers and the inter-socket link
    (round 1: 43 GB/s per GPU each way at N = 1, 8 GB/s at N = 8).  Returns a description for the JSON line."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(device_index)
        bdf = f"{getattr(pr, 'pci_domain_id', 0):04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0:
            return {"pci": bdf, "numa_node": node, "bound": False, "why": "no NUMA information for the device"}
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpulist = f.read().strip()
        cpus = set()
        for part in cpulist.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        if not use:
            return {"pci": bdf, "numa_node": node, "bound": False, "why": "node's cores are outside this process's cpuset"}
        os.sched_setaffinity(0, use)
        return {"pci": bdf, "numa_node": node, "bound": True, "cores": len(use)}
    except Exception as exc:
        return {"bound": False, "why": f"{type(exc).__name__}: {exc}"}


# ----------------------------------------------------------------------------- our arm
FP64_PEAKS = None   # (DFMA, DMMA) TFLOP/s measured by bd_fp64_peak at the start of the run


class Ctx:
    """Per-process plumbing: rank layout, barrier, reductions."""

    def __init__(self, torch, dist, rank, world, comm):
        self.torch, self.dist, self.rank, self.world, self.comm = torch, dist, rank, world, comm

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def _red(self, v, op):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max(self, v):
        return self._red(v, self.dist.ReduceOp.MAX if self.world > 1 else None)

    def sum(self, v):
        return self._red(v, self.dist.ReduceOp.SUM if self.world > 1 else None)


def load_traffic(degree, qmode, bits, mode, ncells):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture, scaled to this mesh."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            tr = json.load(f)
        key = f"q{degree}_qmode{qmode}_f{bits}" + ("" if mode == "precomputed" else "_onthefly")
        if key in tr and tr[key].get("ncells"):
            return tr[key]["dram_bytes_per_launch"] * (ncells / tr[key]["ncells"])
    except Exception:
        pass
    return None


def action_leg(ctx, bd, mesh, op, u, y, steps, warmup, degree, qmode, bits, sample_clocks=False, local_rank=0):
    """W warm-up + K timed operator applications; roofline of the stiffness kernel from per-launch events."""
    torch = ctx.torch
    tsize = bits // 8
    nq = degree + 1 + qmode
    for _ in range(warmup):
        op.apply(u, y)
    ctx.barrier()
    sampler = ClockSampler(local_rank) if (sample_clocks and ctx.rank == 0) else None
    op.profile(True)
    if op.halo is not None:
        op.halo.profile(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        op.apply(u, y)
    ev1.record()
    ctx.barrier()
    secs = ctx.max(ev0.elapsed_time(ev1) / 1e3)
    kernel_ms, kernel_launches = op.profile_read()
    op.profile(False)
    halo_rec = None
    if op.halo is not None:
        h = op.halo.profile_read()
        op.halo.profile(False)
        napp = max(1, h["applications"])
        halo_rec = {"pack_ms": ctx.max(h["pack_ms"] / napp), "exchange_ms": ctx.max(h["exchange_ms"] / napp),
                    "exposed_ms": ctx.max(h["exposed_ms"] / napp), "ghost_dofs_max": int(ctx.max(float(mesh.nghost))),
                    "note": "per operator application, max over ranks; exposed = how long the ghost values arrived "
                            "after the interior-cell kernel had finished (CUDA events on the plan's stream and the "
                            "caller's stream)"}
    clocks = sampler.stop() if sampler is not None else None
    mode = op.geometry_mode()[0]
    peak, peak_src = measured_peak_gbs()
    if mode == "precomputed":
        bytes_apply = algorithmic_bytes_per_apply(mesh.ndofs_local, mesh.ncells, degree, nq, tsize)
    else:
        bytes_apply = algorithmic_bytes_per_apply_onthefly(mesh.ndofs_local, mesh.ncells, mesh.nnodes, degree, tsize)
    kernel_s = ctx.max(kernel_ms / 1e3 / steps)
    achieved = bytes_apply / kernel_s / 1e9
    gflops = algorithmic_flops_per_cell(degree, nq, qmode) * mesh.ncells / kernel_s / 1e9
    hbm = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
           "peak_source": peak_src, "frac_of_8TBs_nominal": achieved / 8000.0}
    common = {
        "traffic": load_traffic(degree, qmode, bits, mode, mesh.ncells), "kernel": op.describe(), "geometry": mode,
        "kernel_ms_per_apply": 1e3 * kernel_s, "kernel_launches_per_apply": kernel_launches // steps,
        "algorithmic_bytes_per_apply": bytes_apply, "algorithmic_bytes_per_dof": bytes_apply / mesh.ndofs_local,
        "algorithmic_flops_per_dof": algorithmic_flops_per_cell(degree, nq, qmode) * mesh.ncells / mesh.ndofs_local,
        "fp_gflops_algorithmic": gflops,
    }
    # The dof-space kernel for parallelepiped cells runs a different (cheaper) algorithm than the quadrature loop
    # SURVEY.md section 8(d) counts: 17 applications of an nd x nd matrix per dof plane = 34 nd^4 flops per cell
    # (7 applications = 14 nd^4 when every cell is a brick and G_cell is diagonal).
    # Its pipe fraction is computed from its OWN flop count; the rate in the reference algorithm's flops is kept as
    # `effective_gflops_reference_flop_count`.
    own_gflops = gflops
    if "stiffness_affine" in common["kernel"]:     # both dof-space kernels
        per_cell = (14 if mode == "onthefly-brick" else 34) * (degree + 1) ** 4
        own_gflops = per_cell * mesh.ncells / kernel_s / 1e9
        common["kernel_flops_per_dof"] = per_cell * mesh.ncells / mesh.ndofs_local
        common["effective_gflops_reference_flop_count"] = gflops
    fp = None
    if FP64_PEAKS is not None:
        # kernels with fp64 arithmetic: DMMA for nq = 7, 8 (also with float storage), DFMA for fp64 otherwise
        dmma_kernel = nq in (7, 8) and ("mma8" in common["kernel"])
        if bits == 64 or dmma_kernel:
            pk = FP64_PEAKS[1] if dmma_kernel else FP64_PEAKS[0]
            fp = {"bound": "fp64 " + ("tensor (DMMA m8n8k4)" if dmma_kernel else "pipe (DFMA)"),
                  "achieved": own_gflops / 1e3, "peak": pk, "unit": "TFLOP/s", "frac": own_gflops / 1e3 / pk,
                  "peak_source": "measured in this run by bd_fp64_peak (dependent-chain-free DFMA / DMMA loops, "
                                 "csrc/peaks.cu); MEASURED_PEAKS.json has no fp64 figure",
                  "note": "achieved = algorithmic flops of the kernel's algorithm (SURVEY.md section 8d for the "
                          "quadrature-space kernels, 34 nd^4 per cell for the dof-space kernel) / kernel time; the "
                          "kernels execute somewhat more (idle lanes, padding, index arithmetic)"}
    # element contributions scattered per second (one RED each unless two cells' shared face was merged first): the
    # third ceiling of the on-the-fly kernels (measured limit on this part: about 200 G/s for fp64)
    common["scatter_adds_per_second_g"] = (degree + 1) ** 3 * mesh.ncells / kernel_s / 1e9
    if mode == "precomputed" or fp is None or hbm["frac"] >= fp["frac"]:
        # stored G: the G stream bounds the kernel; on the fly: whichever ceiling is nearer leads, the other is kept
        roofline = dict(hbm, **common)
        roofline["fp64"] = fp
    else:
        roofline = dict(fp, **common)
        roofline["hbm"] = hbm
    return {"value": mesh.ndofs_global * steps / (1e9 * secs), "ms_per_step": 1e3 * secs / steps, "secs": secs,
            "roofline": roofline, "clocks": clocks, "launches": kernel_launches, "geometry": mode, "halo": halo_rec}


def cg_leg(ctx, bd, mesh, op, u, steps, warmup, bytes_apply, tsize):
    """`steps` CG iterations (timed region includes the initial apply, as src/cg.hpp:108)."""
    torch = ctx.torch
    x = torch.zeros_like(u)
    bd.cg_solve(op, x, u, max(1, warmup), 0.0, ctx.comm)
    ctx.barrier()
    x.zero_()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record()
    its = bd.cg_solve(op, x, u, steps, 0.0, ctx.comm)
    c1.record()
    ctx.barrier()
    cg_secs = ctx.max(c0.elapsed_time(c1) / 1e3)
    return {"value": mesh.ndofs_global * its / (1e9 * cg_secs), "unit": "GDoF/s", "iterations": its,
            "ms_per_iter": 1e3 * cg_secs / its, "x_norm": bd.norm(x, mesh.ndofs_owned, ctx.comm),
            "hbm_gbs_algorithmic": (bytes_apply * (its + 1) + 10 * tsize * mesh.ndofs_local * its) / cg_secs / 1e9,
            "geometry": op.geometry_mode()[0],
            "note": "timed region = initial apply + iterations, as the reference's cg_solve"}


def parity_leg(ctx, bd, geometry):
    """Known answers of the reference through the SAME communicator / partition as the timed legs:
    examples/mat_comp-16.json (Q3, qmode 1, 100 048 dofs: u_norm, y_norm), src/test_output.py:19 (Q3, qmode 0,
    1000 dofs: y_norm) and a 10 000-dof 50-iteration CG whose answer is pinned by the oracle's CG
    (tests/test_gpu_nccl.py).  Norms are partition independent."""
    torch = ctx.torch
    out = {}
    worst = 0.0
    for name, ndofs, qmode, unorm_ref, ynorm_ref in (
            ("mat_comp_16", 100000, 1, 0.6895773850559623, 0.14150257625641838),
            ("test_output", 1000, 0, None, 9.912865833415553)):
        n = bd.compute_mesh_size(ndofs, 3)
        mesh = bd.BoxMesh(n, 3, ctx.rank, ctx.world)
        halo = bd.Halo(ctx.comm, mesh) if ctx.world > 1 else None
        for geo in sorted({"precomputed", geometry, "auto-full"}):
            op = bd.MatFreeLaplacianGPU(mesh, 3, qmode, 2.0, halo=halo, geometry=geo)
            u = mesh.assemble_rhs(qmode)
            y = torch.zeros_like(u)
            op.apply(u, y)
            ynorm = bd.norm(y, mesh.ndofs_owned, ctx.comm)
            err = abs(ynorm - ynorm_ref) / ynorm_ref
            if unorm_ref is not None:
                err = max(err, abs(bd.norm(u, mesh.ndofs_owned, ctx.comm) - unorm_ref) / unorm_ref)
            out[f"{name}_{op.geometry_mode()[0]}"] = {"y_norm": ynorm, "expected": ynorm_ref, "rel_err": err}
            worst = max(worst, err)
    n = bd.compute_mesh_size(10000, 3)
    mesh = bd.BoxMesh(n, 3, ctx.rank, ctx.world)
    halo = bd.Halo(ctx.comm, mesh) if ctx.world > 1 else None
    op = bd.MatFreeLaplacianGPU(mesh, 3, 1, 2.0, halo=halo, geometry=geometry)
    u = mesh.assemble_rhs(1)
    x = torch.zeros_like(u)
    its = bd.cg_solve(op, x, u, 50, 0.0, ctx.comm)
    xnorm = bd.norm(x, mesh.ndofs_owned, ctx.comm)
    cg_ref = 167.59244725655134
    cg_err = abs(xnorm - cg_ref) / cg_ref
    out["cg_50_10000dofs"] = {"x_norm": xnorm, "expected": cg_ref, "rel_err": cg_err, "iterations": its}
    out["rel_err"] = worst
    out["cg_rel_err"] = cg_err
    out["tolerance"] = 1e-12
    out["cg_tolerance"] = 1e-11     # 50 iterations amplify the summation-order differences between partitions
    out["ok"] = bool(worst <= 1e-12 and cg_err <= 1e-11 and its == 50)
    return out


def e2e_leg(ctx, mesh, op, u, steps, tsize):
    """End to end through the host-buffer C-ABI call: H2D u, (halo +) apply, D2H y every step."""
    torch = ctx.torch
    uh = torch.empty(mesh.ndofs_local, dtype=u.dtype, pin_memory=True)
    yh = [torch.empty(mesh.ndofs_local, dtype=u.dtype, pin_memory=True) for _ in range(2)]
    uh.copy_(u)
    torch.cuda.synchronize()
    uh_np, yh_np = uh.numpy(), [t.numpy() for t in yh]
    e_steps = max(2, min(steps, 20))
    s_steps = max(2, min(steps, 4))
    op.apply_host(uh_np, yh_np[0])                        # warm-up (allocates staging)
    op.apply_host_batch([uh_np, uh_np], yh_np)
    ctx.barrier()
    # (a) one synchronous call per step: H2D, apply, D2H strictly in sequence
    t0 = time.perf_counter()
    for _ in range(s_steps):
        op.apply_host(uh_np, yh_np[0])
    torch.cuda.synchronize()
    seq_local = time.perf_counter() - t0
    seq_secs = ctx.max(seq_local)
    ctx.barrier()
    # (b) the batch call: every step still copies its input in and its result out, but the copies of
    #     neighbouring steps overlap with the operator (PCIe is full duplex)
    t0 = time.perf_counter()
    op.apply_host_batch([uh_np] * e_steps, [yh_np[i & 1] for i in range(e_steps)])
    torch.cuda.synchronize()
    e_local = time.perf_counter() - t0
    e_secs = ctx.max(e_local)
    ctx.barrier()
    bytes_rank = mesh.ndofs_local * tsize
    rate_local = bytes_rank * e_steps / e_local / 1e9
    rate_min = -ctx.max(-rate_local)
    rate_max = ctx.max(rate_local)
    ndg = mesh.ndofs_global
    return {"value": ndg * e_steps / (1e9 * e_secs), "unit": "GDoF/s",
            "h2d_bytes_per_step": int(ctx.sum(bytes_rank)), "d2h_bytes_per_step": int(ctx.sum(bytes_rank)),
            "steps": e_steps, "ms_per_step": 1e3 * e_secs / e_steps, "geometry": op.geometry_mode()[0],
            "per_rank_copy_gbs_each_way": {"min": rate_min, "max": rate_max,
                                           "note": "bytes of one direction / wall time of the pipelined call, per rank"},
            "call": "bd_laplacian_apply_host_batch (pinned host in/out vectors; H2D, operator and D2H of "
                    "consecutive steps pipelined on three streams)",
            "sequential": {"value": ndg * s_steps / (1e9 * seq_secs), "ms_per_step": 1e3 * seq_secs / s_steps,
                           "steps": s_steps, "call": "bd_laplacian_apply_host, one synchronous call per step"},
            "y_norm_host": float(np.linalg.norm(yh_np[(e_steps - 1) & 1][:mesh.ndofs_owned].astype(np.float64)))
            if ctx.world == 1 else None}


def build_case(ctx, bd, degree, qmode, dtype, ndofs_per_rank, ndofs_global, geometry, perturb=0.0):
    """Mesh, halo plan, operator and input vector of one configuration (global box as src/main.cpp:237-240,272-273)."""
    torch = ctx.torch
    n = bd.compute_mesh_size(ndofs_global or ndofs_per_rank * ctx.world, degree)
    t0 = time.perf_counter()
    mesh = bd.BoxMesh(n, degree, ctx.rank, ctx.world, geom_perturb_fact=perturb, dtype=dtype)
    halo = bd.Halo(ctx.comm, mesh) if ctx.world > 1 else None
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype, halo=halo, geometry=geometry)
    u = mesh.assemble_rhs(qmode)
    y = torch.zeros_like(u)
    if halo is not None:
        halo.scatter_fwd(u)
    torch.cuda.synchronize()
    return n, mesh, halo, op, u, y, time.perf_counter() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--degree", type=int, default=3)
    ap.add_argument("--ndofs", type=int, default=300_000_000, help="dofs per GPU (reference --ndofs)")
    ap.add_argument("--ndofs_global", type=int, default=0,
                    help="fixed global size (strong scaling, reference --ndofs_global); overrides --ndofs")
    ap.add_argument("--float", type=int, default=64, choices=[32, 64])
    ap.add_argument("--qmode", type=int, default=1, choices=[0, 1])
    ap.add_argument("--geometry", default="auto", choices=["auto", "precomputed", "onthefly", "auto-full", "onthefly-full"],
                    help="geometry variant of the headline legs: G precomputed and streamed (what the reference does), "
                         "rebuilt on the fly in the stiffness kernel, or auto = the operator's own choice (on the fly "
                         "when every cell is a parallelepiped; on meshes of bricks -- the benchmark's own box -- the dof-space "
                         "kernels then also use that G_cell is diagonal; *-full = without that shortcut).  The other "
                         "variants are measured too and reported under `variants`")
    ap.add_argument("--geom_perturb_fact", type=float, default=0.0,
                    help="random perturbation of the mesh vertices as a fraction of the cell size (reference option "
                         "of the same name, src/main.cpp:163); 0 = the benchmark's unperturbed box")
    ap.add_argument("--no-cg", action="store_true", help="skip the CG legs")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg")
    ap.add_argument("--no-variants", action="store_true", help="skip the legs of the other geometry variant")
    ap.add_argument("--no-q6", action="store_true", help="skip the Q6 / 500M-dof sub-record")
    ap.add_argument("--no-parity", action="store_true", help="skip the known-answer parity leg")
    ap.add_argument("--no-reference-gpu-kernel", action="store_true",
                    help="skip timing the reference's own CUDA kernel (oracle/_ref/libref_gpu.so, test "
                         "infrastructure, the kernel to beat of SURVEY.md section 8d)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist

    import benchmark_dolfinx_b200 as bd

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU fallback")
    torch.cuda.set_device(local_rank)
    numa = bind_to_gpu_numa_node(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    comm = bd.Comm.from_torch_distributed() if world > 1 else None
    ctx = Ctx(torch, dist, rank, world, comm)

    global FP64_PEAKS
    FP64_PEAKS = bd.fp64_peak()
    dtype = np.float64 if args.float == 64 else np.float32
    tsize = args.float // 8
    degree, qmode = args.degree, args.qmode
    if args.ndofs_global:                                          # src/main.cpp:237-240
        args.ndofs = args.ndofs_global // world

    # ---- known answers through this communicator, before anything is timed
    parity = None
    if not args.no_parity:
        parity = parity_leg(ctx, bd, args.geometry)
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"error": "parity leg failed", "parity": parity}), flush=True)
            raise SystemExit(3)

    # ---- headline legs: operator action, CG, end to end
    n, mesh, halo, op, u, y, t_setup = build_case(ctx, bd, degree, qmode, dtype, args.ndofs, args.ndofs_global,
                                                  args.geometry, args.geom_perturb_fact)
    ndofs_global = mesh.ndofs_global
    head = action_leg(ctx, bd, mesh, op, u, y, args.steps, args.warmup, degree, qmode, args.float,
                      sample_clocks=True, local_rank=local_rank)
    unorm = bd.norm(u, mesh.ndofs_owned, comm)
    ynorm = bd.norm(y, mesh.ndofs_owned, comm)
    launches_per_apply = head["roofline"]["kernel_launches_per_apply"]
    gpu_launches = args.steps * launches_per_apply + (args.steps if world > 1 else 0)
    bytes_apply = head["roofline"]["algorithmic_bytes_per_apply"]
    cg = None if args.no_cg else cg_leg(ctx, bd, mesh, op, u, args.steps, args.warmup, bytes_apply, tsize)
    e2e = None if args.no_e2e else e2e_leg(ctx, mesh, op, u, args.steps, tsize)
    head_mode = head["geometry"]
    del op

    # ---- the other geometry variant on the same mesh (both are part of the measurement)
    variants = {head_mode: {"value": head["value"], "ms_per_step": head["ms_per_step"], "roofline": head["roofline"],
                            "cg": cg and cg["value"], "y_norm": ynorm, "headline": True}}
    if not args.no_variants:
        others = ["precomputed" if head_mode != "precomputed" else "onthefly"]
        if head_mode == "onthefly-brick":
            others.append("onthefly-full")   # the general parallelepiped kernel on the same mesh
        for other in others:
            try:
                op2 = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype, halo=halo, geometry=other)
                leg = action_leg(ctx, bd, mesh, op2, u, y, args.steps, args.warmup, degree, qmode, args.float)
                yn2 = bd.norm(y, mesh.ndofs_owned, comm)
                cg2 = None if args.no_cg else cg_leg(ctx, bd, mesh, op2, u, args.steps, args.warmup,
                                                     leg["roofline"]["algorithmic_bytes_per_apply"], tsize)
                variants[leg["geometry"]] = {"value": leg["value"], "ms_per_step": leg["ms_per_step"],
                                             "roofline": leg["roofline"], "cg": cg2 and cg2["value"], "y_norm": yn2,
                                             "halo": leg["halo"],
                                             "rel_diff_y_norm_vs_headline": abs(yn2 - ynorm) / ynorm, "headline": False}
                gpu_launches += args.steps * leg["roofline"]["kernel_launches_per_apply"]
                del op2
            except bd.BenchDolfinxError as exc:
                variants[other] = {"value": None, "note": str(exc)}

    # ---- the GPU at qmode 0, beside the qmode-0 CPU baseline
    gpu_q0 = None
    if qmode != 0 and world == 1 and not args.no_cpu:
        del mesh, halo, u, y
        _, mesh0, halo0, op0, u0, y0, _ = build_case(ctx, bd, degree, 0, dtype, args.ndofs, args.ndofs_global,
                                                     args.geometry, args.geom_perturb_fact)
        leg = action_leg(ctx, bd, mesh0, op0, u0, y0, max(3, args.steps // 2), args.warmup, degree, 0, args.float)
        gpu_q0 = {"value": leg["value"], "unit": "GDoF/s", "ms_per_step": leg["ms_per_step"],
                  "roofline_frac": leg["roofline"]["frac"], "roofline_bound": leg["roofline"]["bound"], "geometry": leg["geometry"], "kernel": leg["roofline"]["kernel"],
                  "note": "same mesh and degree at qmode 0, the only mode the reference CPU kernel supports"}
        del op0, mesh0, halo0, u0, y0
    else:
        del mesh, halo, u, y
    torch.cuda.empty_cache()

    # ---- Q6 at 500M dofs per GPU: the other half of BASELINE.json's metric ("Q3/Q6")
    q6 = None
    default_headline = (degree == 3 and qmode == 1 and args.float == 64 and not args.ndofs_global
                        and args.ndofs == 300_000_000)
    if default_headline and not args.no_q6:
        n6, mesh6, halo6, op6, u6, y6, t6 = build_case(ctx, bd, 6, 1, np.float64, 500_000_000, 0, args.geometry)
        leg = action_leg(ctx, bd, mesh6, op6, u6, y6, args.steps, args.warmup, 6, 1, 64)
        q6 = {"value": leg["value"], "unit": "GDoF/s", "ms_per_step": leg["ms_per_step"], "roofline": leg["roofline"],
              "geometry": leg["geometry"], "ndofs_global": mesh6.ndofs_global, "mesh_cells": n6, "setup_seconds": t6,
              "halo": leg["halo"],
              "y_norm": bd.norm(y6, mesh6.ndofs_owned, comm),
              "cg": None if args.no_cg else cg_leg(ctx, bd, mesh6, op6, u6, args.steps, args.warmup,
                                                   leg["roofline"]["algorithmic_bytes_per_apply"], 8)}
        gpu_launches += args.steps * leg["roofline"]["kernel_launches_per_apply"]
        if not args.no_variants:
            others = ["precomputed" if leg["geometry"] != "precomputed" else "onthefly"]
            if leg["geometry"] == "onthefly-brick":
                others.append("onthefly-full")
            q6["variants"] = {leg["geometry"]: {"value": leg["value"], "roofline_frac": leg["roofline"]["frac"]}}
            for other in others:
                del op6
                op6 = bd.MatFreeLaplacianGPU(mesh6, 6, 1, 2.0, halo=halo6, geometry=other)
                leg2 = action_leg(ctx, bd, mesh6, op6, u6, y6, args.steps, args.warmup, 6, 1, 64)
                q6["variants"][leg2["geometry"]] = {"value": leg2["value"], "ms_per_step": leg2["ms_per_step"],
                                                    "roofline": leg2["roofline"],
                                                    "y_norm": bd.norm(y6, mesh6.ndofs_owned, comm)}
                gpu_launches += args.steps * leg2["roofline"]["kernel_launches_per_apply"]
        del op6, mesh6, halo6, u6, y6
        torch.cuda.empty_cache()

    # ---- the kernel to beat: the reference's own stiffness_operator_gpu (unmodified source recompiled for
    #      sm_100a, oracle/_ref/libref_gpu.so) with its own launch shape, on a smaller mesh of the same
    #      degree (its G layout needs a second copy of G), timed with CUDA events on the same GPU
    ref_gpu = None
    if world == 1 and rank == 0 and not args.no_reference_gpu_kernel:
        try:
            ref_gpu = reference_gpu_kernel_leg(bd, degree, qmode, dtype, min(args.ndofs, 100_000_000),
                                               max(3, min(args.steps, 10)))
        except Exception as exc:
            ref_gpu = {"value": None, "note": f"{type(exc).__name__}: {exc}"}

    # ---- CPU baseline: the reference's CPU kernels on this box's host cores (rank 0, N = 1)
    cpu = None
    if world == 1 and rank == 0 and not args.no_cpu:
        try:
            g, cores, kind, sample, _, _ = cpu_reference_run(degree, args.float, 1000, 1, target_seconds=15.0)
            cpu = {"value": g, "unit": "GDoF/s", "cores": cores, "kind": kind, "sample": sample}
        except Exception as exc:  # the baseline is reported, never required for the GPU number
            cpu = {"value": None, "unit": "GDoF/s", "cores": os.cpu_count(), "kind": "unavailable",
                   "sample": f"{type(exc).__name__}: {exc}"}

    if rank == 0:
        cfg = dict(workload_config(args, world), ndofs_global_requested=args.ndofs_global or None, mesh_cells=n,
                   ndofs_global=ndofs_global, setup_seconds=t_setup, geometry=head_mode,
                   geom_perturb_fact=args.geom_perturb_fact,
                   geometry_requested=args.geometry)
        line = {
            "metric": "gdof_per_second_laplacian_action", "value": head["value"], "unit": "GDoF/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "strong" if args.ndofs_global else "weak",
            "vs_baseline": None, "dtype": f"f{args.float}", "data": "synthetic", "config": cfg,
            "roofline": head["roofline"], "cpu_baseline": cpu, "gpu_qmode0": gpu_q0, "e2e": e2e, "cg": cg,
            "halo": head["halo"], "variants": variants, "q6": q6, "parity": parity, "reference_gpu_kernel": ref_gpu,
            "host_binding_rank0": numa, "fp64_peaks_measured_tflops": {"dfma": FP64_PEAKS[0], "dmma": FP64_PEAKS[1]},
            "gpu_launches": gpu_launches, "clocks": head["clocks"], "u_norm": unorm, "y_norm": ynorm,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
