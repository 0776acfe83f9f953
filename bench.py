#!/usr/bin/env python
"""bench.py — WENO-reconstructed faces/sec of the correction() hot path.

  python bench.py --gpus N --steps K --warmup W              (our arm, N=1)
  python -m torch.distributed.run ... bench.py --gpus N ...  (our arm, N>1)
  python bench.py --impl reference ...                       (CPU arm: the oracle
        port of the reference loop on all host threads; the reference library
        itself needs OpenFOAM and cannot be built)

Workload (BASELINE.json configs[2], the one the metric is quoted on): 128^3 hex
box per GPU, `WENOUpwindFit 2 1` (order 2, bounded), vector field (3 components
share every LSmatrix read). N>1: weak scaling, the global box is decomposed
`simple`-style into N blocks of 128^3 (2x1x1, 2x2x1, 2x2x2), halo exchange of
stencil cell values and processor-patch faces over NCCL.

One JSON line on stdout (rank 0). A "step" = one correction() call.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

NPROC = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=128, help="cells per direction per GPU")
    ap.add_argument("--order", type=int, default=2)
    ap.add_argument("--ncomp", type=int, default=3)
    ap.add_argument("--bounded", type=int, default=1)
    ap.add_argument("--tets", type=int, default=0, choices=[0, 6],
                    help="6: split every hex into 6 tets (irregular stencils, BASELINE config 4 variant)")
    ap.add_argument("--best-conditioned", action="store_true", help="WENODict bestConditioned true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def workload_name(a, N):
    px, py, pz = NPROC[N]
    g = (a.n * px, a.n * py, a.n * pz)
    mesh = "hex box" if not a.tets else "tet-split box (6 tets per hex" + \
        (", bestConditioned" if a.best_conditioned else "") + ")"
    return (f"{mesh} {g[0]}x{g[1]}x{g[2]} ({a.n}^3 per GPU), WENOUpwindFit {a.order} {a.bounded}, "
            f"{'vector U' if a.ncomp == 3 else str(a.ncomp) + ' component(s)'}, fp64")


def case_kwargs(a):
    kw = {}
    if a.tets:
        kw["tetSplit"] = a.tets
    if a.best_conditioned:
        kw["bestConditioned"] = True
    return kw


def make_fields(case, ncomp, rank=0):
    """synthetic fields of SURVEY.md 8(d): psi = sum sin(pi x_i) family + 1e-3 noise,
    flux of a rotating velocity (mixed signs) with a few exact zeros"""
    from tests import helpers as H
    lo, hi = case.C.min(axis=0), case.C.max(axis=0)
    v = H.vector_field(case.C, seed=1 + rank)
    if ncomp <= 3:
        vf = np.ascontiguousarray(v[:, :ncomp])
    else:
        vf = np.ascontiguousarray(np.concatenate([v] * ((ncomp + 2) // 3), axis=1)[:, :ncomp])
    phi = H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.001, seed=7 + rank)
    return vf, phi


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def mark(self):
        """start of the timed region: earlier samples are dropped (unless nothing follows)"""
        t0 = time.time()
        while self.proc and not self.lines and time.time() - t0 < 5.0:
            time.sleep(0.01)  # nvidia-smi still starting
        self.first = len(self.lines)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        timed = self.lines[getattr(self, "first", 0):]
        for ln in (timed if timed else self.lines[-3:]):
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(mx)) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def oracle_for_cpu():
    """the checker used as the timed CPU leg: Blaze-backed variant (the reference's
    own arithmetic library) when oracle/_ref was built, else the plain-C port"""
    from tests.oracle_binding import BLAZE_SO, Oracle
    if os.path.exists(BLAZE_SO):
        return Oracle("blaze"), "oracle/_ref Blaze-backed restatement of WENOCoeff.C/WENOUpwindFit.C"
    return Oracle("c"), "oracle/ plain-C restatement of WENOCoeff.C/WENOUpwindFit.C"


def run_reference(a):
    """--impl reference: the CPU implementation of the path on all host threads.
    libWENOEXT itself cannot be compiled without OpenFOAM, so this is
    the oracle port, cell and face loops split over the host threads the way
    `mpirun -np <cores>` splits the reference."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import wenoext_b200 as wx
    from tests.oracle_binding import correction_threaded
    t0 = time.time()
    case = wx.BoxCase((a.n,) * 3, a.order, **case_kwargs(a))
    log(f"[reference] setup {time.time() - t0:.1f}s, {case.nCells} cells")
    vf, phi = make_fields(case, a.ncomp)
    orc, what = oracle_for_cpu()
    T = host_threads()
    for _ in range(max(a.warmup, 1)):
        correction_threaded(orc, case, a.bounded, vf, phi, T)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        correction_threaded(orc, case, a.bounded, vf, phi, T)
    dt = (time.perf_counter() - t0) / a.steps
    val = case.nInternalFaces / dt
    out = {
        "impl": "reference", "metric": "WENO-reconstructed faces/sec", "value": val,
        "unit": "faces/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a, a.gpus), "sample_cells": case.nCells,
                   "sample_faces": case.nInternalFaces},
        "cpu_baseline": {"value": val, "unit": "faces/s", "cores": T, "kind": "port",
                         "cpu": cpu_model(),
                         "sample": f"one full {a.n}^3 block (the per-GPU share of the workload) per step, {what}, "
                                   f"de-duplicated matrices ({case.desc.nUniqueMatrices} unique)"},
        "e2e": {"value": val, "unit": "faces/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(out), flush=True)


def run_ours(a):
    import torch
    import torch.distributed as dist

    import wenoext_b200 as wx
    from wenoext_b200 import SCHEME_UPWINDFIT as FIT

    N = a.gpus
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != N:
        if N == 1 and world == 1:
            pass
        else:
            raise SystemExit(f"--gpus {N} needs torchrun with {N} ranks (WORLD_SIZE={world})")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; wenoext_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if N > 1:
        dist.init_process_group("nccl", device_id=dev)
    px, py, pz = NPROC[N]
    t0 = time.time()
    case = wx.BoxCase((a.n * px, a.n * py, a.n * pz), a.order, nProc=(px, py, pz), myProc=rank,
                      **case_kwargs(a))
    if N > 1:
        wx.exchange_halo_lists_distributed(case)
    t_setup = time.time() - t0
    t0 = time.time()
    plan = wx.Plan(case.desc, device=local)
    if N > 1:
        uid = [wx.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        plan.attach_nccl(uid[0], rank, N)
    t_plan = time.time() - t0
    info = plan.info
    log(f"[rank {rank}] setup {t_setup:.1f}s plan {t_plan:.1f}s cells {info.nCells} "
        f"records {info.recordBytes / 1e9:.2f} GB halo {info.nHaloCells} interior {info.nInteriorCells}")
    vf_h, phi_h = make_fields(case, a.ncomp, rank)
    vf = torch.from_numpy(vf_h).to(dev)
    phi = torch.from_numpy(phi_h).to(dev)
    out = torch.empty((plan.nFaces, a.ncomp), dtype=torch.float64, device=dev)
    faces = int(info.nInternalFaces + info.nCoupledFaces)
    alg_bytes = plan.alg_bytes(a.ncomp)

    def barrier():
        if N > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        plan.correction(FIT, a.bounded, vf, phi, out=out)

    # every rank samples its own GPU; rank 0 reports the slowest. nvidia-smi needs a while to
    # start (longer with 8 of them), so it is started before the warm-up and only the samples
    # taken during the timed region are used.
    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(a.warmup):
        step()
    # ---- timed region: device-resident inputs, K steps, CUDA events ----
    plan.enable_kernel_timing(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    sampler.mark()
    barrier()
    e0.record()
    for _ in range(a.steps):
        step()
        launches += plan.last_launch_count()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    if N > 1:
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        cm = torch.tensor([clocks["sm_mhz"] or 0.0] + [float(nm in clocks["reasons"]) for nm in names],
                          dtype=torch.float64, device=dev)
        lo, hi = cm.clone(), cm.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        clocks["sm_mhz"] = lo[0].item() or None  # median under load of the slowest GPU
        clocks["reasons"] = [nm for i, nm in enumerate(names) if hi[1 + i].item() > 0]
        clocks["ranks"] = N
    k_ms, k_n = plan.kernel_time()
    plan.enable_kernel_timing(False)
    t = torch.tensor([ms, k_ms / max(a.steps, 1)], dtype=torch.float64, device=dev)
    tot = torch.tensor([faces, alg_bytes], dtype=torch.float64, device=dev)
    if N > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_step = t[0].item() / a.steps
    value = tot[0].item() / (ms_step * 1e-3)

    # ---- e2e: the same call through the C ABI with HOST (pinned) buffers ----
    e2e = None
    if not a.no_e2e:
        pin = lambda x: torch.from_numpy(x).pin_memory().numpy()  # noqa: E731
        vf_p, phi_p = pin(vf_h), pin(phi_h)
        out_p = torch.empty((plan.nFaces, a.ncomp), dtype=torch.float64).pin_memory().numpy()
        for _ in range(2):
            plan.correction(FIT, a.bounded, vf_p, phi_p, out=out_p)
        ke = max(3, min(a.steps, 10))
        barrier()
        w0 = time.perf_counter()
        for _ in range(ke):
            plan.correction(FIT, a.bounded, vf_p, phi_p, out=out_p)
        torch.cuda.synchronize()
        w = torch.tensor([(time.perf_counter() - w0) / ke], dtype=torch.float64, device=dev)
        if N > 1:
            dist.all_reduce(w, op=dist.ReduceOp.MAX)
        e2e = {"value": tot[0].item() / w.item(), "unit": "faces/s",
               "ms_per_step": w.item() * 1e3,
               "h2d_bytes_per_step": int(vf_p.nbytes + phi_p.nbytes),
               "d2h_bytes_per_step": int(out_p.nbytes)}

    if rank != 0:
        if N > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # DRAM traffic of the dominant kernel: taken from the committed ncu capture of this
    # workload (profiles/traffic.json), null for workloads that were not captured
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj.get(workload_name(a, 1), {}).get("traffic_bytes_per_launch")
    except OSError:
        pass
    k_ms_launch = t[1].item()  # fused cell kernel(s) per step, max over ranks
    achieved = alg_bytes / (k_ms_launch * 1e-3) / 1e9 if k_ms_launch > 0 else 0.0
    res = {
        "metric": "WENO-reconstructed faces/sec", "value": value, "unit": "faces/s",
        "n_gpus": N, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": workload_name(a, N), "cells_per_gpu": int(info.nCells),
                   "faces_per_step": int(tot[0].item()),
                   "l2": "inputs (per-cell records, %.1f GB) are far larger than the 126 MB L2; "
                         "no flush needed" % (info.recordBytes / 1e9),
                   "setup_s": round(t_setup, 1)},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic,
                     "kernel": "weno_cell_kernel", "kernel_ms": k_ms_launch,
                     "alg_bytes_per_launch": alg_bytes,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks
                     else "fallback 6650 GB/s (of fallback)"},
    }
    if e2e:
        res["e2e"] = e2e
    if N == 1 and not a.no_cpu_baseline:
        from tests.oracle_binding import correction_threaded
        orc, what = oracle_for_cpu()
        t0 = time.perf_counter()
        ref = correction_threaded(orc, case, a.bounded, vf_h, phi_h, 1)
        dt = time.perf_counter() - t0
        from tests import helpers as H
        err = H.normwise_err(out.cpu().numpy(), ref)
        res["cpu_baseline"] = {"value": case.nInternalFaces / dt, "unit": "faces/s", "cores": 1,
                               "kind": "port", "cpu": cpu_model(),
                               "sample": f"one full correction() of the same {a.n}^3 workload, "
                                         f"serial, {what}"}
        res["parity_vs_oracle"] = err
    print(json.dumps(res), flush=True)
    if N > 1:
        dist.barrier()
        dist.destroy_process_group()


def _quiet_stdout():
    """Libraries (NCCL's version banner, torchrun) write to fd 1; the contract is ONE JSON
    line on stdout. Route fd 1 to stderr for the run and keep the real stdout for the result."""
    real = os.fdopen(os.dup(1), "w")
    sys.stdout.flush()
    os.dup2(2, 1)
    return real


if __name__ == "__main__":
    args = parse()
    _real_stdout = _quiet_stdout()
    sys.stdout = _real_stdout
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
